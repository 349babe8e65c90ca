"""Where does the HOST spend its time in a time step?  cProfile over three RK5BS steps of the bench problem."""
import cProfile
import io
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

sys.argv = [sys.argv[0]] + sys.argv[1:]
args = bench.parse_args()
from pynama_b200.petsc import Context  # noqa: E402
ctx = Context.get()
fem = bench.build_problem(args)
for _ in range(3):
    fem.ts.step()
ctx.sync()
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    fem.ts.step()
ctx.sync()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(22)
print(s.getvalue())
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(30)
print(s.getvalue())
