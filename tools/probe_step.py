"""Where does a time step go?  Wall-clock (with device syncs) per piece of evalRHS (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import argparse
import bench
from pynama_b200 import _lib
from pynama_b200.petsc import Context

args = argparse.Namespace(nelem=int(sys.argv[1]) if len(sys.argv) > 1 else 16, ngl=int(sys.argv[2]) if len(sys.argv) > 2 else 4,
                          rtol=1e-10, dt=1e-3, pc="bjacobi")
ctx = Context.get()
fem = bench.build_problem(args)
for _ in range(2):
    fem.ts.step()
ctx.sync()
T = {"bc": 0.0, "solve": 0.0, "ops": 0.0}
vel = fem.solverKLE.getSolution()
def rhs(ts, t, vort, f):
    w = fem.vort
    ctx.sync(); t0 = time.perf_counter()
    fem.dom.applyBoundaryConditions(w, "vorticity", t, fem.nu)
    fem.dom.applyBoundaryConditions(vel, "velocity", t, fem.nu)
    ctx.sync(); t1 = time.perf_counter()
    fem.solverKLE.solve(w)
    ctx.sync(); t2 = time.perf_counter()
    fem.operator.SrT.mult(vel, fem._Aux1)
    _lib.call("pyn_stress_combine", fem._Aux1._h, vel._h, float(fem.mu), float(fem.rho), fem.dim)
    _lib.call("pyn_mat_mult_axpby", fem.operator.DivSrT._h, 1.0 / fem.rho, fem._Aux1._h, 0.0, None, fem._rhsWork._h)
    fem.operator.Curl.mult(fem._rhsWork, f)
    ctx.sync(); t3 = time.perf_counter()
    T["bc"] += t1 - t0; T["solve"] += t2 - t1; T["ops"] += t3 - t2
    T.setdefault("its", []).append(fem.solverKLE.solver.getIterationNumber())
fem.ts._rhs = rhs
ctx.prof_enable(True)
t0 = time.perf_counter()
n = 3
for _ in range(n):
    fem.ts.step()
ctx.sync()
wall = time.perf_counter() - t0
print(f"step wall {wall/n*1e3:.1f} ms; per step: bc {T['bc']/n*1e3:.2f} ms, solve {T['solve']/n*1e3:.2f} ms, ops {T['ops']/n*1e3:.2f} ms")
print("iterations per solve:", T["its"])
for k in ("spmv_ksp", "spmv", "cg_update", "cg_direction"):
    c, ms = ctx.prof_get(k)
    print(f"  {k}: {c/n:.0f} launches/step, {ms/n:.2f} ms/step, avg {ms/max(c,1)*1e3:.1f} us")
