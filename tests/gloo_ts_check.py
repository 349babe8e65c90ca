"""CPU, world_size 2 (gloo): the adaptive RK controller of pynama_b200/ts.py on a row-partitioned
state.  Each rank owns a slice of a decoupled linear ODE whose local error differs from rank to rank;
the error norm is reduced over the ranks, so every rank must take the same accepted / rejected step
sequence as a single rank integrating the whole vector.  Launched by tests/test_dist_gloo.py."""
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pynama_b200.petsc import COMM_WORLD, Options                 # noqa: E402
from pynama_b200.solver.ts_solver import TsSolver                  # noqa: E402
from tests.test_ts_host import NpVec                               # noqa: E402

LAM = -np.concatenate([np.linspace(0.5, 3.0, 40), np.linspace(5.0, 30.0, 24)])     # faster tail on the last rank


def run(lam, world_size_override=None):
    if world_size_override is not None:
        saved = os.environ["WORLD_SIZE"]
        os.environ["WORLD_SIZE"] = str(world_size_override)
    try:
        ts = TsSolver(COMM_WORLD)
        ts.setUpTimes(0.0, 1.0, 10000)
        log = []
        ts.initSolver(lambda ts_, t, u, f: np.copyto(f.a, lam * u.a * (1.0 + 0.5 * np.sin(5 * t))),
                      lambda ts_: log.append((ts_.time, ts_._last_dt, ts_.rejections)))
        u = NpVec(np.ones_like(lam))
        ts.solve(u)
        return log, u.a
    finally:
        if world_size_override is not None:
            os.environ["WORLD_SIZE"] = saved


def main():
    dist.init_process_group("gloo")
    rank, size = dist.get_rank(), dist.get_world_size()
    o = Options()
    o.setValue("ts_adapt_type", "basic")
    o.setValue("ts_dt", 0.5)            # too large: the first steps are rejected
    cut = np.linspace(0, len(LAM), size + 1).astype(int)
    log, mine = run(LAM[cut[rank]:cut[rank + 1]])
    assert log[-1][2] > 0, "the first step must be rejected"
    logs = COMM_WORLD.allgather(log)
    for other in logs[1:]:
        assert other == logs[0], "ranks took different step sequences"
    whole_log, whole = run(LAM, world_size_override=1)             # the same system on one rank
    assert len(whole_log) == len(log)
    # (the split sum of squares rounds differently from the whole one: agreement to rounding, not bits)
    for a, b in zip(whole_log, log):
        assert abs(a[0] - b[0]) <= 1e-8 and abs(a[1] - b[1]) <= 1e-8 * abs(a[1]) and a[2] == b[2], (a, b)
    assert np.abs(whole[cut[rank]:cut[rank + 1]] - mine).max() <= 1e-9
    print("OK", rank, len(log), log[-1][2], flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
