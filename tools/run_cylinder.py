"""BASELINE.json configs[3]: 3-D uniform flow past a (blocked-cell) cylinder, unstructured hexahedral mesh,
ngl = 4, on the GPUs of the run (RCB partition).  One JSON line from rank 0.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29660 tools/run_cylinder.py --h 0.25 --steps 2
"""
import argparse
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from pynama_b200.cases.base_problem import BaseProblem  # noqa: E402
from pynama_b200.domain.cylinder_mesh import write_cylinder_msh  # noqa: E402
from pynama_b200.petsc import COMM_WORLD, Context, Options  # noqa: E402
from pynama_b200.run_case import loadCase  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--h", type=float, default=0.25)
    ap.add_argument("--ngl", type=int, default=4)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--dt", type=float, default=1e-3)
    ap.add_argument("--rtol", type=float, default=1e-8)
    a = ap.parse_args()
    rank, size = COMM_WORLD.rank, COMM_WORLD.size
    ctx = Context.get()
    o = Options()
    for k, v in (("ksp_type", "gmres"), ("pc_type", "bjacobi"), ("ksp_rtol", a.rtol), ("ksp_max_it", 20000),
                 ("ksp_gmres_restart", 50), ("ts_adapt_type", "none"), ("ts_dt", a.dt)):
        o.setValue(k, v)
    path = os.path.join(tempfile.mkdtemp(prefix=f"pyn_cyl_{rank}_"), "cylinder-channel.msh")
    t0 = time.perf_counter()
    nel, nv = write_cylinder_msh(path, h=a.h)
    cfg = loadCase("cylinder")
    cfg["domain"]["gmsh-file"] = path
    fem = BaseProblem(cfg, case="cylinder", ngl=a.ngl)
    fem.setUp()
    t_mesh = time.perf_counter() - t0
    t0 = time.perf_counter()
    fem.setUpSolver()
    ctx.sync()
    t_setup = time.perf_counter() - t0
    fem.ts._post = None
    fem.ts.setSolution(fem.vort)
    mx = lambda v: float(ctx.allreduce([v], "max")[0]) if size > 1 else float(v)      # noqa: E731
    sm = lambda v: float(ctx.allreduce([v], "sum")[0]) if size > 1 else float(v)      # noqa: E731
    K = fem.mat.K
    ndof = K.getSize()[0]
    stored, traffic = K.bytes()
    dm = fem.dom.getDeviceMesh()
    x, y = K.createVecRight(), K.createVecLeft()
    x.set(1.0)
    for _ in range(3):
        K.mult(x, y)
    ctx.sync()
    ctx.timer_start()
    for _ in range(20):
        K.mult(x, y)
    us = mx(ctx.timer_stop() / 20 * 1e3)
    its0 = fem.kspIterations
    ctx.sync()
    ctx.timer_start()
    for _ in range(a.steps):
        fem.ts.step()
    ms = mx(ctx.timer_stop()) / a.steps
    vort = fem.vort.getGlobalArray()
    vel = fem.solverKLE.getSolution().getGlobalArray()
    own = [int(v) for v in COMM_WORLD.allgather(int(K.getLocalSize()[0]))]
    rec = {"case": "BASELINE.json configs[3]: uniform flow past a blocked-cell cylinder, channel [-4,12]x[-4,4]x[0,2]",
           "h": a.h, "ngl": a.ngl, "gpus": size, "elements": nel, "dofs": int(ndof),
           "borders": fem.dom.getMesh().sides, "bc": "all borders no-slip (MatNS, 8 matrices): channel sides at U=(1,0,0), cylinder walls at rest",
           "ksp": f"gmres(50) + bjacobi, rtol {a.rtol:g} (K + Kfs is not definite)", "dt": a.dt,
           "host_mesh_and_domain_s": round(mx(t_mesh), 2), "setup_and_assembly_s": round(mx(t_setup), 2),
           "assembly_phase_ms": {p: round(mx(dm.phase_ms(p)), 2) for p in ("symbolic", "elem_kle", "kle_numeric", "elem_ops", "ops_numeric")},
           "congruence_classes": dm.congruence()[1], "K_GB_total": round(sm(stored) / 1e9, 2),
           "mem_GiB_per_gpu_max": round(mx(ctx.mem_used()) / 2 ** 30, 1),
           "spmv_K_us": round(us, 1), "spmv_K_GBs_aggregate": round(sm(traffic) / us / 1e3, 1),
           "spmv_K_frac_of_hbm_per_gpu": round(sm(traffic) / size / us / 1e3 / 6548.8, 3),
           "ms_per_step": round(ms, 1), "ksp_iterations_per_step": (fem.kspIterations - its0) / a.steps,
           "owned_dofs_per_rank": own, "rcb_load_imbalance": round(max(own) / (sum(own) / size), 3),
           "checksum": {"vort_l2": float(np.sqrt(vort @ vort)), "vel_l2": float(np.sqrt(vel @ vel)),
                        "vort_max": float(np.abs(vort).max())}}
    if rank == 0:
        print(json.dumps(rec), flush=True)
    if size > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
