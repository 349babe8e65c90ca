"""BASELINE.json configs[2]: 3-D lid-driven cavity (no-slip walls, MatNS: 8 KLE matrices, two solves per RHS),
Re = 1000.  One process per GPU (torchrun) or a single process.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29700 \
        tools/run_cavity.py 32 5

Reports (rank 0, one JSON line): set-up and assembly times, device memory, SpMV(K) bandwidth of the local rows,
and ONE evaluation of the vorticity-transport right-hand side (free-slip pre-solve with K + Kfs, KLE solve with
K, stress and curl products) with its Krylov iteration counts."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

nel = int(sys.argv[1]) if len(sys.argv) > 1 else 32
ngl = int(sys.argv[2]) if len(sys.argv) > 2 else 5
rtol = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-6
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"
from pynama_b200.petsc import COMM_WORLD, Context, Options, Vec
from pynama_b200.cases.base_problem import BaseProblem
from pynama_b200.run_case import loadCase

rank, size = COMM_WORLD.rank, COMM_WORLD.size
ctx = Context.get()
o = Options()
for k, v in (("ksp_type", "gmres"), ("pc_type", "bjacobi"), ("ksp_rtol", rtol), ("ksp_max_it", 20000),
             ("ksp_gmres_restart", 50), ("ts_adapt_type", "none"), ("ts_dt", 1e-3)):
    o.setValue(k, v)
cfg = loadCase("cavity")
t0 = time.perf_counter()
fem = BaseProblem(cfg, case="cavity", nelem=[nel] * 3, ngl=ngl)
fem.setUp()
t_dom = time.perf_counter() - t0
t0 = time.perf_counter()
fem.setUpSolver()
ctx.sync()
t_asm = time.perf_counter() - t0
dm = fem.dom.getDeviceMesh()
K = fem.mat.K
ndof = K.getSize()[0]
rec = {"case": f"cavity {nel}^3 ngl={ngl} Re=1000 (BASELINE configs[2])", "ranks": size, "dofs": int(ndof),
       "host_domain_s": round(t_dom, 2), "setUpSolver_wall_s": round(t_asm, 2),
       "phase_ms": {p: round(dm.phase_ms(p), 1) for p in ("symbolic", "elem_kle", "kle_numeric", "elem_ops", "ops_numeric")},
       "device_mem_GiB_rank0": round(ctx.mem_used() / 2 ** 30, 1)}
# SpMV(K) on the local rows
x = K.createVecRight(); y = K.createVecLeft()
x.set(1.0)
for _ in range(3):
    K.mult(x, y)
ctx.sync(); ctx.timer_start()
reps = 10
for _ in range(reps):
    K.mult(x, y)
ms = ctx.timer_stop() / reps
stored, traffic = K.bytes()
rec["spmv_K"] = {"ms": round(ms, 3), "local_bytes_GB": round(traffic / 1e9, 2), "GBs_rank0": round(traffic / ms / 1e6, 1),
                 "frac_of_hbm_6548": round(traffic / ms / 1e6 / 6548.5, 3)}
# one RHS evaluation (two Krylov solves + the operator products)
fem.ts.setSolution(fem.vort)
f = fem.vort.duplicate()
its0, s0 = fem.kspIterations, fem.kspSolves
ctx.sync(); t0 = time.perf_counter()
fem.evalRHS(fem.ts, 0.0, fem.vort, f)
ctx.sync()
rec["evalRHS"] = {"wall_s": round(time.perf_counter() - t0, 2), "ksp_solves": fem.kspSolves - s0,
                  "ksp_iterations": fem.kspIterations - its0, "ksp": f"gmres(50)+bjacobi rtol {rtol:g}",
                  "max_abs_velocity": float(ctx.allreduce([np.abs(fem.solverKLE.getSolution().getArray()).max()], "max")[0]),
                  "rhs_norm": float(f.norm())}
nsteps = int(sys.argv[4]) if len(sys.argv) > 4 else 0
if nsteps > 0:                       # full RK5BS steps (8 RHS evaluations in the first, 7 afterwards)
    its0, s0, e0 = fem.kspIterations, fem.kspSolves, fem.ts.rhs_evals
    ctx.sync(); t0 = time.perf_counter()
    for _ in range(nsteps):
        fem.ts.step()
    ctx.sync()
    wall = time.perf_counter() - t0
    rec["time_steps"] = {"steps": nsteps, "wall_s_per_step": round(wall / nsteps, 2), "dt": 1e-3,
                         "rhs_evals": fem.ts.rhs_evals - e0, "ksp_iterations_main_solve": fem.kspIterations - its0,
                         "dof_per_s": round(ndof * nsteps / wall, 1),
                         "max_abs_vorticity": float(ctx.allreduce([np.abs(fem.vort.getArray()).max()], "max")[0])}
if rank == 0:
    print(json.dumps(rec), flush=True)
try:
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
except Exception:
    pass
