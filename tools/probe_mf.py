"""Matrix-free vs stored-block K product and CG on one GPU (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pynama_b200.device_mesh import DeviceMesh
from pynama_b200.domain.box_mesh import BoxMesh
from pynama_b200.petsc import Context, KSP, PC, Vec

nel = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ngl = int(sys.argv[2]) if len(sys.argv) > 2 else 4
mesh = BoxMesh([nel] * 3, [0] * 3, [1] * 3, ngl, perturb=0.15)
ctx = Context.get()
dm = DeviceMesh(3, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
cls = np.zeros(mesh.nnodes * 3, np.uint8)
b = mesh.getBordersNodes()
for d in range(3):
    cls[b * 3 + d] = 1
dm.set_dof_class(cls)
mats = dm.assemble_kle("FS")
K = mats["K"]
rng = np.random.default_rng(12345)
n = mesh.nnodes * 3
x = Vec().createWithArray(rng.uniform(-1, 1, n)); y = Vec(n)
ref = None
for mf in (False, True):
    K.setMatFree(mf)
    for _ in range(3):
        K.mult(x, y)
    reps = 50
    ctx.sync(); ctx.timer_start()
    for _ in range(reps):
        K.mult(x, y)
    ms = ctx.timer_stop() / reps
    ya = y.getArray()
    if ref is None:
        ref = ya
    print(f"{nel}^3 ngl={ngl} K*x matfree={mf}: {ms*1e3:8.1f} us  maxdiff {np.abs(ya-ref).max()/np.abs(ref).max():.2e}", flush=True)
xs = np.where(cls == 0, rng.uniform(-1, 1, n), 0.5)
K.setMatFree(False)
rhs = K * Vec().createWithArray(xs)
for mf in (False, True):
    K.setMatFree(mf)
    ksp = KSP().create(); ksp.setType("cg"); pc = PC().create(); pc.setType("bjacobi"); ksp.setPC(pc)
    ksp.setOperators(K); ksp.setTolerances(rtol=1e-10, max_it=20000); ksp.setUp()
    for rep in range(2):
        sol = Vec(n)
        ctx.sync(); ctx.timer_start()
        ksp.solve(rhs, sol)
        ms = ctx.timer_stop()
    ctx.prof_enable(True); sol = Vec(n); ksp.solve(rhs, sol); ctx.sync()
    print("   per-launch us:", {k: round(ctx.prof_get(k)[1] / max(ctx.prof_get(k)[0], 1) * 1e3, 2) for k in ("spmv_ksp", "cg_update", "cg_direction")}); ctx.prof_enable(False)
    print(f"CG matfree={mf}: its {ksp.its} {ms:.2f} ms ({ms/max(ksp.its,1)*1e3:.1f} us/it) err {np.abs(sol.getArray()-xs).max():.2e}", flush=True)
for mf in (False, True):
    K.setMatFree(mf)
    ksp = KSP().create(); ksp.setType("cg"); pc = PC().create(); pc.setType("bjacobi"); ksp.setPC(pc)
    ksp.setCGSingleReduction(True)
    ksp.setOperators(K); ksp.setTolerances(rtol=1e-10, max_it=20000); ksp.setUp()
    for rep in range(2):
        sol = Vec(n)
        ctx.sync(); ctx.timer_start()
        ksp.solve(rhs, sol)
        ms = ctx.timer_stop()
    ctx.prof_enable(True); sol = Vec(n); ksp.solve(rhs, sol); ctx.sync()
    print("   per-launch us (single reduction):", {k: round(ctx.prof_get(k)[1] / max(ctx.prof_get(k)[0], 1) * 1e3, 2) for k in ("spmv_ksp", "cg_update", "cg_direction")}); ctx.prof_enable(False)
    print(f"CG single-reduction matfree={mf}: its {ksp.its} {ms:.2f} ms ({ms/max(ksp.its,1)*1e3:.1f} us/it) err {np.abs(sol.getArray()-xs).max():.2e}", flush=True)
