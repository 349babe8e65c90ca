"""Quick performance probe of the hot path pieces on one GPU (development aid)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pynama_b200.device_mesh import DeviceMesh
from pynama_b200.domain.box_mesh import BoxMesh
from pynama_b200.petsc import Context, KSP, PC, Vec

nel = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ngl = int(sys.argv[2]) if len(sys.argv) > 2 else 4
compact = int(sys.argv[3]) if len(sys.argv) > 3 else 1
dim = 3
t0 = time.time()
mesh = BoxMesh([nel] * 3, [0] * 3, [1] * 3, ngl, perturb=0.15)
print(f"mesh {nel}^3 ngl={ngl}: nodes {mesh.nnodes} host gen {time.time()-t0:.2f}s", flush=True)
ctx = Context.get()
t0 = time.time()
dm = DeviceMesh(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
print(f"symbolic: {dm.phase_ms('symbolic'):.1f} ms (wall {time.time()-t0:.2f}s) {dm.graph_info()}", flush=True)
cls = np.zeros(mesh.nnodes * dim, np.uint8)
b = mesh.getBordersNodes()
for d in range(dim):
    cls[b * dim + d] = 1
dm.set_dof_class(cls)
t0 = time.time()
mats = dm.assemble_kle("FS")
ctx.sync()
print(f"assemble_kle wall {time.time()-t0:.3f}s elem {dm.phase_ms('elem_kle'):.2f} ms numeric {dm.phase_ms('kle_numeric'):.2f} ms", flush=True)
t0 = time.time()
curl, srt, div, w = dm.assemble_ops(compact=bool(compact))
ctx.sync()
print(f"assemble_ops wall {time.time()-t0:.3f}s elem {dm.phase_ms('elem_ops'):.2f} ms numeric {dm.phase_ms('ops_numeric'):.2f} ms", flush=True)
print(f"device memory in use {ctx.mem_used()/2**30:.2f} GiB", flush=True)
mats.update({"Curl": curl, "SrT": srt, "DivSrT": div})
rng = np.random.default_rng(12345)
for name, m in mats.items():
    nr, nc = m.getSize()
    if m.nblocks() == 0:
        continue
    x = Vec().createWithArray(rng.uniform(-1, 1, nc)); y = Vec(nr)
    for _ in range(3):
        m.mult(x, y)
    reps = 20
    ctx.sync(); ctx.timer_start()
    for _ in range(reps):
        m.mult(x, y)
    ms = ctx.timer_stop() / reps
    stored, traffic = m.bytes()
    print(f"SpMV {name:7s} blocks {m.nblocks():10d} bs {m.getBlockSizes()} {ms*1e3:8.1f} us  {traffic/ms/1e6:8.1f} GB/s (alg bytes {traffic/1e6:.1f} MB)", flush=True)

X = 2 * np.pi * mesh.coords
vel = np.stack([np.cos(X[:, 0]) * np.sin(X[:, 1]) * np.sin(X[:, 2]), np.sin(X[:, 0]) * np.cos(X[:, 1]) * np.sin(X[:, 2]), -2 * np.sin(X[:, 0]) * np.sin(X[:, 1]) * np.cos(X[:, 2])], 1).ravel()
vort = np.stack([-6 * np.pi * np.sin(X[:, 0]) * np.cos(X[:, 1]) * np.cos(X[:, 2]), 6 * np.pi * np.cos(X[:, 0]) * np.sin(X[:, 1]) * np.cos(X[:, 2]), 0 * X[:, 0]], 1).ravel()
vbc = np.where(cls == 1, vel, 0.0)
for pcname in ("jacobi", "bjacobi"):
    for rtol in (1e-5, 1e-10):
        ksp = KSP().create(); ksp.setType("cg"); pc = PC(); pc.setType(pcname); ksp.setPC(pc)
        ksp.setOperators(mats["K"]); ksp.setTolerances(rtol=rtol, max_it=20000)
        rhs = mats["Rw"] * Vec().createWithArray(vort) + mats["Krhs"] * Vec().createWithArray(vbc)
        x = Vec().createWithArray(vbc)
        ksp.setUp(); ctx.sync(); ctx.timer_start()
        ksp.solve(rhs, x)
        ms = ctx.timer_stop()
        err = np.linalg.norm(x.getArray() - vel) / np.linalg.norm(vel)
        print(f"CG {pcname:8s} rtol {rtol:g}: its {ksp.its} reason {ksp.reason} {ms:.2f} ms ({ms/max(ksp.its,1)*1e3:.1f} us/it) err vs exact {err:.2e}", flush=True)
