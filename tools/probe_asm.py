"""Assembly phase timings on one GPU (development aid): python tools/probe_asm.py [nel] [ngl] [reps] [perturb]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pynama_b200.device_mesh import DeviceMesh
from pynama_b200.domain.box_mesh import BoxMesh
from pynama_b200.petsc import Context

nel = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ngl = int(sys.argv[2]) if len(sys.argv) > 2 else 4
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
perturb = float(sys.argv[4]) if len(sys.argv) > 4 else 0.15
dim = 3
mesh = BoxMesh([nel] * 3, [0] * 3, [1] * 3, ngl, perturb=perturb)
ctx = Context.get()
dm = DeviceMesh(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
cls = np.zeros(mesh.nnodes * dim, np.uint8)
b = mesh.getBordersNodes()
for d in range(dim):
    cls[b * dim + d] = 1
dm.set_dof_class(cls)
print(f"{nel}^3 ngl={ngl} perturb={perturb}: {dm.congruence()[1]} congruence classes, symbolic {dm.phase_ms('symbolic'):.2f} ms", flush=True)
for r in range(reps):
    ctx.prof_enable(True)
    mats = dm.assemble_kle("FS")
    ops = dm.assemble_ops(compact=True)
    ctx.sync()
    prof = {k: ctx.prof_get(k) for k in ("elem_kle", "numeric")}
    ctx.prof_enable(False)
    print(f"rep {r}: elem_kle {dm.phase_ms('elem_kle'):.3f} kle_numeric {dm.phase_ms('kle_numeric'):.3f} "
          f"elem_ops {dm.phase_ms('elem_ops'):.3f} ops_numeric {dm.phase_ms('ops_numeric'):.3f} ms | prof {prof}", flush=True)
    del mats, ops
