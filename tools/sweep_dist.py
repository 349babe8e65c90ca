"""BASELINE.json configs[4]: KLE assembly + KSP solve strong / weak scaling sweep on synthetic box meshes,
1 to 8 GPUs of one box (row partition, one process per GPU).

    python tools/sweep_dist.py "64,64,64,3 32,32,32,4"                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29650 tools/sweep_dist.py "128,128,128,3"

Every point "nx,ny,nz,ngl" prints ONE JSON line from rank 0.  A point whose assembled operators do not
fit the GPUs of the run (SURVEY.md section 8 feasibility table) is printed as
{"status": "exceeds box memory", ...} with the estimate, not skipped and not shrunk.  Meshes are uniform
boxes (one congruence class: the element block store of a perturbed 128^3 mesh alone would need 290 GB).
Times are CUDA-event times, max over ranks; SpMV bytes are the blocks of the rows each rank owns.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from pynama_b200.device_mesh import DeviceMesh  # noqa: E402
from pynama_b200.domain.box_mesh import BoxMesh  # noqa: E402
from pynama_b200.petsc import COMM_WORLD, KSP, PC, Context, Layout, PynamaError, Vec  # noqa: E402

GPU_BYTES = 120e9          # what the estimate may reach per B200: r2k showed 128^3 / ngl=4 on 8 GPUs (estimate
                           # 161 GB) failing in a 66 GB allocation -- graphs, the halo-layer elements and the
                           # transient buffers of the symbolic phase are not in the estimate
PEAK = 6548.8


def load_peak():
    try:
        with open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return PEAK


def estimate_bytes(nelem, ngl, size):
    """per-GPU bytes of K, Krhs (~ boundary only), Rw in node-block BSR + graphs, from the pair count
    prod(nel_d (p+1)^2 - (nel_d - 1)) of SURVEY.md section 8 (one extra element layer per rank ignored)"""
    pairs = float(np.prod([n * ngl * ngl - (n - 1) for n in nelem]))
    kle = pairs * (2 * (9 * 8 + 4))                       # K and Rw: 3x3 blocks + column id
    graph = pairs * (4 + 8 + 8) + float(np.prod(nelem)) * ngl ** 6 * 8       # slots, segment pointers, contributions
    return (kle + graph) / size


def main():
    pts = [tuple(int(v) for v in p.split(",")) for p in (sys.argv[1] if len(sys.argv) > 1 else "16,16,16,4").split()]
    ctx = Context.get()
    rank, size = COMM_WORLD.rank, COMM_WORLD.size
    peak = load_peak()
    mx = lambda v: float(ctx.allreduce([v], "max")[0]) if size > 1 else float(v)      # noqa: E731
    sm = lambda v: float(ctx.allreduce([v], "sum")[0]) if size > 1 else float(v)      # noqa: E731
    for nx, ny, nz, ngl in pts:
        nelem = [nx, ny, nz]
        rec = {"nelem": nelem, "ngl": ngl, "gpus": size}
        est = estimate_bytes(nelem, ngl, size)
        rec["estimated_GB_per_gpu"] = round(est / 1e9, 1)
        if est > GPU_BYTES:
            rec["status"] = "exceeds box memory"
            rec["dofs"] = int(np.prod([n * (ngl - 1) + 1 for n in nelem]) * 3)
            if rank == 0:
                print(json.dumps(rec), flush=True)
            continue
        try:
            t0 = time.perf_counter()
            mesh = BoxMesh(nelem, [0.0] * 3, [float(n) / max(nelem) for n in nelem], ngl)
            rec["dofs"] = int(mesh.nnodes * 3)
            cls = np.zeros(mesh.nnodes * 3, np.uint8)
            b = mesh.getBordersNodes()
            for d in range(3):
                cls[b * 3 + d] = 1
            layout = None
            if size > 1:
                from pynama_b200.parallel import Partition
                part = Partition(mesh, rank, size)
                layout = Layout(part, mesh.nnodes)
                dm = DeviceMesh(3, ngl, part.corner_xyz, part.elem_nodes, part.n_loc, layout=layout)
                ldofs = (part.l2g[:, None] * 3 + np.arange(3)[None, :]).ravel()
                cls_l = cls[ldofs]
                rec["partition"] = "x".join(str(p) for p in part.pgrid) if hasattr(part, "pgrid") else "rcb"
                nown = part.n_own
            else:
                dm = DeviceMesh(3, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
                cls_l = cls
                nown = mesh.nnodes
            rec["host_mesh_s"] = round(mx(time.perf_counter() - t0), 2)
            dm.set_dof_class(cls_l)
            ctx.sync()
            t0 = time.perf_counter()
            mats = dm.assemble_kle("FS")
            ctx.sync()
            rec["assemble_kle_wall_ms"] = round(mx((time.perf_counter() - t0) * 1e3), 1)
            rec["phase_ms"] = {p: round(mx(dm.phase_ms(p)), 2) for p in ("symbolic", "elem_kle", "kle_numeric")}
            rec["assemble_Mdof_per_s"] = round(rec["dofs"] / rec["assemble_kle_wall_ms"] / 1e3, 1)
            K = mats["K"]
            stored, traffic = K.bytes()
            rec["K_GB_total"] = round(sm(stored) / 1e9, 2)
            rec["mem_GiB_per_gpu_max"] = round(mx(ctx.mem_used()) / 2 ** 30, 1)
            rng = np.random.default_rng(12345 + rank)
            x, y = K.createVecRight(), K.createVecLeft()
            x.setArray(rng.uniform(-1, 1, nown * 3))
            for _ in range(3):
                K.mult(x, y)
            reps = 20
            ctx.sync()
            ctx.timer_start()
            for _ in range(reps):
                K.mult(x, y)
            us = mx(ctx.timer_stop() / reps * 1e3)
            rec["spmv_us"] = round(us, 1)
            rec["spmv_GBs_aggregate"] = round(sm(traffic) / us / 1e3, 1)
            rec["spmv_frac_of_hbm_per_gpu"] = round(sm(traffic) / size / us / 1e3 / peak, 3)
            xs = np.where(cls_l[:nown * 3] == 0, rng.uniform(-1, 1, nown * 3), 0.5)
            xv = K.createVecRight()
            xv.setArray(xs)
            rhs = K * xv
            for mf in (False, True):
                if mf and ngl > 6:
                    continue
                K.setMatFree(mf)
                ksp = KSP().create()
                ksp.setType("cg")
                pc = PC().create()
                pc.setType("bjacobi")
                ksp.setPC(pc)
                ksp.setCGSingleReduction(size > 1)
                ksp.setOperators(K)
                ksp.setTolerances(rtol=1e-8, max_it=50000)
                ksp.setUp()
                sol = K.createVecRight()
                ctx.sync()
                ctx.timer_start()
                ksp.solve(rhs, sol)
                ms = mx(ctx.timer_stop())
                err = mx(float(np.abs(sol.getArray() - xs).max()))
                rec["cg_matfree" if mf else "cg_stored"] = {"its": ksp.its, "ms": round(ms, 2),
                                                           "us_per_it": round(ms / max(ksp.its, 1) * 1e3, 1), "err": err}
                ksp.destroy()
            K.setMatFree(False)
            rec["status"] = "ok"
            del mats, K, x, y, xv, rhs, sol, dm
        except (PynamaError, MemoryError, RuntimeError) as e:
            rec["status"] = "error"
            rec["error"] = str(e)[:300]
        if rank == 0:
            print(json.dumps(rec), flush=True)
    if size > 1:
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.barrier()
                dist.destroy_process_group()
        except Exception:
            pass


if __name__ == "__main__":
    main()
