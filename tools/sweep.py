"""BASELINE.json configs[4]: KLE assembly + KSP solve sweep on synthetic (perturbed) box meshes, one GPU.
python tools/sweep.py "nel,ngl nel,ngl ..."  ->  one JSON line per point (profiles/r1_sweep_n1.jsonl)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pynama_b200.device_mesh import DeviceMesh
from pynama_b200.domain.box_mesh import BoxMesh
from pynama_b200.petsc import Context, KSP, PC, PynamaError, Vec

points = [tuple(int(v) for v in p.split(",")) for p in (sys.argv[1] if len(sys.argv) > 1 else "16,4").split()]
ctx = Context.get()
PEAK = 6548.5
for nel, ngl in points:
    rec = {"nel": f"{nel}^3", "ngl": ngl}
    try:
        t0 = time.perf_counter()
        mesh = BoxMesh([nel] * 3, [0] * 3, [1] * 3, ngl, perturb=0.15)
        rec["dofs"] = int(mesh.nnodes * 3)
        rec["host_mesh_s"] = round(time.perf_counter() - t0, 2)
        dm = DeviceMesh(3, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
        cls = np.zeros(mesh.nnodes * 3, np.uint8)
        b = mesh.getBordersNodes()
        for d in range(3):
            cls[b * 3 + d] = 1
        dm.set_dof_class(cls)
        ctx.sync(); t0 = time.perf_counter()
        mats = dm.assemble_kle("FS")
        ctx.sync(); rec["assemble_kle_wall_ms"] = round((time.perf_counter() - t0) * 1e3, 1)
        rec["phase_ms"] = {p: round(dm.phase_ms(p), 2) for p in ("symbolic", "elem_kle", "kle_numeric")}
        K = mats["K"]
        n = mesh.nnodes * 3
        stored, traffic = K.bytes()
        rec["K_GB"] = round(stored / 1e9, 2)
        rec["nnz_ref"] = int(K.nnz()[0])
        rec["mem_GiB"] = round(ctx.mem_used() / 2**30, 1)
        rng = np.random.default_rng(12345)
        x = Vec().createWithArray(rng.uniform(-1, 1, n)); y = Vec(n)
        for mf in (False, True):
            if mf and ngl > 6:
                continue
            K.setMatFree(mf)
            for _ in range(3):
                K.mult(x, y)
            reps = 20
            ctx.sync(); ctx.timer_start()
            for _ in range(reps):
                K.mult(x, y)
            us = ctx.timer_stop() / reps * 1e3
            if mf:
                rec["matfree_us"] = round(us, 1)
            else:
                rec["spmv_us"] = round(us, 1)
                rec["spmv_GBs"] = round(traffic / us / 1e3, 1)
                rec["spmv_frac_of_hbm"] = round(traffic / us / 1e3 / PEAK, 3)
        xs = np.where(cls == 0, rng.uniform(-1, 1, n), 0.5)
        K.setMatFree(False)
        rhs = K * Vec().createWithArray(xs)
        for mf in (False, True):
            if mf and ngl > 6:
                continue
            K.setMatFree(mf)
            ksp = KSP().create(); ksp.setType("cg"); pc = PC().create(); pc.setType("bjacobi"); ksp.setPC(pc)
            ksp.setOperators(K); ksp.setTolerances(rtol=1e-8, max_it=50000); ksp.setUp()
            sol = Vec(n)
            ctx.sync(); ctx.timer_start()
            ksp.solve(rhs, sol)
            ms = ctx.timer_stop()
            rec["cg_matfree" if mf else "cg_stored"] = {"its": ksp.its, "ms": round(ms, 2), "us_per_it": round(ms / max(ksp.its, 1) * 1e3, 1),
                                                       "err": float(np.abs(sol.getArray() - xs).max())}
        K.setMatFree(False)
        del mats, K, x, y, rhs, sol, ksp, dm
    except (PynamaError, MemoryError, RuntimeError) as e:
        rec["error"] = str(e)[:200]
    print(json.dumps(rec), flush=True)
