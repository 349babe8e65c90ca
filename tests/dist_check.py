"""Multi-GPU parity check, one process per GPU (launched by tests/test_gpu_dist.py or by hand):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tests/dist_check.py

Row-partitioned assembly, SpMV with halo exchange, CG with allreduced inner products, the
vorticity-transport RHS and RK5BS steps are compared with the CPU oracle on the GLOBAL mesh.
Exit code 0 = all ranks agree with the oracle."""
import os
import sys

import numpy as np
import scipy.sparse.linalg as spl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import assembly, sem, timestep                      # noqa: E402
from pynama_b200.petsc import COMM_WORLD, Context, Options      # noqa: E402


def main():
    nelem = [int(a) for a in os.environ.get("DIST_NELEM", "4,4,4").split(",")]
    ngl = int(os.environ.get("DIST_NGL", "3"))
    rank, size = COMM_WORLD.rank, COMM_WORLD.size
    ctx = Context.get()
    import bench
    from pynama_b200.cases.base_problem import BaseProblem
    o = Options()
    for k, v in (("ksp_type", os.environ.get("DIST_KSP", "cg")), ("pc_type", "bjacobi"), ("ksp_rtol", 1e-13), ("ksp_max_it", 20000),
                 ("ts_adapt_type", "none"), ("ts_dt", 1e-3)):
        o.setValue(k, v)
    if os.environ.get("DIST_FISCHER", "0") != "0":
        o.setValue("ksp_fischer_guess", "1," + os.environ["DIST_FISCHER"])
    for name in ("ksp_cg_single_reduction", "kle_matfree"):          # DIST_OPTS="name=value,name=value"
        o.delValue(name)
    for kv in filter(None, os.environ.get("DIST_OPTS", "").split(",")):
        name, val = kv.split("=")
        o.setValue(name, val)
    case = os.environ.get("DIST_CASE", "taylor-green3d")
    if case in ("cavity", "cylinder"):
        return main_cavity(nelem, ngl, rank, size, ctx, case)
    cfg = bench.case_config(nelem[0], ngl)
    cfg["domain"]["box-mesh"]["nelem"] = nelem
    if os.environ.get("DIST_GMSH", "0") != "0":
        # the same box as an unstructured .msh file (scrambled numbering, rotated and mirrored
        # cells): recursive-coordinate-bisection partition; every rank writes its own identical copy
        import tempfile
        from tests.test_gmsh_mesh import box_as_msh
        path = os.path.join(tempfile.mkdtemp(prefix=f"pyn_gmsh_{rank}_"), "mesh.msh")
        box_as_msh(path, nelem, perturb=0.1)
        del cfg["domain"]["box-mesh"]
        cfg["domain"]["gmsh-file"] = path
    fem = BaseProblem(cfg, case="taylor-green3d")
    fem.setUp()
    fem.setUpSolver()
    fem.ts._post = None
    fem.ts.setSolution(fem.vort)
    mesh = fem.dom.getMesh()
    dim = 3
    part = getattr(fem.dom, "_partition", None)
    assert (part is not None) == (size > 1)

    # ---- oracle on the global mesh
    orc = timestep.OracleProblem(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.coords,
                                 mesh.getBordersNodes(), rho=fem.rho, mu=fem.mu, solver="direct")
    rng = np.random.default_rng(7)
    fails = []

    def check(name, got, want, tol):
        err = np.abs(got - want).max() / max(np.abs(want).max(), 1e-300)
        if rank == 0:
            print(f"[dist_check] {name:28s} rel.err {err:.2e} (tol {tol:g})", flush=True)
        if not err <= tol:
            fails.append(name)

    # ---- SpMV of every matrix on a random global vector
    mats = {"K": (fem.mat.K, orc.mats["K"]), "Krhs": (fem.mat.Krhs, orc.mats["Krhs"]),
            "Rw": (fem.mat.Rw, orc.mats["Rw"]), "Curl": (fem.operator.Curl, orc.ops["Curl"]),
            "SrT": (fem.operator.SrT, orc.ops["SrT"]), "DivSrT": (fem.operator.DivSrT, orc.ops["DivSrT"])}
    for name, (m, ref) in mats.items():
        xg = rng.uniform(-1, 1, ref.shape[1])
        x = m.createVecRight()
        x.setValues(np.arange(ref.shape[1]), xg)
        y = m.createVecLeft()
        m.mult(x, y)
        assert m.getSize() == ref.shape, (name, m.getSize(), ref.shape)
        check(f"SpMV {name}", y.getGlobalArray(), ref @ xg, 1e-12)
        # distributed reductions
        d = x.dot(x)
        check(f"dot {name}", np.array([d]), np.array([xg @ xg]), 1e-13)

    # ---- KLE solve and RHS at t = 0, then two RK5BS steps
    f = fem.vort.duplicate()
    fem.evalRHS(fem.ts, 0.0, fem.vort, f)
    fo = orc.eval_rhs(0.0)
    check("KLE solve", fem.solverKLE.getSolution().getGlobalArray(), orc.vel, 1e-9)
    check("evalRHS", f.getGlobalArray(), fo, 1e-8)
    t, k0 = 0.0, None
    for _ in range(2):
        fem.ts.step()
        t, k0 = orc.step(t, 1e-3, k0)
    check("vorticity after 2 steps", fem.vort.getGlobalArray(), orc.vort, 1e-8)
    check("velocity after 2 steps", fem.solverKLE.getSolution().getGlobalArray(), orc.vel, 1e-8)
    nf = ctx.allreduce([len(fails)], "max")[0]
    if rank == 0:
        print(f"[dist_check] ranks {size}, collectives issued on rank 0: {ctx.collectives()}, "
              f"{'OK' if nf == 0 else 'FAILED ' + str(fails)}", flush=True)
    return 0 if nf == 0 else 1


def main_cavity(nelem, ngl, rank, size, ctx, case="cavity"):
    """no-slip lid-driven cavity (MatNS, the two-solve branch of evalRHS) on the partitioned mesh
    against OracleProblemNS on the global mesh.  case == "cylinder": BASELINE.json configs[3] on a
    small channel (blocked-cell cylinder, unstructured file, 8 borders, RCB partition)"""
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.run_case import loadCase
    Options().setValue("ksp_type", os.environ.get("DIST_KSP", "gmres"))
    cfg = loadCase(case)
    kw = {"nelem": nelem}
    if case == "cylinder":
        import tempfile
        from pynama_b200.domain.cylinder_mesh import write_cylinder_msh
        path = os.path.join(tempfile.mkdtemp(prefix=f"pyn_cyl_{rank}_"), "cyl.msh")      # every rank its own identical copy
        write_cylinder_msh(path, h=0.5, lower=(-2.0, -1.5, 0.0), upper=(3.0, 1.5, 1.0))
        cfg["domain"]["gmsh-file"] = path
        kw = {}
    fem = BaseProblem(cfg, case=case, ngl=ngl, **kw)
    fem.setUp()
    fem.setUpSolver()
    fem.ts._post = None
    fem.ts.setSolution(fem.vort)
    mesh = fem.dom.getMesh()
    walls = cfg["boundary-conditions"]["no-slip"]
    sides = list(walls.keys())
    orc = timestep.OracleProblemNS(3, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, sides,
                                   {s: mesh.getBorderNodes(s) for s in sides}, walls, rho=fem.rho, mu=fem.mu)
    fails = []

    def check(name, got, want, tol):
        err = np.abs(got - want).max() / max(np.abs(want).max(), 1e-300)
        if rank == 0:
            print(f"[dist_check] {name:28s} rel.err {err:.2e} (tol {tol:g})", flush=True)
        if not err <= tol:
            fails.append(name)

    t, k0 = 0.0, None
    for _ in range(2):
        fem.ts.step()
        t, k0 = orc.step(t, 1e-3, k0)
    check("cavity velocity", fem.solverKLE.getSolution().getGlobalArray(), orc.vel, 1e-8)
    check("cavity free-slip velocity", fem.solverKLE.getFreeSlipSolution().getGlobalArray(), orc.velfs, 1e-8)
    check("cavity vorticity", fem.vort.getGlobalArray(), orc.vort, 1e-8)
    nf = ctx.allreduce([len(fails)], "max")[0]
    if rank == 0:
        print(f"[dist_check] {case}, ranks {size}: {'OK' if nf == 0 else 'FAILED ' + str(fails)}", flush=True)
    return 0 if nf == 0 else 1


if __name__ == "__main__":
    rc = main()
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
    except Exception:
        pass
    sys.exit(rc)
