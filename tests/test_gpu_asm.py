"""GPU parity of -pc_type asm (element-block additive Schwarz, pynama_b200/csrc/asm_pc.cu) through
the C ABI: same solutions as the direct solve of the oracle system, iteration counts of the
oracle's CG with the same preconditioner, fewer iterations than node-block Jacobi.

First run on a B200 in round 2 (gpurun_out r2a / r2b: all single-GPU cases green); since then part of
the normal GPU suite.  The kernels are also exercised on CPU threads by tests/test_asm_emulated.py."""
import os

import numpy as np
import pytest
import scipy.sparse.linalg as spl

from oracle import assembly, sem, timestep
from tests.util import fs_classes

pytestmark = pytest.mark.gpu


def _setup(nelem, ngl, perturb=0.0):
    from pynama_b200.device_mesh import DeviceMesh
    from pynama_b200.domain.box_mesh import BoxMesh
    dim = len(nelem)
    mesh = BoxMesh(nelem, [0.0] * dim, [1.0] * dim, ngl, perturb=perturb)
    dm = DeviceMesh(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
    cls = fs_classes(mesh)
    dm.set_dof_class(cls)
    mats = dm.assemble_kle("FS")
    want = assembly.assemble_kle(sem.SpectralTables(ngl, dim), mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    return mesh, dm, mats, want


def _ksp(K, ktype, pc, rtol, **kw):
    from pynama_b200.petsc import KSP, PC
    ksp = KSP().create()
    ksp.setType(ktype)
    p = PC().create()
    p.setType(pc)
    ksp.setPC(p)
    ksp.setOperators(K)
    ksp.setTolerances(rtol=rtol, max_it=5000)
    for name, val in kw.items():
        getattr(ksp, name)(*val)
    return ksp


@pytest.mark.parametrize("nelem,ngl,perturb", [([3, 3, 3], 3, 0.0), ([4, 4, 4], 4, 0.0), ([2, 1, 1], 5, 0.0),
                                                ([3, 3, 2], 3, 0.1), ([6, 5], 4, 0.0), ([5, 4], 3, 0.1)])
def test_cg_asm_vs_direct_and_oracle_iterations(nelem, ngl, perturb):
    from pynama_b200.petsc import Vec
    mesh, dm, mats, want = _setup(nelem, ngl, perturb)
    dim = len(nelem)
    n = mesh.nnodes * dim
    rng = np.random.default_rng(3)
    b = rng.uniform(-1, 1, n)
    Kref = want["K"].tocsr()
    ref = spl.spsolve(Kref.tocsc(), b)
    its = {}
    for pc in ("bjacobi", "asm"):
        ksp = _ksp(mats["K"], "cg", pc, 1e-11)
        x = Vec(n)
        ksp.solve(Vec().createWithArray(b), x)
        assert ksp.getConvergedReason() > 0, (pc, ksp.getConvergedReason(), ksp.getIterationNumber())
        assert np.linalg.norm(x.getArray() - ref) <= 1e-8 * np.linalg.norm(ref), pc
        its[pc] = ksp.getIterationNumber()
    assert its["asm"] < its["bjacobi"], its
    # the oracle's CG with the same preconditioner (scipy tests the true residual, PETSc's default
    # and ours the preconditioned one: a few iterations apart)
    prob = timestep.OracleProblem.__new__(timestep.OracleProblem)
    prob.K, prob.dim, prob.pc, prob.elem_nodes = Kref, dim, "asm", np.asarray(mesh.elem_nodes, dtype=np.int64)
    prob._setup_pc()
    count = [0]

    def cb(_):
        count[0] += 1
    spl.cg(Kref, b, rtol=1e-11, atol=0.0, maxiter=5000, M=prob._M, callback=cb)
    # first GPU run (r2a): 29 iterations here against scipy's 35 on the 3^3 / ngl=3 box -- the preconditioned
    # norm this solver tests (PETSc's default) falls below rtol a few iterations before the true one does
    assert 0.6 * count[0] - 4 <= its["asm"] <= 1.15 * count[0] + 4, (its, count[0])


def test_asm_is_reproducible_and_composes_with_the_other_options():
    from pynama_b200.petsc import Vec
    mesh, dm, mats, want = _setup([3, 3, 3], 3)
    n = mesh.nnodes * 3
    rng = np.random.default_rng(8)
    b = rng.uniform(-1, 1, n)
    ksp = _ksp(mats["K"], "cg", "asm", 1e-10)
    xs = []
    for _ in range(2):
        x = Vec(n)
        ksp.solve(Vec().createWithArray(b), x)
        xs.append((x.getArray(), ksp.getIterationNumber()))
    assert np.array_equal(xs[0][0], xs[1][0]) and xs[0][1] == xs[1][1]    # fixed summation orders
    # warm start from the solution: at most one iteration
    x = Vec().createWithArray(xs[0][0])
    ksp.solve(Vec().createWithArray(b), x)
    assert ksp.getIterationNumber() <= 1
    # the single-reduction request is dropped under asm, not an error
    ksp2 = _ksp(mats["K"], "cg", "asm", 1e-10, setCGSingleReduction=(True,))
    x2 = Vec(n)
    ksp2.solve(Vec().createWithArray(b), x2)
    assert ksp2.getConvergedReason() > 0 and np.array_equal(x2.getArray(), xs[0][0])
    # with the Fischer guess on a family of right-hand sides
    ksp3 = _ksp(mats["K"], "cg", "asm", 1e-10, setUseFischerGuess=(1, 4))
    lu = spl.splu(want["K"].tocsc())
    b1 = rng.uniform(-1, 1, n)
    x3 = Vec(n)
    for a in (1.0, 0.8, 0.6, 0.5, 0.4):
        bb = b + a * b1
        ksp3.solve(Vec().createWithArray(bb), x3)
        assert ksp3.getConvergedReason() > 0
        r = lu.solve(bb)
        assert np.linalg.norm(x3.getArray() - r) <= 1e-7 * np.linalg.norm(r)


@pytest.mark.parametrize("nelem,ngl,perturb", [([4, 4, 4], 3, 0.0), ([4, 4, 4], 4, 0.0), ([3, 3, 2], 3, 0.1),
                                                ([8, 6], 4, 0.0)])
def test_cg_two_level_asm_vs_direct(nelem, ngl, perturb):
    """-pc_asm_coarse 5: same solution, no more iterations than the one-level preconditioner"""
    from pynama_b200.petsc import Vec
    mesh, dm, mats, want = _setup(nelem, ngl, perturb)
    dim = len(nelem)
    n = mesh.nnodes * dim
    # a smooth right-hand side (what the time loop produces): the coarse space has something to do
    X = 2 * np.pi * mesh.coords
    xs = np.stack([np.sin(X[:, d]) * np.cos(X[:, (d + 1) % dim]) for d in range(dim)], 1).ravel()
    Kref = want["K"].tocsr()
    b = Kref @ xs
    its = {}
    for steps in (0, 5):
        ksp = _ksp(mats["K"], "cg", "asm", 1e-11, setASMCoarse=(steps,))
        x = Vec(n)
        ksp.solve(Vec().createWithArray(b), x)
        assert ksp.getConvergedReason() > 0, (steps, ksp.getConvergedReason(), ksp.getIterationNumber())
        assert np.linalg.norm(x.getArray() - xs) <= 1e-8 * np.linalg.norm(xs), steps
        its[steps] = ksp.getIterationNumber()
        x2 = Vec(n)                                            # reproducible
        ksp.solve(Vec().createWithArray(b), x2)
        assert np.array_equal(x.getArray(), x2.getArray())
    assert its[5] <= its[0] + 3, its            # (small meshes: little for the coarse space to do)


@pytest.mark.parametrize("nelem,ngl", [([4, 4, 4], 3), ([3, 4, 5], 4), ([7, 5], 4)])
def test_fused_coarse_level_equals_the_separate_launches(nelem, ngl):
    """one cooperative launch (restriction + all Chebyshev steps, a warp per coarse vertex) against the
    seven-launch path it replaces: same iterates up to the order of the row sums"""
    from pynama_b200 import _lib
    from pynama_b200.petsc import Vec
    mesh, dm, mats, want = _setup(nelem, ngl)
    n = mesh.nnodes * len(nelem)
    b = np.random.default_rng(21).uniform(-1, 1, n)
    res = {}
    for unfused in (1, 0):
        _lib.call("pyn_tune", b"asm_coarse_unfused", unfused)
        try:
            ksp = _ksp(mats["K"], "cg", "asm", 1e-11, setASMCoarse=(5,))
            x = Vec(n)
            ksp.solve(Vec().createWithArray(b), x)
            assert ksp.getConvergedReason() > 0
            res[unfused] = (x.getArray(), ksp.getIterationNumber())
            x2 = Vec(n)
            ksp.solve(Vec().createWithArray(b), x2)
            assert np.array_equal(x2.getArray(), res[unfused][0])            # reproducible
        finally:
            _lib.call("pyn_tune", b"asm_coarse_unfused", 0)
    assert abs(res[0][1] - res[1][1]) <= 1
    assert np.abs(res[0][0] - res[1][0]).max() <= 1e-8 * np.abs(res[1][0]).max()


@pytest.mark.parametrize("ktype", ["gmres", "bcgs"])
def test_gmres_and_bcgs_with_asm(ktype):
    from pynama_b200.petsc import Vec
    mesh, dm, mats, want = _setup([3, 3, 3], 3, 0.1)
    n = mesh.nnodes * 3
    b = np.random.default_rng(11).uniform(-1, 1, n)
    ref = spl.spsolve(want["K"].tocsc(), b)
    ksp = _ksp(mats["K"], ktype, "asm", 1e-12)
    x = Vec(n)
    ksp.solve(Vec().createWithArray(b), x)
    assert ksp.getConvergedReason() > 0
    assert np.linalg.norm(x.getArray() - ref) <= 1e-8 * np.linalg.norm(ref)


def test_asm_with_matrix_free_products():
    from pynama_b200.petsc import Vec
    mesh, dm, mats, want = _setup([3, 3, 3], 4)
    n = mesh.nnodes * 3
    b = np.random.default_rng(12).uniform(-1, 1, n)
    ref = spl.spsolve(want["K"].tocsc(), b)
    mats["K"].setMatFree(True)
    ksp = _ksp(mats["K"], "cg", "asm", 1e-11)
    x = Vec(n)
    ksp.solve(Vec().createWithArray(b), x)
    assert ksp.getConvergedReason() > 0
    assert np.linalg.norm(x.getArray() - ref) <= 1e-8 * np.linalg.norm(ref)


@pytest.mark.parametrize("case,nelem,ngl", [("taylor-green", [5, 4], 4), ("taylor-green3d", [3, 3, 3], 3)])
def test_time_steps_with_asm_match_the_oracle(case, nelem, ngl):
    """the bench path (BaseProblem / TsSolver / KleSolver) under -pc_type asm: fields after three
    RK5BS steps against the oracle's direct solves, as tests/test_gpu_timestep.py does for bjacobi"""
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.petsc import Options
    from pynama_b200.run_case import loadCase
    o = Options()
    for k, v in (("ksp_type", "cg"), ("pc_type", "asm"), ("ksp_rtol", 1e-13), ("ksp_max_it", 50000),
                 ("ts_adapt_type", "none"), ("ts_dt", 2e-3)):
        o.setValue(k, v)
    o.delValue("ksp_fischer_guess")
    try:
        fem = BaseProblem(loadCase(case), case=case, nelem=nelem, ngl=ngl)
        fem.setUp()
        fem.setUpSolver()
        fem.ts._post = None
        fem.ts.setSolution(fem.vort)
        mesh = fem.dom.getMesh()
        dim = len(nelem)
        orc = timestep.OracleProblem(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes(),
                                     rho=fem.rho, mu=fem.mu, solver="direct")
        t, k0 = 0.0, None
        for _ in range(3):
            fem.ts.step()
            t, k0 = orc.step(t, 2e-3, k0)
        rel = lambda a, b: np.abs(np.asarray(a) - b).max() / np.abs(b).max()     # noqa: E731
        assert rel(fem.vort.getArray(), orc.vort) < 1e-8
        assert rel(fem.solverKLE.getSolution().getArray(), orc.vel) < 1e-8
    finally:
        o.setValue("pc_type", "bjacobi")


@pytest.mark.parametrize("nproc,coarse", [(2, 5), (4, 5), (8, 3)])
def test_partitioned_run_with_two_level_asm(nproc, coarse):
    """row-partitioned operator with the coarse level: global coarse tables on every rank (pyn_mesh_set_global_tables),
    this rank's part of P^T r summed over the ranks, the Chebyshev steps replicated; same oracle comparison"""
    import subprocess
    import sys
    from tests.test_gpu_dist import SCRIPT, _ngpus
    if _ngpus() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    env = dict(os.environ, DIST_OPTS=f"pc_type=asm,pc_asm_coarse={coarse}", DIST_NELEM="6,6,6", DIST_KSP="cg")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(29590 + nproc), SCRIPT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
    assert "OK" in r.stdout


@pytest.mark.parametrize("nproc,ksp", [(2, "cg"), (2, "gmres"), (4, "cg"), (8, "cg")])
def test_partitioned_run_with_asm(nproc, ksp):
    """row-partitioned operator: ghost refresh of r, all local elements, owned nodes summed, blocks of
    elements with ghost nodes borrowed by signature; fields after the steps against the oracle
    (tests/dist_check.py).  Blocks are 3 elements thick so that every class has a complete twin."""
    import subprocess
    import sys
    from tests.test_gpu_dist import SCRIPT, _ngpus
    if _ngpus() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    env = dict(os.environ, DIST_OPTS="pc_type=asm", DIST_NELEM="6,6,6", DIST_KSP=ksp)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(29570 + nproc), SCRIPT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
    assert "OK" in r.stdout
