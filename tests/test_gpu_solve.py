"""GPU parity: Vec operations, SpMV (k6), the KLE solve (k7) and the pointwise RHS
pieces (k8, k9) through the C ABI.  Includes the reference's own known-answer
tests: KAT-1/KAT-2 (src/tests/test_solver.py:7-37) and KAT-3 (:45-76)."""
import numpy as np
import pytest
import scipy.sparse.linalg as spl

from oracle import assembly, sem
from tests.util import csr_of, fs_classes, ns_classes

pytestmark = pytest.mark.gpu


def _setup(nelem, ngl, perturb=0.0):
    from pynama_b200.device_mesh import DeviceMesh
    from pynama_b200.domain.box_mesh import BoxMesh
    dim = len(nelem)
    mesh = BoxMesh(nelem, [0.0] * dim, [1.0] * dim, ngl, perturb=perturb)
    dm = DeviceMesh(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
    return mesh, dm, sem.SpectralTables(ngl, dim)


def test_vec_operations():
    from pynama_b200.petsc import Vec
    rng = np.random.default_rng(1)
    n = 100003
    a, b = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
    x, y = Vec().createWithArray(a), Vec().createWithArray(b)
    assert abs(x.dot(y) - a @ b) <= 1e-12 * n ** 0.5
    assert abs(x.norm() - np.linalg.norm(a)) <= 1e-12 * n ** 0.5
    assert abs(x.sum() - a.sum()) <= 1e-11
    y.axpy(2.5, x)
    b = b + 2.5 * a
    assert np.allclose(y.getArray(), b, rtol=0, atol=1e-15)
    y.aypx(-0.5, x)
    b = -0.5 * b + a
    assert np.allclose(y.getArray(), b, rtol=0, atol=1e-15)
    y *= 3.0
    b *= 3.0
    y.scale(0.25)
    b *= 0.25
    assert np.allclose(y.getArray(), b, rtol=0, atol=1e-15)
    w = x * y
    assert np.allclose(w.getArray(), a * b, rtol=0, atol=1e-15)
    d = x - y
    assert np.allclose(d.getArray(), a - b, rtol=0, atol=1e-15)
    w.waxpy(1.5, x, y)
    assert np.allclose(w.getArray(), 1.5 * a + b, rtol=0, atol=1e-15)
    r = x.copy()
    r.reciprocal()
    assert np.allclose(r.getArray(), 1.0 / a, rtol=1e-15)
    idx = rng.choice(n, 500, replace=False)
    vals = rng.uniform(-1, 1, 500)
    a = x.getArray()
    x.setValues(idx, vals)
    a[idx] = vals
    assert np.array_equal(x.getArray(), a)
    x.setValues(idx, vals, addv=True)
    a[idx] += vals
    assert np.array_equal(x.getArray(), a)
    assert np.array_equal(x.getValues(idx[:10]), a[idx[:10]])
    vs = [Vec().createWithArray(rng.uniform(-1, 1, n)) for _ in range(7)]
    al = rng.uniform(-1, 1, 7)
    ref = a + sum(c * v.getArray() for c, v in zip(al, vs))
    x.maxpy(al, vs)
    assert np.allclose(x.getArray(), ref, rtol=0, atol=1e-14)
    z = Vec(0)
    assert z.norm() == 0.0


def test_vtensv_kat3():
    """computeVtensV on [1..18] (src/tests/test_solver.py:57-76)"""
    from pynama_b200 import _lib
    from pynama_b200.petsc import Vec
    v = Vec().createWithArray(np.arange(1, 19, dtype=float))
    out = Vec(27)
    _lib.call("pyn_vtensv", v._h, out._h, 2)
    ref = []
    for k in range(9):
        a, b = 2 * k + 1, 2 * k + 2
        ref += [a * a, a * b, b * b]
    np.testing.assert_array_almost_equal(np.array(ref, dtype=float), out.getArray(), decimal=10)
    # 3-D packing [xx, xy, yy, yz, zz, zx] (base_problem.py:146-153)
    v3 = Vec().createWithArray(np.array([1.0, 2.0, 3.0, -1.0, 0.5, 4.0]))
    o3 = Vec(12)
    _lib.call("pyn_vtensv", v3._h, o3._h, 3)
    assert np.array_equal(o3.getArray(), [1, 2, 4, 6, 9, 3, 1, -0.5, 0.25, 2.0, 16.0, -4.0])
    # aux <- 2 mu aux - rho v(x)v  (base_problem.py:129-130)
    aux = Vec().createWithArray(np.arange(12, dtype=float))
    _lib.call("pyn_stress_combine", aux._h, v3._h, 0.01, 0.5, 3)
    assert np.allclose(aux.getArray(), 0.02 * np.arange(12) - 0.5 * o3.getArray(), rtol=0, atol=1e-15)


@pytest.mark.parametrize("nelem,ngl,mode", [([5, 4], 3, "FS"), ([3, 3], 6, "FS"), ([3, 3, 3], 3, "FS"),
                                            ([2, 3, 2], 4, "NS"), ([2, 2], 2, "NS"),
                                            ([6, 5, 4], 4, "FS")])
def test_spmv_all_matrices(nelem, ngl, mode):
    from pynama_b200.petsc import Vec
    mesh, dm, tab = _setup(nelem, ngl, 0.1)
    cls = fs_classes(mesh) if mode == "FS" else ns_classes(mesh)
    dm.set_dof_class(cls)
    mats = dm.assemble_kle(mode)
    c, s, d, _ = dm.assemble_ops(compact=True)
    mats.update({"Curl": c, "SrT": s, "DivSrT": d})
    rng = np.random.default_rng(12345)
    for name, m in mats.items():
        ref = csr_of(m)
        x = rng.uniform(-1, 1, ref.shape[1])
        y = (m * Vec().createWithArray(x)).getArray()
        want = ref @ x
        assert np.abs(y - want).max() <= 1e-12 * max(np.abs(want).max(), 1e-300), name
        # y = v + A x
        v = rng.uniform(-1, 1, ref.shape[0])
        out = Vec(ref.shape[0])
        m.multAdd(Vec().createWithArray(x), Vec().createWithArray(v), out)
        assert np.abs(out.getArray() - (want + v)).max() <= 1e-12 * max(np.abs(want + v).max(), 1e-300), name


def _solve(dm, mats, vort, vel_bc, pc="bjacobi", rtol=1e-13):
    from pynama_b200.petsc import KSP, PC, Vec
    ksp = KSP().create()
    ksp.setType("cg")
    p = PC().create()
    p.setType(pc)
    ksp.setPC(p)
    ksp.setOperators(mats["K"])
    ksp.setTolerances(rtol=rtol, max_it=20000)
    b = mats["Rw"] * Vec().createWithArray(vort) + mats["Krhs"] * Vec().createWithArray(vel_bc)
    x = Vec().createWithArray(vel_bc)
    ksp.solve(b, x)
    assert ksp.getConvergedReason() > 0, (ksp.getConvergedReason(), ksp.getIterationNumber())
    return x.getArray(), ksp


@pytest.mark.parametrize("pc", ["none", "jacobi", "bjacobi"])
def test_kat1_uniform_flow(pc):
    """uniform flow 3x3, ngl=3, omega=0, v_bc=(4,0): ||v - (4,0)|| < 1e-12 (test_solver.py:7-25)"""
    mesh, dm, _ = _setup([3, 3], 3)
    cls = fs_classes(mesh)
    dm.set_dof_class(cls)
    mats = dm.assemble_kle("FS")
    exact = np.tile([4.0, 0.0], mesh.nnodes)
    vbc = np.where(cls == 1, exact, 0.0)
    x, _ = _solve(dm, mats, np.zeros(mesh.nnodes), vbc, pc=pc, rtol=1e-15)
    assert np.linalg.norm(x - exact) < 1e-12


def test_kat2_taylor_green():
    """Taylor-Green 10x10, ngl=5, t=0: ||v - v_exact|| < 1e-5 (test_solver.py:27-37)"""
    mesh, dm, tab = _setup([10, 10], 5)
    cls = fs_classes(mesh)
    dm.set_dof_class(cls)
    mats = dm.assemble_kle("FS")
    x_, y_ = 2 * np.pi * mesh.coords[:, 0], 2 * np.pi * mesh.coords[:, 1]
    vel = np.stack([np.cos(x_) * np.sin(y_), -np.sin(x_) * np.cos(y_)], 1).ravel()
    vort = -4 * np.pi * np.cos(x_) * np.cos(y_)
    vbc = np.where(cls == 1, vel, 0.0)
    x, ksp = _solve(dm, mats, vort, vbc, rtol=1e-12)
    assert np.linalg.norm(x - vel) < 1e-5
    # and the same discrete solution as the reference path (direct solve of the oracle system)
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    ref = spl.spsolve(want["K"].tocsc(), want["Rw"] @ vort + want["Krhs"] @ vbc)
    assert np.linalg.norm(x - ref) <= 1e-9 * np.linalg.norm(ref)


def test_kle_solve_3d_vs_oracle():
    mesh, dm, tab = _setup([3, 3, 3], 4, 0.1)
    cls = fs_classes(mesh)
    dm.set_dof_class(cls)
    mats = dm.assemble_kle("FS")
    X = 2 * np.pi * mesh.coords
    vel = np.stack([np.cos(X[:, 0]) * np.sin(X[:, 1]) * np.sin(X[:, 2]),
                    np.sin(X[:, 0]) * np.cos(X[:, 1]) * np.sin(X[:, 2]),
                    -2 * np.sin(X[:, 0]) * np.sin(X[:, 1]) * np.cos(X[:, 2])], 1).ravel()
    vort = np.stack([-6 * np.pi * np.sin(X[:, 0]) * np.cos(X[:, 1]) * np.cos(X[:, 2]),
                     6 * np.pi * np.cos(X[:, 0]) * np.sin(X[:, 1]) * np.cos(X[:, 2]),
                     0 * X[:, 0]], 1).ravel()
    vbc = np.where(cls == 1, vel, 0.0)
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    ref = spl.spsolve(want["K"].tocsc(), want["Rw"] @ vort + want["Krhs"] @ vbc)
    for pc in ("jacobi", "bjacobi"):
        x, ksp = _solve(dm, mats, vort, vbc, pc=pc, rtol=1e-12)
        assert np.linalg.norm(x - ref) <= 1e-9 * np.linalg.norm(ref), pc
    # warm start from the solution converges at once
    from pynama_b200.petsc import Vec
    b = mats["Rw"] * Vec().createWithArray(vort) + mats["Krhs"] * Vec().createWithArray(vbc)
    xs = Vec().createWithArray(x)
    ksp.setTolerances(rtol=1e-8)
    ksp.solve(b, xs)
    assert ksp.getIterationNumber() <= 1
    # max_it is honoured and reported as DIVERGED_ITS
    ksp.setTolerances(rtol=1e-14, max_it=3)
    ksp._drop()
    x0 = Vec().createWithArray(vbc)
    ksp.solve(b, x0)
    assert ksp.getConvergedReason() == -3 and ksp.getIterationNumber() == 3



def test_fischer_guess_on_a_family_of_right_hand_sides():
    """-ksp_fischer_guess 1,4 (KSPSetUseFischerGuess): same solutions as the direct solve, far
    fewer iterations once the right-hand sides repeat inside a low-dimensional family"""
    from pynama_b200.petsc import KSP, PC, Vec
    mesh, dm, tab = _setup([3, 3, 3], 3, 0.1)
    cls = fs_classes(mesh)
    dm.set_dof_class(cls)
    mats = dm.assemble_kle("FS")
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    lu = spl.splu(want["K"].tocsc())
    rng = np.random.default_rng(5)
    b0, b1 = rng.uniform(-1, 1, mesh.nnodes * 3), rng.uniform(-1, 1, mesh.nnodes * 3)
    its = {}
    for fischer in (0, 4):
        ksp = KSP().create()
        ksp.setType("cg")
        p = PC().create()
        p.setType("bjacobi")
        ksp.setPC(p)
        ksp.setOperators(mats["K"])
        ksp.setTolerances(rtol=1e-11, max_it=5000)
        if fischer:
            ksp.setUseFischerGuess(1, fischer)
        x = Vec(mesh.nnodes * 3)
        its[fischer] = []
        for a in (1.0, 0.9, 0.8, 0.7, 0.6, 0.5, 0.45, 0.4, 0.35, 0.3):
            b = b0 + a * b1
            ksp.solve(Vec().createWithArray(b), x)
            assert ksp.getConvergedReason() > 0
            ref = lu.solve(b)
            assert np.linalg.norm(x.getArray() - ref) <= 1e-8 * np.linalg.norm(ref), (fischer, a)
            its[fischer].append(ksp.getIterationNumber())
    assert sum(its[4][2:]) < 0.3 * sum(its[0][2:]), its


@pytest.mark.parametrize("ktype", ["gmres", "bcgs"])
@pytest.mark.parametrize("pc", ["jacobi", "bjacobi"])
def test_gmres_and_bcgs_vs_direct(ktype, pc):
    """-ksp_type gmres / bcgs (the reference's KspSolver default type is gmres, kle_solver.py:58)"""
    from pynama_b200.petsc import KSP, PC, Vec
    mesh, dm, tab = _setup([3, 3, 3], 3, 0.1)
    cls = fs_classes(mesh)
    dm.set_dof_class(cls)
    mats = dm.assemble_kle("FS")
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    rng = np.random.default_rng(11)
    b = rng.uniform(-1, 1, mesh.nnodes * 3)
    ref = spl.spsolve(want["K"].tocsc(), b)
    ksp = KSP().create()
    ksp.setType(ktype)
    p = PC().create()
    p.setType(pc)
    ksp.setPC(p)
    ksp.setOperators(mats["K"])
    ksp.setTolerances(rtol=1e-12, max_it=4000)
    x = Vec(mesh.nnodes * 3)
    ksp.solve(Vec().createWithArray(b), x)
    assert ksp.getConvergedReason() > 0, (ksp.getConvergedReason(), ksp.getIterationNumber())
    assert np.linalg.norm(x.getArray() - ref) <= 1e-8 * np.linalg.norm(ref)
    # 2-D (2x2 blocks) too
    mesh2, dm2, tab2 = _setup([4, 4], 4)
    cls2 = fs_classes(mesh2)
    dm2.set_dof_class(cls2)
    m2 = dm2.assemble_kle("FS")
    w2 = assembly.assemble_kle(tab2, mesh2.corner_xyz, mesh2.elem_nodes, mesh2.nnodes, cls2, "FS")
    b2 = rng.uniform(-1, 1, mesh2.nnodes * 2)
    ksp.setOperators(m2["K"])
    x2 = Vec(mesh2.nnodes * 2)
    ksp.solve(Vec().createWithArray(b2), x2)
    assert ksp.getConvergedReason() > 0
    ref2 = spl.spsolve(w2["K"].tocsc(), b2)
    assert np.linalg.norm(x2.getArray() - ref2) <= 1e-8 * np.linalg.norm(ref2)


@pytest.mark.parametrize("pc", ["jacobi", "bjacobi", "none"])
def test_cg_single_reduction_matches_cg(pc):
    """-ksp_cg_single_reduction (KSPCGUseSingleReduction): same solution and iteration count (+-2)
    as the two-reduction recurrence, 3-D and 2-D"""
    from pynama_b200.petsc import KSP, PC, Vec
    for nelem, ngl in (([4, 4, 3], 4), ([6, 5], 5)):
        mesh, dm, tab = _setup(nelem, ngl, 0.1)
        dim = len(nelem)
        cls = fs_classes(mesh)
        dm.set_dof_class(cls)
        K = dm.assemble_kle("FS")["K"]
        rng = np.random.default_rng(5)
        xs = np.where(cls == 0, rng.uniform(-1, 1, mesh.nnodes * dim), -0.25)
        b = K * Vec().createWithArray(xs)
        res = {}
        for sr in (False, True):
            ksp = KSP().create()
            ksp.setType("cg")
            p = PC().create()
            p.setType(pc)
            ksp.setPC(p)
            ksp.setCGSingleReduction(sr)
            ksp.setOperators(K)
            ksp.setTolerances(rtol=1e-11, max_it=20000)
            x = Vec(mesh.nnodes * dim)
            ksp.solve(b, x)
            assert ksp.getConvergedReason() == 2, (sr, ksp.getConvergedReason())
            res[sr] = (x.getArray(), ksp.getIterationNumber(), ksp.getResidualNorm())
        assert np.abs(res[True][0] - xs).max() <= 1e-7
        assert np.abs(res[True][0] - res[False][0]).max() <= 1e-8
        assert abs(res[True][1] - res[False][1]) <= 2, (res[True][1], res[False][1])
    # converged on entry: zero iterations
    x = Vec().createWithArray(xs)
    ksp.solve(b, x)
    assert ksp.getIterationNumber() <= 1
