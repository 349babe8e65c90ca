"""GPU parity of the matrix-free KLE product: K x evaluated element by element
with sum-factorised quadrature against (i) the oracle's assembled K and (ii) the library's own
stored-block product, on affine and perturbed meshes, free-slip and no-slip dof classes; then the
Krylov solve and the time loop through it."""
import numpy as np
import pytest

from oracle import assembly, sem
from tests.util import fs_classes, ns_classes

pytestmark = pytest.mark.gpu


def _setup(nelem, ngl, perturb, ns=False):
    from pynama_b200.device_mesh import DeviceMesh
    from pynama_b200.domain.box_mesh import BoxMesh
    mesh = BoxMesh(nelem, [0.0] * 3, [1.0, 1.5, 0.75], ngl, perturb=perturb)
    dm = DeviceMesh(3, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
    cls = ns_classes(mesh) if ns else fs_classes(mesh)
    dm.set_dof_class(cls)
    return mesh, dm, cls


@pytest.mark.parametrize("ngl", [2, 3, 4, 5, 6])
@pytest.mark.parametrize("perturb", [0.0, 0.15])
def test_matfree_product_matches_oracle_matrix(ngl, perturb):
    from pynama_b200.petsc import Vec
    nelem = [3, 2, 2] if ngl >= 5 else [3, 3, 2]
    mesh, dm, cls = _setup(nelem, ngl, perturb)
    tab = sem.SpectralTables(ngl, 3)
    K_ref = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")["K"].tocsr()
    mats = dm.assemble_kle("FS")
    K = mats["K"]
    rng = np.random.default_rng(7)
    xv = rng.uniform(-1, 1, mesh.nnodes * 3)
    x = Vec().createWithArray(xv)
    y_blocks = (K * x).getArray()
    K.setMatFree(True)
    y_mf = (K * x).getArray()
    want = K_ref @ xv
    scale = np.abs(want).max()
    assert np.abs(y_blocks - want).max() <= 1e-12 * scale
    assert np.abs(y_mf - want).max() <= 1e-12 * scale, np.abs(y_mf - want).max() / scale
    # y = a K x + b z through the same path
    z = Vec().createWithArray(rng.uniform(-1, 1, mesh.nnodes * 3))
    w = Vec(mesh.nnodes * 3)
    K.multAdd(x, z, w)
    assert np.abs(w.getArray() - (want + z.getArray())).max() <= 1e-12 * scale
    K.setMatFree(False)
    assert np.array_equal((K * x).getArray(), y_blocks)


def test_matfree_no_slip_classes():
    """K of MatNS: free x free, identity on normal and tangential wall dofs (mat_ns.py:116-130)"""
    from pynama_b200.petsc import Vec
    mesh, dm, cls = _setup([3, 3, 3], 3, 0.1, ns=True)
    K = dm.assemble_kle("NS")["K"]
    xv = np.random.default_rng(3).uniform(-1, 1, mesh.nnodes * 3)
    x = Vec().createWithArray(xv)
    y0 = (K * x).getArray()
    K.setMatFree(True)
    y1 = (K * x).getArray()
    assert np.abs(y1 - y0).max() <= 1e-12 * np.abs(y0).max()
    assert np.array_equal(y1[cls != 0], xv[cls != 0])


def test_matfree_rejects_other_matrices():
    mesh, dm, cls = _setup([2, 2, 2], 3, 0.0)
    mats = dm.assemble_kle("FS")
    with pytest.raises(RuntimeError):
        mats["Rw"].setMatFree(True)


@pytest.mark.parametrize("ksp_type", ["cg", "gmres"])
def test_solve_through_matfree_operator(ksp_type):
    from pynama_b200.petsc import KSP, PC, Vec
    mesh, dm, cls = _setup([4, 4, 4], 4, 0.1)
    mats = dm.assemble_kle("FS")
    K = mats["K"]
    rng = np.random.default_rng(11)
    xs = np.where(cls == 0, rng.uniform(-1, 1, mesh.nnodes * 3), 0.5)
    b = K * Vec().createWithArray(xs)
    out = {}
    for mf in (False, True):
        K.setMatFree(mf)
        ksp = KSP().create()
        ksp.setType(ksp_type)
        pc = PC().create()
        pc.setType("bjacobi")
        ksp.setPC(pc)
        ksp.setOperators(K)
        ksp.setTolerances(rtol=1e-12, max_it=5000)
        ksp.setUp()
        x = Vec(mesh.nnodes * 3)
        ksp.solve(b, x)
        assert ksp.reason > 0
        out[mf] = (x.getArray(), ksp.its)
    assert np.abs(out[True][0] - xs).max() <= 1e-8
    assert np.abs(out[True][0] - out[False][0]).max() <= 1e-9
    assert abs(out[True][1] - out[False][1]) <= 2


def test_time_loop_with_matfree_option():
    """-kle_matfree 1 through the BaseProblem / KleSolver / TsSolver stack: same fields after 2 steps"""
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.petsc import Options
    from pynama_b200.run_case import loadCase
    res = {}
    for mf in (0, 1):
        o = Options()
        for k, v in (("ksp_type", "cg"), ("pc_type", "bjacobi"), ("ksp_rtol", 1e-12), ("ksp_max_it", 50000),
                     ("ts_adapt_type", "none"), ("ts_dt", 2e-3), ("kle_matfree", mf)):
            o.setValue(k, v)
        o.delValue("ksp_fischer_guess")
        fem = BaseProblem(loadCase("taylor-green3d"), case="taylor-green3d", nelem=[3, 3, 3], ngl=4)
        fem.setUp()
        fem.setUpSolver()
        fem.ts._post = None
        fem.ts.setSolution(fem.vort)
        for _ in range(2):
            fem.ts.step()
        res[mf] = (fem.solverKLE.getSolution().getArray().copy(), fem.vort.getArray().copy())
        assert fem.mat.K._matfree == bool(mf)
    Options().delValue("kle_matfree")
    assert np.abs(res[1][0] - res[0][0]).max() <= 1e-9 * np.abs(res[0][0]).max()
    assert np.abs(res[1][1] - res[0][1]).max() <= 1e-8 * np.abs(res[0][1]).max()
