"""CPU: the oracle assembly + direct solve reproduces the reference's own
known-answer tests (src/tests/test_solver.py), the box-mesh tables reproduce
the conventions src/tests/test_dmplex_wrapper.py pins, and the host-side logic
(options DB, RK tableau) behaves.  No GPU, no product compute path."""
import numpy as np
import pytest
import scipy.sparse.linalg as spl

from oracle import assembly, sem
from pynama_b200.domain.box_mesh import BoxMesh
from tests.util import fs_classes, ns_classes


def _oracle_solve(mesh, ngl, vort, vel_exact):
    tab = sem.SpectralTables(ngl, mesh.dim)
    cls = fs_classes(mesh)
    mats = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    vbc = np.where(cls == 1, vel_exact, 0.0)
    x = spl.spsolve(mats["K"].tocsc(), mats["Rw"] @ vort + mats["Krhs"] @ vbc)
    return x, mats


def test_kat1_uniform_flow():
    """test_solver.py:7-25: 3x3 elements, ngl=3, omega=0, v_bc=(4,0)  ->  ||v-(4,0)|| < 1e-12"""
    mesh = BoxMesh([3, 3], [0, 0], [1, 1], 3)
    exact = np.tile([4.0, 0.0], mesh.nnodes)
    x, mats = _oracle_solve(mesh, 3, np.zeros(mesh.nnodes), exact)
    assert np.linalg.norm(x - exact) < 1e-12
    K = mats["K"]
    assert abs(K - K.T).max() < 1e-11 * abs(K).max()           # SPD after elimination
    # the reference keeps explicit zeros: pattern = node graph (x) dense block on free rows
    assert K.nnz > (K != 0).sum()


def test_kat2_taylor_green():
    """test_solver.py:27-37: 10x10, ngl=5, t=0  ->  ||v - v_exact|| < 1e-5"""
    mesh = BoxMesh([10, 10], [0, 0], [1, 1], 5)
    x_, y_ = 2 * np.pi * mesh.coords[:, 0], 2 * np.pi * mesh.coords[:, 1]
    vel = np.stack([np.cos(x_) * np.sin(y_), -np.sin(x_) * np.cos(y_)], 1).ravel()
    vort = -4 * np.pi * np.cos(x_) * np.cos(y_)
    x, _ = _oracle_solve(mesh, 5, vort, vel)
    err = np.linalg.norm(x - vel)
    assert err < 1e-5
    assert abs(err - 5.01e-6) < 2e-7                            # value measured in the survey probe


def test_kat1_3d_uniform_flow():
    """the 3-D variant the reference left commented out (test_solver.py:39-42)"""
    mesh = BoxMesh([2, 2, 2], [0, 0, 0], [1, 1, 1], 3)
    exact = np.tile([1.0, -2.0, 0.5], mesh.nnodes)
    x, _ = _oracle_solve(mesh, 3, np.zeros(mesh.nnodes * 3), exact)
    assert np.linalg.norm(x - exact) < 1e-11


def test_ns_assembly_shapes_and_diagonals():
    """mat_ns.py:129-142 diagonal fix-ups"""
    mesh = BoxMesh([3, 3], [0, 0], [1, 1], 3)
    cls = ns_classes(mesh)
    tab = sem.SpectralTables(3, 2)
    m = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "NS")
    bc = np.nonzero(cls >= 2)[0]
    tang = np.nonzero(cls == 3)[0]
    normal = np.nonzero(cls == 2)[0]
    assert np.allclose(m["K"].diagonal()[bc], 1.0) and np.allclose(m["Krhs"].diagonal()[bc], 1.0)
    assert np.allclose(m["Krhsfs"].diagonal()[normal], 1.0)
    # Kfs tangential diagonal = K_e sums - 1
    assert m["Kfs"].shape == m["K"].shape and len(tang) > 0
    KK = (m["K"] + m["Kfs"]).tocsr()
    assert abs(KK - KK.T).max() < 1e-10 * abs(KK).max()


# ---- mesh conventions pinned by src/tests/test_dmplex_wrapper.py -------------------------------
def test_corner_order_of_cell0():
    m2 = BoxMesh([2, 2], [0, 0], [1, 1], 2)
    assert np.allclose(m2.getCellCornersCoords(0), [0, 0, 0.5, 0, 0.5, 0.5, 0, 0.5])      # :23-27
    m3 = BoxMesh([2, 2, 2], [0, 0, 0], [1, 1, 1], 2)
    h = 0.5
    want = [0, 0, 0, 0, h, 0, h, h, 0, h, 0, 0, 0, 0, h, h, 0, h, h, h, h, 0, h, h]       # :105-115
    assert np.allclose(m3.getCellCornersCoords(0), want)


def test_vertex_numbering_is_lexicographic_ngl2():
    m = BoxMesh([2, 2], [0, 0], [1, 1], 2)                                               # :223-227
    want = [[0, 0], [.5, 0], [1, 0], [0, .5], [.5, .5], [1, .5], [0, 1], [.5, 1], [1, 1]]
    assert np.allclose(m.coords, want)


@pytest.mark.parametrize("ngl", [2, 3, 5, 8])
def test_node_counts(ngl):
    m = BoxMesh([2, 2], [0, 0], [1, 1], ngl)                                             # :212-221
    assert m.nnodes == (2 * (ngl - 1) + 1) ** 2
    for s in m.sides:                                                                    # :61-88
        assert len(m.getBorderNodes(s)) == 2 * (ngl - 1) + 1
    m = BoxMesh([2, 2, 2], [0, 0, 0], [1, 1, 1], min(ngl, 5))
    g = 2 * (min(ngl, 5) - 1) + 1
    assert m.nnodes == g ** 3
    for s in m.sides:                                                                    # :161-195
        assert len(m.getBorderNodes(s)) == g * g


def test_node_coordinates_follow_local_order():
    """local node i sits at HCooOp[i] @ corners (dmplex.py:83-85)"""
    m = BoxMesh([2, 3, 2], [0, 0, 0], [1, 2, 1], 4, perturb=0.1)
    tab = sem.SpectralTables(4, 3)
    for e in (0, 5, m.nel - 1):
        xyz = sem.node_coords(tab, m.corner_xyz[e])
        assert np.abs(m.coords[m.elem_nodes[e]] - xyz).max() < 1e-13


# ---- host logic ------------------------------------------------------------------------------
def test_options_db():
    from pynama_b200.petsc import Options
    db = Options._parse(["-case", "taylor-green3d", "-ksp_type", "cg", "-ksp_rtol", "1e-8", "-flag",
                         "-ts_dt", "-0.5", "-pc_type", "bjacobi"])
    assert db == {"case": "taylor-green3d", "ksp_type": "cg", "ksp_rtol": "1e-8", "flag": "true",
                  "ts_dt": "-0.5", "pc_type": "bjacobi"}


def test_rk5bs_tableau_order_conditions():
    from pynama_b200.ts import TABLEAUX
    t = TABLEAUX["5bs"]
    s = len(t["b"])
    A = np.zeros((s, s))
    for i, row in enumerate(t["A"]):
        A[i, :len(row)] = row
    b, be, c = np.array(t["b"]), np.array(t["be"]), np.array(t["c"])
    assert np.allclose(A.sum(1), c, atol=1e-14)
    for bb, order in ((b, 5), (be, 4)):
        assert abs(bb.sum() - 1) < 1e-14
        for k in range(1, order):
            assert abs(bb @ c ** k - 1 / (k + 1)) < 1e-13, (order, k)
        assert abs(bb @ (A @ c) - 1 / 6) < 1e-13
        assert abs(bb @ (A @ c ** 2) - 1 / 12) < 1e-13
        assert abs(bb @ (c * (A @ c)) - 1 / 8) < 1e-13
        assert abs(bb @ (A @ (A @ c)) - 1 / 24) < 1e-13
    assert abs(b @ c ** 4 - 1 / 5) < 1e-13
    assert np.allclose(A[-1], b)            # FSAL


def test_taylor_green3d_field_is_consistent():
    """vorticity = curl(velocity) for the analytic field used as IC/BC (functions/taylor_green3d.py)"""
    from pynama_b200.functions import taylor_green3d as tg
    rng = np.random.default_rng(0)
    X = rng.uniform(0, 1, (20, 3))
    a = tg.alpha(0.02, 0.3)
    w = np.asarray(tg.vorticity(X, a)).reshape(-1, 3)
    eps = 1e-6
    J = np.zeros((20, 3, 3))
    for d in range(3):
        Xp, Xm = X.copy(), X.copy()
        Xp[:, d] += eps
        Xm[:, d] -= eps
        J[:, :, d] = (np.asarray(tg.velocity(Xp, a)).reshape(-1, 3) -
                      np.asarray(tg.velocity(Xm, a)).reshape(-1, 3)) / (2 * eps)
    curl = np.stack([J[:, 2, 1] - J[:, 1, 2], J[:, 0, 2] - J[:, 2, 0], J[:, 1, 0] - J[:, 0, 1]], 1)
    assert np.abs(curl - w).max() < 1e-6 * max(np.abs(w).max(), 1)


# ---- oracle time loop ---------------------------------------------------------------------------
def test_oracle_timestep_tracks_taylor_green_decay():
    """RK5BS steps of the restated evalRHS keep following the analytic Taylor-Green decay"""
    from oracle import timestep
    mesh = BoxMesh([6, 6], [0, 0], [1, 1], 5)
    p = timestep.OracleProblem(2, 5, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes(),
                               rho=0.5, mu=0.01, solver="direct")
    t, k0 = 0.0, None
    for _ in range(3):
        t, k0 = p.step(t, 5e-3, k0)
    a = timestep.tg_alpha(p.nu, t, 2)
    assert a < 0.99
    assert np.abs(p.vel - timestep.tg_velocity(mesh.coords, a)).max() < 2e-2      # 6x6, ngl 5 mesh
    assert np.abs(p.vort - timestep.tg_vorticity(mesh.coords, a)).max() < 0.2     # |w| ~ 12.6


def test_oracle_fischer_guess_same_solution_fewer_iterations():
    from oracle import timestep
    mesh = BoxMesh([3, 3, 3], [0, 0, 0], [1, 1, 1], 3)
    out = {}
    for f in (0, 6):
        p = timestep.OracleProblem(3, 3, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes(),
                                   rho=0.5, mu=0.01, solver="cg", rtol=1e-11, pc="bjacobi", fischer=f)
        t, k0 = 0.0, None
        for _ in range(3):
            t, k0 = p.step(t, 1e-3, k0)
        out[f] = (p.vel.copy(), p.iterations)
    assert np.abs(out[0][0] - out[6][0]).max() < 1e-8
    assert out[6][1] < 0.6 * out[0][1]


def test_ns_dof_classes_corner_rule():
    """boundary_conditions.py:227-243: up/down listed first -> the x dof of a left/right corner node
    is tangential, not normal"""
    from oracle import timestep
    mesh = BoxMesh([3, 3], [0, 0], [1, 1], 3)
    sides = ["up", "down", "left", "right"]
    cls = timestep.ns_dof_classes(sides, {s: mesh.getBorderNodes(s) for s in sides}, mesh.nnodes, 2)
    corner = 0                                  # node (0, 0): on 'down' and 'left'
    assert cls[corner * 2 + 1] == 2 and cls[corner * 2 + 0] == 3
    assert np.array_equal(ns_classes(mesh) == 0, cls == 0)
    sides = ["left", "right", "up", "down"]     # sides in the other order: normal wins everywhere
    cls2 = timestep.ns_dof_classes(sides, {s: mesh.getBorderNodes(s) for s in sides}, mesh.nnodes, 2)
    assert np.array_equal(cls2, ns_classes(mesh))
