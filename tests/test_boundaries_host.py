"""Host mirror of the boundary-condition layer and the analytic field libraries,
on CPU: pynama_b200/functions/* against fixtures recorded from the reference
(tests/golden/functions.npz, make_golden_functions.py), and the reference's own tests
src/tests/test_boundary_conditions.py, test_boundary_ns.py and the Vec-free parts of
test_domain.py restated on pynama_b200/boundaries and pynama_b200/domain."""
import importlib
import os

import numpy as np
import pytest

from pynama_b200.boundaries.boundary_conditions import Boundary, BoundaryConditions, FunctionBoundary

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---- analytic fields --------------------------------------------------------------------------
@pytest.mark.parametrize("lib_name", ["taylor_green", "taylor_green3d", "taylor_green_3d"])
def test_function_libraries_match_reference(lib_name):
    g = np.load(os.path.join(GOLD, "functions.npz"))
    lib = importlib.import_module(f"pynama_b200.functions.{lib_name}")
    coords = g[f"{lib_name}__coords"]
    alphas = np.array([lib.alpha(nu, t) for nu, t in g["nu_t"]])
    assert np.array_equal(alphas, g[f"{lib_name}__alpha"])
    for name in g[f"{lib_name}__names"]:
        for k, a in enumerate(alphas):
            mine = np.asarray(getattr(lib, str(name))(coords.copy(), a), dtype=np.float64).ravel()
            ref = g[f"{lib_name}__{name}"][k]
            assert mine.shape == ref.shape, name
            assert np.abs(mine - ref).max() <= 1e-14 * max(np.abs(ref).max(), 1e-300), (name, k)
    if f"{lib_name}__velocity_test" in g:
        for k, (nu, t) in enumerate(g["nu_t"]):
            mine = np.array([lib.velocity_test(c, t, nu) for c in coords])
            assert np.abs(mine - g[f"{lib_name}__velocity_test"][k]).max() <= 1e-14


# ---- test_boundary_conditions.py:9-64 -----------------------------------------------------------
@pytest.mark.parametrize("dim,vel,vort", [(2, [1, 4], [0]), (3, [1, 6, 8], [32, 12, 124])])
def test_basic_boundary(dim, vel, vort):
    b = Boundary("left", "free-slip", dim)
    b.setValues("velocity", vel)
    b.setValues("vorticity", vort)
    assert b.getName() == "left" and b.getType() == "free-slip"
    nodes = [0, 11, 24, 78]
    b.setNodes(nodes)
    assert np.array_equal(b.getDofsConstrained(), [i * dim + d for i in nodes for d in range(dim)])
    assert np.array_equal(b.getValues("velocity"), np.tile(vel, len(nodes)))
    assert np.array_equal(b.getValues("vorticity"), np.tile(vort, len(nodes)))
    b.setNodes([123, 12415, 1566, 121])                       # kept in the order given
    assert np.array_equal(b.getNodes(), [123, 12415, 1566, 121])
    with pytest.raises(Exception):
        b.getValues("pressure")


# ---- test_boundary_conditions.py:66-209 ---------------------------------------------------------
SIDES2 = ["up", "down", "right", "left"]
VALS_FS = {"velocity": [1, 0], "vorticity": [0]}
CUST = {"name": "taylor_green", "attributes": ["velocity", "vorticity"]}


def _bcs(data, sides=SIDES2):
    bcs = BoundaryConditions(sides)
    bcs.setBoundaryConditions(data)
    return bcs


def test_bc_types():
    bcs = _bcs({"free-slip": {s: VALS_FS for s in SIDES2}})
    assert bcs.getType() == "FS"
    assert len(bcs.getNamesByType("free-slip")) == 4 and bcs.getNamesByType("no-slip") == []

    bcs = _bcs({"custom-func": CUST})
    assert bcs.getType() == "FS" and len(bcs.getNamesByType("free-slip")) == 4
    assert set(SIDES2) <= set(bcs.getBordersNeedsCoords())

    bcs = _bcs({"free-slip": {"down": VALS_FS, "right": {"custom-func": CUST}, "left": {"custom-func": CUST},
                              "up": VALS_FS}})
    assert bcs.getType() == "FS" and len(bcs.getNamesByType("free-slip")) == 4
    assert {"right", "left"} <= set(bcs.getBordersNeedsCoords())

    bcs = _bcs({"no-slip": {s: {"velocity": [1, 0]} for s in SIDES2}})
    assert bcs.getType() == "NS"
    assert len(bcs.getNamesByType("no-slip")) == 4 and bcs.getNamesByType("free-slip") == []

    bcs = _bcs({"free-slip": {"down": VALS_FS, "right": {"custom-func": CUST}},
                "no-slip": {"left": {"velocity": [1, 0]}, "up": {"velocity": [1, 0]}}})
    assert bcs.getType() == "FS-NS"
    assert {"down", "right"} <= set(bcs.getNamesByType("free-slip"))
    assert {"up", "left"} <= set(bcs.getNamesByType("no-slip"))
    assert "right" in bcs.getBordersNeedsCoords()


def test_bc_indices_by_type():
    down, right, up, left = [0, 1, 2, 3], [3, 4, 5, 6], [6, 7, 8], [8, 9, 10, 11, 0]
    bcs = _bcs({"free-slip": {"down": VALS_FS, "right": VALS_FS},
                "no-slip": {"left": {"velocity": [2, 0]}, "up": {"velocity": [2, 0]}}})
    for name, nodes in (("down", down), ("up", up), ("left", left), ("right", right)):
        bcs.setBoundaryNodes(name, nodes)
    ns = {n * 2 + d for n in left + up for d in range(2)}
    fs = {n * 2 + d for n in down + right for d in range(2)}
    assert set(bcs.getNoSlipIndices()) == ns and set(bcs.getFreeSlipIndices()) == fs
    assert set(bcs.getIndicesByType("no-slip")) == ns and set(bcs.getIndicesByType("free-slip")) == fs


def test_uniform_from_reynolds_number():
    """boundary_conditions.py:86-112: velRef = re (mu / rho) / Lref along `direction`"""
    bcs = _bcs({"uniform": {"re": 10, "direction": 0, "Lref": "1", "rho": 0.5, "mu": 0.01}})
    assert bcs.getType() == "FS"
    v = np.asarray(bcs.getFreeStreamVelocity(), dtype=float)
    assert np.allclose(v, [10 * 0.02, 0.0], rtol=0, atol=1e-16)
    with pytest.raises(Exception, match="not defined"):
        _bcs({"uniform": {"re": 10, "direction": 0}})


# ---- test_boundary_conditions.py:211-254: function boundaries -----------------------------------
def test_function_boundary_taylor_green_2d():
    from pynama_b200.functions.taylor_green import velocity_test, vorticity_test
    coords = np.array([0, 0, 0.1, 0.1, 0.3, 0.3, 0.6, 0.6, 0.8, 0.8], dtype=float)
    b = FunctionBoundary("up", "taylor_green", ["velocity", "vorticity", "alpha"], dim=2)
    b.setNodes([0, 1, 2, 3, 4])
    b.setNodesCoordinates(coords)
    assert np.array_equal(np.ravel(b.getNodesCoordinates()), coords)
    for t, nu in ((0, 100), (0.3, 0.02)):
        xy = coords.reshape(5, 2)
        vel_ref = np.concatenate([velocity_test(c, t, nu) for c in xy])
        assert np.abs(b.getValues("velocity", t, nu) - vel_ref).max() < 1e-14
        vort_ref = np.array([vorticity_test(c, nu, t) for c in xy]).ravel()
        assert np.abs(b.getValues("vorticity", t, nu) - vort_ref).max() < 1e-13


# ---- test_boundary_ns.py ------------------------------------------------------------------------
@pytest.mark.parametrize("dim,nodes", [(2, [0, 11, 24, 78]), (3, [0, 11, 24])])
def test_no_slip_normal_and_tangential_dofs(dim, nodes):
    sides = {"down": 1, "left": 0} if dim == 2 else {"down": 1, "left": 0, "front": 2}
    for name, normal in sides.items():
        b = Boundary(name, "no-slip", dim)
        b.setValues("velocity", list(range(1, dim + 1)))
        b.setNodes(nodes)
        assert b.getNormalDofs() == {n * dim + normal for n in nodes}
        tang = [d for d in range(dim) if d != normal]
        assert b.getTangDofs() == {n * dim + d for d in tang for n in nodes}
        assert b.getTangDirections() == tang


# ---- test_domain.py (the parts that need no Vec) -------------------------------------------------
BOX = {"ngl": 3, "box-mesh": {"nelem": [2, 2], "lower": [0, 0], "upper": [1, 1]}}
UNIFORM = {"uniform": {"velocity": [1, 5], "vorticity": [8]}}


def _dom(bc=None, box=BOX, **opts):
    from pynama_b200.domain.domain import Domain
    data = {"domain": box}
    if bc is not None:
        data["boundary-conditions"] = bc
    dom = Domain()
    dom.configure(data)
    dom.setOptions(**opts)
    if bc is None:
        dom.create()
    else:
        dom.setUp()
    return dom


def test_domain_box_mesh_counts_and_options():
    dom = _dom()
    assert (dom.getMeshType(), dom.getNumOfElements(), dom.getNGL(), dom.getNumOfNodes()) == ("box", 4, 3, 25)
    dom = _dom(ngl=7)                                            # test_domain.py:53-63
    assert dom.getNGL() == 7 and dom.getNumOfNodes() == (7 * 2 - 1) * (7 * 2 - 1)
    assert _dom(nelem=[4, 4]).getNumOfElements() == 16


def test_domain_boundary_indices():
    dom = _dom(UNIFORM)                                          # test_domain.py:131-139
    assert len(dom.getGlobalIndicesDirichlet()) == 32 and len(dom.getGlobalIndicesNoSlip()) == 0
    ns = {"no-slip": {s: {"velocity": [1, 5], "vorticity": [8]} for s in ["left", "right", "up", "down"]}}
    dom = _dom(ns)                                               # :141-153
    assert len(dom.getGlobalIndicesDirichlet()) == 0 and len(dom.getGlobalIndicesNoSlip()) == 32
    dom = _dom({"custom-func": {"name": "taylor_green", "attributes": ["velocity", "vorticity", "alpha"]}})
    assert len(dom.getGlobalIndicesDirichlet()) == 32            # :104-116


def test_domain_boundary_values_without_a_vec():
    """test_domain.py:155-223 on the merged (indices, values) pairs the device scatter consumes"""
    from pynama_b200.functions.taylor_green import velocity_test
    dom = _dom(UNIFORM)
    bc = dom.getBoundaryConditions()
    n = dom.getNumOfNodes()
    for name, width, ref in (("velocity", 2, [1, 5]), ("vorticity", 1, [8])):
        idx, vals = bc.getValuesMerged(name, 0.0, 0.02)
        got = np.zeros(n * width)
        got[np.asarray(idx)] = np.asarray(vals)
        want = np.zeros(n * width)
        border = sorted({int(i) // 2 for i in dom.getGlobalIndicesDirichlet()})
        for node in border:
            want[node * width:(node + 1) * width] = ref
        assert np.array_equal(got, want), name
    dom = _dom({"custom-func": {"name": "taylor_green", "attributes": ["velocity", "vorticity", "alpha"]}})
    bc = dom.getBoundaryConditions()
    t, nu = 0.0, 0.5 / 0.01
    idx, vals = bc.getValuesMerged("velocity", t, nu)
    got = np.zeros(n * 2)
    got[np.asarray(idx)] = np.asarray(vals)
    want = np.zeros(n * 2)
    for node in sorted({int(i) // 2 for i in dom.getGlobalIndicesDirichlet()}):
        want[node * 2:node * 2 + 2] = velocity_test(dom.getNodesCoordinates([node])[0], t, nu)
    assert np.abs(got - want).max() < 1e-15


def test_domain_node_separation_ibm():
    box = {"ngl": 3, "box-mesh": {"nelem": [5, 5], "lower": [-5, -5], "upper": [8, 8]}}
    dom = _dom(UNIFORM, box=box)                                 # test_domain.py:245-255
    assert dom.getNodeSeparationIBM() == pytest.approx((8 - -5) / 5 / 2, abs=1e-15)


# ---- test_dmplex_wrapper.py on the box generator: sides on asymmetric meshes ---------------------
def _box(nelem, upper, ngl):
    from pynama_b200.domain.box_mesh import BoxMesh
    return BoxMesh(nelem, [0.0] * len(nelem), upper, ngl)


def test_box_2d_borders_and_corners():
    m = _box([3, 4], [0.6, 0.8], 3)                               # test_dmplex_wrapper.py:10-59
    assert m.nel == 12
    assert np.allclose(m.getCellCornersCoords(0).ravel(), [0, 0, 0.2, 0, 0.2, 0.2, 0, 0.2], atol=1e-13)
    assert len(set(np.asarray(m.getBordersNodes()).tolist())) == 28
    for b in m.getBordersNames():
        assert len(m.getBorderNodes(b)) == (7 if b in ("up", "down") else 9), b
    # which side is which: 'down' is y = lower, 'left' is x = lower (boundary.py:151-162)
    assert np.allclose(m.coords[m.getBorderNodes("down")][:, 1], 0.0)
    assert np.allclose(m.coords[m.getBorderNodes("up")][:, 1], 0.8)
    assert np.allclose(m.coords[m.getBorderNodes("left")][:, 0], 0.0)
    assert np.allclose(m.coords[m.getBorderNodes("right")][:, 0], 0.6)


@pytest.mark.parametrize("ngl", [2, 4, 6, 8])
def test_box_2d_border_counts_per_ngl(ngl):
    m = _box([2, 3], [0.6, 0.8], ngl)                             # :61-88
    assert len(set(np.asarray(m.getBordersNodes()).tolist())) == 10 + 10 * (ngl - 2)
    for b in m.getBordersNames():
        want = 3 + 2 * (ngl - 2) if b in ("up", "down") else 4 + 3 * (ngl - 2)
        assert len(m.getBorderNodes(b)) == want, b


def test_box_3d_borders_and_corners():
    m = _box([3, 4, 5], [0.6, 0.8, 1.0], 3)                       # :90-147
    assert m.getDimension() == 3 and m.nel == 60
    want = [0, 0, 0, 0, 0.2, 0, 0.2, 0.2, 0, 0.2, 0, 0, 0, 0, 0.2, 0.2, 0, 0.2, 0.2, 0.2, 0.2, 0, 0.2, 0.2]
    assert np.allclose(m.getCellCornersCoords(0).ravel(), want, atol=1e-13)
    names = m.getBordersNames()
    assert len(names) == 6 and set(names) == {"up", "down", "left", "right", "front", "back"}
    assert len(set(np.asarray(m.getBordersNodes()).tolist())) == 28 * 11 + 35 * 2
    for b in names:
        want = 7 * 11 if b in ("up", "down") else (9 * 11 if b in ("left", "right") else 7 * 9)
        assert len(m.getBorderNodes(b)) == want, b
    for b, axis, val in (("down", 1, 0.0), ("up", 1, 0.8), ("left", 0, 0.0), ("right", 0, 0.6)):
        assert np.allclose(m.coords[m.getBorderNodes(b)][:, axis], val), b
    z = {b: np.unique(np.round(m.coords[m.getBorderNodes(b)][:, 2], 12)) for b in ("front", "back")}
    assert all(len(v) == 1 for v in z.values()) and {float(z["front"][0]), float(z["back"][0])} == {0.0, 1.0}


@pytest.mark.parametrize("ngl", [2, 4, 6])
def test_box_3d_border_counts_per_ngl(ngl):
    m = _box([2, 3, 4], [0.6, 0.8, 1.0], ngl)                     # :161-195
    q = ngl - 2
    assert len(set(np.asarray(m.getBordersNodes()).tolist())) == 54 + (36 + 68) * q + 52 * q * q
    for b in m.getBordersNames():
        if b in ("up", "down"):
            want = 15 + 22 * q + 8 * q * q
        elif b in ("left", "right"):
            want = 20 + 31 * q + 12 * q * q
        else:
            want = 12 + 17 * q + 6 * q * q
        assert len(m.getBorderNodes(b)) == want, b


def test_edge_width_is_the_node_separation_base():
    for nelem, upper in (([4, 4], [0.6, 0.8]), ([4, 4, 4], [0.8, 0.8, 0.8])):    # :49-59, :149-159
        box = {"ngl": 3, "box-mesh": {"nelem": nelem, "lower": [0] * len(nelem), "upper": upper}}
        dom = _dom(box=box)
        assert dom.getNodeSeparationIBM() * 2 == pytest.approx(upper[0] / nelem[0], abs=1e-16)
