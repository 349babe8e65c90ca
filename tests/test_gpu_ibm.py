"""GPU parity of the immersed-boundary path (csrc/ibm.cu behind cases/immersed_boundary.py):
the device-built H, S, A against the fixtures recorded from the reference
(tests/golden/ibm_rows_*.npz) and against the oracle (oracle/ibm.py), the reference's own IBM
tests (src/tests/test_ibm.py) through the case class, and the velocity correction."""
import ctypes
import os

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import ibm
from tests.test_oracle_ibm import GOLD, ROW_CASES, load_rows
from tests.util import csr_of

pytestmark = pytest.mark.gpu


def build(lag_xy, xy, h, dl, dynamic=False, kind=0, nodes=None):
    from pynama_b200._lib import call, ptr, vp, c_i64
    from pynama_b200.petsc import Context, Mat
    n = len(xy)
    nodes = np.arange(n, dtype=np.int32) if nodes is None else np.ascontiguousarray(nodes, np.int32)
    lag = np.ascontiguousarray(lag_xy, np.float64)
    cand = np.ascontiguousarray(xy[nodes], np.float64)
    H, S, A, mx = vp(), vp(), vp(), c_i64()
    mark = np.zeros(n, np.uint8)
    call("pyn_ibm_build", Context.get().handle, n, len(lag), ptr(lag), len(cand), ptr(nodes), ptr(cand),
         float(h), float(dl), float(h) ** 2, float(dl) * float(h), kind, int(dynamic), ctypes.byref(H),
         ctypes.byref(S), ctypes.byref(A), ctypes.byref(mx), ptr(mark))
    return Mat(H, "H"), Mat(S, "S"), Mat(A, "A"), mx.value, mark


@pytest.mark.parametrize("name", ROW_CASES)
def test_rows_match_reference_fixtures(name):
    """pattern bit-exact, delta weights equal to the reference's Python floats"""
    g, xy = load_rows(name)
    h, dl = float(g["h"]), float(g["dl"])
    H, S, A, mx, mark = build(g["lag_xy"], xy, h, dl, dynamic=bool(int(g["dynamic"])))
    ptr_, nodes, w = g["rowptr"], g["nodes"], g["weights"]
    csr = csr_of(H)
    assert csr.shape == (2 * len(g["lag_xy"]), 2 * len(xy))
    assert mx == np.diff(ptr_).max()
    for dof in range(2):
        rows = csr[dof::2]
        assert np.array_equal(rows.indptr, ptr_)
        assert np.array_equal(rows.indices, nodes * 2 + dof)            # explicit zeros kept
        want = w * h ** 2
        assert np.allclose(rows.data, want, rtol=1e-14, atol=0)
    exact = np.array_equal(csr[0::2].data, w * h ** 2)
    print(f"{name}: weights bit-identical to the reference: {exact}")
    assert set(np.nonzero(mark)[0]) == set(nodes[w > 0].tolist())


@pytest.mark.parametrize("name", ["coarse", "jitter", "dynamic"])
def test_matrices_match_oracle(name):
    g, xy = load_rows(name)
    h, dl = float(g["h"]), float(g["dl"])
    dyn = bool(int(g["dynamic"]))
    H, S, A, _, _ = build(g["lag_xy"], xy, h, dl, dynamic=dyn)
    rows = ibm.build_rows(g["lag_xy"], np.arange(len(xy)), xy, h, dl, dynamic=dyn)
    Ho, So, Ao = ibm.ibm_matrices(rows, len(xy), h, dl)
    Hg, Sg, Ag = csr_of(H), csr_of(S), csr_of(A)
    assert np.abs(Hg.toarray() - Ho).max() <= 1e-14 * np.abs(Ho).max()
    assert np.abs(Sg.toarray() - So).max() <= 1e-14 * np.abs(So).max()
    assert np.abs(Ag.toarray() - Ao).max() <= 1e-13 * np.abs(Ao).max()
    # S has the pattern of H^T; A the pattern of PETSc's MatMatMult: rows sharing a kept node
    assert (Sg != 0).nnz <= Sg.nnz and np.array_equal(sp.csr_matrix(Hg.T).indices, Sg.indices)
    D = sp.csr_matrix((np.ones(Hg.nnz), Hg.indices, Hg.indptr), shape=Hg.shape)
    want = sp.csr_matrix(D @ D.T)
    want.sort_indices()
    assert np.array_equal(Ag.indptr, want.indptr) and np.array_equal(Ag.indices, want.indices)
    # products through the SpMV of the KLE path
    from pynama_b200.petsc import Vec
    rng = np.random.default_rng(5)
    v = rng.uniform(-1, 1, 2 * len(xy))
    q = rng.uniform(-1, 1, 2 * len(rows))
    assert np.abs((H * Vec().createWithArray(v)).getArray() - Ho @ v).max() <= 1e-13 * np.abs(Ho @ v).max()
    assert np.abs((S * Vec().createWithArray(q)).getArray() - So @ q).max() <= 1e-13 * np.abs(So @ q).max()
    assert np.abs((A * Vec().createWithArray(q)).getArray() - Ao @ q).max() <= 1e-12 * np.abs(Ao @ q).max()


def test_other_delta_functions_and_subset_of_nodes():
    g, xy = load_rows("coarse")
    h, dl = float(g["h"]), float(g["dl"])
    sub = np.nonzero((np.abs(xy[:, 0]) < 3.2) & (np.abs(xy[:, 1]) < 3.2))[0]
    for kind, name in ((1, "three"), (2, "linear")):
        H, _, _, _, _ = build(g["lag_xy"], xy, h, dl, kind=kind, nodes=sub)
        rows = ibm.build_rows(g["lag_xy"], sub, xy[sub], h, dl, kind=name)
        Ho = ibm.dense_H(rows, len(xy)) * h ** 2
        assert np.abs(csr_of(H).toarray() - Ho).max() <= 1e-14 * np.abs(Ho).max()


def test_marker_without_euler_node_raises():
    g, xy = load_rows("coarse")
    lag = np.array([[40.0, 40.0]])
    with pytest.raises(RuntimeError, match="Lag Node without Euler"):
        build(lag, xy, float(g["h"]), float(g["dl"]))


def _case(cls_name="ImmersedBoundaryStatic", case="ibm-static", **opts):
    from pynama_b200.cases import immersed_boundary as ib
    from pynama_b200.run_case import loadCase
    fem = getattr(ib, cls_name)(loadCase(case), case=case, **opts)
    fem.setUp()
    fem.setUpSolver()
    return fem


COARSE = {"nelem": [10, 10], "ngl": 3, "lower": [-5, -5], "upper": [5, 5]}


def test_reference_ibm_tests_through_the_case_class():
    """src/tests/test_ibm.py: TestSearch :29-66 and TestDiracRegularGrid :77-98"""
    fem = _case(**COARSE)
    assert len(fem.getAffectedCells(1)) == 2 * 2
    assert len(fem.getAffectedCells(xSide=2, ySide=2)) == 4 * 4
    assert len(fem.getAffectedCells(xSide=1, ySide=2)) == 2 * 4
    off = np.array([0.5, 0.5])
    assert len(fem.getAffectedCells(1, center=off)) == 1
    assert len(fem.getAffectedCells(xSide=2, ySide=2, center=off)) == 3 * 3
    assert len(fem.getAffectedCells(xSide=1, ySide=2, center=off)) == 1 * 3
    for args, want in (((1,), 5 * 5), ((2, 2), 9 * 9), ((1, 2), 5 * 9)):
        nodes, _ = fem.dom.getNodesCoordsFromEntities(fem.getAffectedCells(*args))
        assert len(nodes) == want
    D = fem.H
    nrows = D.getSizes()[0][1]
    dim = fem.dom.getDimension()
    for i in range(nrows):
        assert abs(D.getRow(i)[1].sum() - 1) < 1e-10                      # mass
    for lagInd in range(0, nrows, dim):
        eulerInd, diracs = D.getRow(lagInd)
        dist = fem.dom.getNodesCoordinates(eulerInd // dim) - fem.body.getNodeCoordinates(lagInd // dim)
        assert abs((dist[:, 0] * dist[:, 1] * diracs).sum()) < 1e-10      # momentum
    # the marker -> nodes mapping the reference keeps in collectedNodes
    g, xy = load_rows("coarse")
    assert np.allclose(fem.body.getAllCoordinates(), g["lag_xy"], rtol=0, atol=0)
    for l in (0, 3, len(g["lag_xy"]) - 1):
        assert np.array_equal(fem.collectedNodes[l]["nodes"], g["nodes"][g["rowptr"][l]:g["rowptr"][l + 1]])
    assert fem.maxNodesPerLag == np.diff(g["rowptr"]).max()


def test_velocity_correction_matches_oracle():
    fem = _case(nelem=[20, 20], ngl=3, lower=[-3, -3], upper=[3, 3])
    xy = fem.dom.getFullCoordArray()
    h, dl = fem.nodeSeparation, fem.body.getElementLong()
    rows = ibm.build_rows(fem.body.getAllCoordinates(), fem.ibmNodes, fem.eulerCoords, h, dl)
    Ho, So, Ao = ibm.ibm_matrices(rows, len(xy), h, dl)
    rng = np.random.default_rng(11)
    v0 = rng.uniform(-1, 1, 2 * len(xy))
    fem.vel.setValues(np.arange(v0.size), v0)
    body = np.tile([0.3, -0.1], fem.body.getTotalNodes())
    fem.body.bodies[0].setVelocity(np.arange(body.size), body)
    fem.ksp.setTolerances(rtol=1e-13)
    fem.computeVelocityCorrection()
    new, corr, q = ibm.velocity_correction(Ho, So, Ao, v0, body)
    got = fem.vel.getArray()
    assert np.abs(Ho @ got - body).max() < 1e-9                           # no slip on the markers
    assert np.abs(got - new).max() <= 1e-8 * np.abs(new).max()
    assert np.abs(fem.virtualFlux.getArray() - q).max() <= 1e-7 * np.abs(q).max()
    assert set(np.nonzero(fem.vel_correction.getArray().reshape(-1, 2).any(axis=1))[0]) <= set(fem.mNodes)


def test_dynamic_body_rebuilds_the_matrices():
    fem = _case("ImmersedBoundaryDynamic", "ibm-dynamic", nelem=[30, 30], ngl=3, lower=[-3, -3], upper=[3, 3])
    h, dl = fem.nodeSeparation, fem.body.getElementLong()
    xy = fem.dom.getFullCoordArray()
    for t in (0.0, 0.7):
        fem.vel.set(0.0)
        fem.computeVelocityCorrection(t=t)
        centre, vel = ibm.moving_body_state(t, fem.U_ref, [0, 0])
        lag = ibm.body_circle(h, 0.5)[0] + np.asarray(centre, dtype=float)
        assert np.array_equal(fem.body.getAllCoordinates(), lag)
        rows = ibm.build_rows(lag, fem.ibmNodes, fem.eulerCoords, h, dl, dynamic=True)
        Ho, So, Ao = ibm.ibm_matrices(rows, len(xy), h, dl)
        assert np.abs(csr_of(fem.H).toarray() - Ho).max() <= 1e-14 * np.abs(Ho).max()
        want = np.tile(vel, len(lag))
        assert np.abs(Ho @ fem.vel.getArray() - want).max() < 1e-8 * max(np.abs(want).max(), 1e-30)


def test_static_case_advances_in_time():
    """a few steps of the ibm-static loop on a small mesh: fields stay finite, the markers keep
    the body velocity after each correction"""
    from pynama_b200.petsc import Options
    o = Options()
    o.setValue("ts_adapt_type", "none")
    o.setValue("ts_dt", 5e-3)
    try:
        fem = _case(nelem=[12, 12], ngl=3, lower=[-3, -3], upper=[3, 3])
        fem.config["save-n-steps"] = 10 ** 9
        fem.viewer.saveData = lambda *a, **k: None
        fem.viewer.writeXmf = lambda *a, **k: None
        fem.computeInitialCondition()
        fem.ts.setSolution(fem.vort)
        for i in range(2):
            fem.ts.step()
            fem._afterStep(i)
        hv = (fem.H * fem.vel).getArray()
        assert np.isfinite(fem.vort.getArray()).all() and np.abs(hv).max() < 1e-8
    finally:
        o.delValue("ts_adapt_type")
        o.delValue("ts_dt")
