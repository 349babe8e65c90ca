"""The immersed-boundary oracle (oracle/ibm.py) against the fixtures recorded from the reference
(tests/golden/make_golden_ibm.py), plus the reference's own IBM tests restated on the oracle
(src/tests/test_ibm.py: affected cells, mass and momentum conservation of the delta weights)."""
import os

import numpy as np
import pytest

from oracle import ibm

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROW_CASES = ["coarse", "static", "jitter", "dynamic"]


def load_rows(name):
    g = np.load(os.path.join(GOLD, f"ibm_rows_{name}.npz"))
    from tests.golden.make_golden_ibm import grid_nodes
    xy = g["euler_xy"] if float(g["jitter"]) else grid_nodes(g["nelem"].tolist(), g["lower"], g["upper"], int(g["ngl"]))
    return g, xy


def test_delta_functions_match_reference():
    g = np.load(os.path.join(GOLD, "ibm_dirac.npz"))
    for name in ("fourGrid", "threeGrid", "linear"):
        mine = np.array([getattr(ibm, name)(float(r)) for r in g["r"]])
        assert np.array_equal(mine, g[name]), name


def test_bodies_match_reference():
    g = np.load(os.path.join(GOLD, "ibm_bodies.npz"))
    for i in range(3):
        dh = float(g[f"circle{i}_dh"])
        for kind, make in (("circle", lambda: ibm.body_circle(dh, 0.5)), ("line", lambda: ibm.body_line(dh)),
                           ("box", lambda: ibm.body_box(dh))):
            xy, dl = make()
            assert np.array_equal(xy, g[f"{kind}{i}_xy"]), (kind, i)
            assert dl == float(g[f"{kind}{i}_dl"])
    for t, c, v in zip(g["mover_t"], g["mover_center"], g["mover_vel"]):
        cc, vv = ibm.moving_body_state(float(t), float(g["mover_uref"]), g["mover_start"])
        assert np.array_equal(np.array(cc, dtype=float), c) and np.array_equal(np.array(vv, dtype=float), v)


@pytest.mark.parametrize("name", ROW_CASES)
def test_rows_match_reference(name):
    g, xy = load_rows(name)
    rows = ibm.build_rows(g["lag_xy"], np.arange(len(xy)), xy, float(g["h"]), float(g["dl"]),
                          dynamic=bool(int(g["dynamic"])))
    ptr = g["rowptr"]
    assert len(rows) == len(ptr) - 1
    for l, (nodes, w) in enumerate(rows):
        assert np.array_equal(nodes, g["nodes"][ptr[l]:ptr[l + 1]]), l
        assert np.array_equal(w, g["weights"][ptr[l]:ptr[l + 1]]), l          # bit for bit


def test_affected_cells_reference_counts():
    """src/tests/test_ibm.py:29-66 on the same 10x10 mesh of [-5,5]^2"""
    c = (np.stack(np.meshgrid(np.arange(10), np.arange(10), indexing="xy"), -1).reshape(-1, 2) + 0.5) - 5.0
    assert len(ibm.affected_cells(c, 1)) == 2 * 2
    assert len(ibm.affected_cells(c, 2, 2)) == 4 * 4
    assert len(ibm.affected_cells(c, 1, 2)) == 2 * 4
    off = np.array([0.5, 0.5])
    assert len(ibm.affected_cells(c, 1, center=off)) == 1
    assert len(ibm.affected_cells(c, 2, 2, center=off)) == 3 * 3
    assert len(ibm.affected_cells(c, 1, 2, center=off)) == 1 * 3


@pytest.mark.parametrize("name", ["coarse", "static"])
def test_mass_and_momentum_conservation(name):
    """src/tests/test_ibm.py:77-98: rows of H = h^2 D sum to one; first cross moment vanishes"""
    g, xy = load_rows(name)
    h = float(g["h"])
    rows = ibm.build_rows(g["lag_xy"], np.arange(len(xy)), xy, h, float(g["dl"]))
    for (nodes, w), lp in zip(rows, g["lag_xy"]):
        assert abs((w * h * h).sum() - 1) < 1e-10
        d = xy[nodes] - lp
        assert abs((d[:, 0] * d[:, 1] * w * h * h).sum()) < 1e-10


def test_velocity_correction_enforces_body_velocity():
    g, xy = load_rows("static")
    h, dl = float(g["h"]), float(g["dl"])
    rows = ibm.build_rows(g["lag_xy"], np.arange(len(xy)), xy, h, dl)
    H, S, A = ibm.ibm_matrices(rows, len(xy), h, dl)
    assert np.allclose(A, A.T, rtol=0, atol=1e-15) and np.linalg.eigvalsh(A).min() > 0
    rng = np.random.default_rng(3)
    vel = rng.uniform(-1, 1, 2 * len(xy))
    body = np.tile([0.3, -0.1], len(rows))
    new, corr, q = ibm.velocity_correction(H, S, A, vel, body)
    assert np.abs(H @ new - body).max() < 1e-10
    assert set(np.nonzero(corr.reshape(-1, 2).any(axis=1))[0]) <= set(ibm.affected_nodes(rows))
