"""Host side of the immersed-boundary path without a GPU: the bodies' markers, the scalar delta
functions and the prescribed motion of pynama_b200/domain/immersed_body.py against the fixtures
recorded from the reference (tests/golden/ibm_bodies.npz, ibm_dirac.npz), and the cell search of
the case class (src/tests/test_ibm.py:29-66)."""
import os

import numpy as np

from pynama_b200.domain import immersed_body as ib

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_delta_functions_bit_identical():
    g = np.load(os.path.join(GOLD, "ibm_dirac.npz"))
    for name in ("fourGrid", "threeGrid", "linear"):
        mine = np.array([getattr(ib, name)(float(r)) for r in g["r"]])
        assert np.array_equal(mine, g[name]), name


def test_body_markers_bit_identical():
    g = np.load(os.path.join(GOLD, "ibm_bodies.npz"))
    for i in range(3):
        dh = float(g[f"circle{i}_dh"])
        for kind, body in (("circle", ib.Circle([0, 0], [0, 0], 0.5)), ("line", ib.Line([0, 0], [0, 0])),
                           ("box", ib.OpenBox([0, 0], [0, 0]))):
            xy, dl = body.generateBody(dh)
            assert np.array_equal(xy, g[f"{kind}{i}_xy"]), (kind, i)
            assert dl == float(g[f"{kind}{i}_dl"])


def test_prescribed_motion_bit_identical():
    g = np.load(os.path.join(GOLD, "ibm_bodies.npz"))
    body = ib.Circle([0, 0], list(g["mover_start"]), 0.5)
    body.setIsMoving()
    body.setVelRef(float(g["mover_uref"]))
    body.updateVelocity = lambda: None            # the velocity Vec lives on the device
    for t, c, v in zip(g["mover_t"], g["mover_center"], g["mover_vel"]):
        body.updateBodyParameters(float(t))
        assert np.array_equal(np.array(body.getCenterBody(), dtype=float), c)
        assert np.array_equal(np.array(body._vel, dtype=float), v)


def test_affected_cells_of_the_case_class():
    from pynama_b200.cases.immersed_boundary import ImmersedBoundaryStatic
    from pynama_b200.run_case import loadCase
    fem = ImmersedBoundaryStatic(loadCase("ibm-static"), case="ibm-static", nelem=[10, 10], ngl=3,
                                 lower=[-5, -5], upper=[5, 5])
    fem.setUpDomain()
    assert abs(fem.nodeSeparation - 0.5) < 1e-15
    assert len(fem.getAffectedCells(1)) == 4 and len(fem.getAffectedCells(2, 2)) == 16
    assert len(fem.getAffectedCells(1, 2)) == 8
    off = np.array([0.5, 0.5])
    assert len(fem.getAffectedCells(1, center=off)) == 1 and len(fem.getAffectedCells(2, 2, center=off)) == 9
    nodes, xy = fem.dom.getNodesCoordsFromEntities(fem.getAffectedCells(1, 2))
    assert len(nodes) == 5 * 9 and np.all(np.diff(nodes) > 0) and xy.shape == (45, 2)
