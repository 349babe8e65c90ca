"""CPU: the oracle (oracle/sem.py) against the golden vectors generated from the
reference's own Spectral class (tests/golden/make_golden.py), and against the
closed-form rules the reference's tests pin (src/tests/test_element.py:181-229).
Tolerance 1e-12 relative to the largest entry (north_star)."""
import numpy as np
import pytest

from oracle import sem
from tests.util import golden

TOL = 1e-12
CASES = [(2, 2), (2, 3), (2, 4), (2, 5), (2, 8), (3, 2), (3, 3), (3, 4), (3, 5), (3, 6)]


def _rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300)


def test_rules_1d_vs_reference():
    g = golden("rules_1d.npz")
    for n in range(2, 9):
        x, w = sem.gauss_legendre(n)
        assert np.abs(x - g[f"gauss{n}_x"]).max() < 1e-14
        assert np.abs(w - g[f"gauss{n}_w"]).max() < 1e-14
        x, w = sem.gauss_lobatto(n)
        assert np.abs(x - g[f"lobatto{n}_x"]).max() < 1e-14
        assert np.abs(w - g[f"lobatto{n}_w"]).max() < 1e-14


def test_rules_closed_forms():
    """src/tests/test_element.py:181-229"""
    x, w = sem.gauss_legendre(2)
    assert np.allclose(np.sort(x), [-1 / np.sqrt(3), 1 / np.sqrt(3)], atol=1e-14) and np.allclose(w, 1.0)
    x, w = sem.gauss_legendre(3)
    assert np.allclose(np.sort(x), [-np.sqrt(0.6), 0, np.sqrt(0.6)], atol=1e-14)
    assert np.allclose(np.sort(w), [5 / 9, 5 / 9, 8 / 9], atol=1e-14)
    x, w = sem.gauss_lobatto(3)
    assert np.allclose(x, [-1, 0, 1], atol=1e-14) and np.allclose(w, [1 / 3, 4 / 3, 1 / 3], atol=1e-14)
    x, w = sem.gauss_lobatto(4)
    assert np.allclose(x, [-1, -np.sqrt(0.2), np.sqrt(0.2), 1], atol=1e-14)
    assert np.allclose(w, [1 / 6, 5 / 6, 5 / 6, 1 / 6], atol=1e-14)


@pytest.mark.parametrize("dim,ngl", CASES)
def test_tables_vs_reference(dim, ngl):
    g = golden(f"elem_{dim}d_ngl{ngl}.npz")
    tab = sem.SpectralTables(ngl, dim)
    # local node order: reference coordinates of every node
    assert np.abs(tab.nodal.pts - g["node_rst"]).max() < 1e-14
    for rule, key in ((tab.full, "gps"), (tab.red, "gpsRed"), (tab.nodal, "gpsOp")):
        assert np.abs(rule.pts - g[key][:, :dim]).max() < 1e-14
        assert np.abs(rule.w - g[key][:, dim]).max() < 1e-14
    for rule, key in ((tab.full, "HrsCoo"), (tab.red, "HrsCooRed"), (tab.nodal, "HrsCooOp")):
        assert _rel(rule.Hrsc, g[key]) < 1e-14
    assert _rel(tab.nodal.Hc, g["HCooOp"].reshape(tab.nodal.Hc.shape)) < 1e-14
    if "H" in g:
        for rule, hk, dk in ((tab.full, "H", "Hrs"), (tab.red, "HRed", "HrsRed"), (tab.nodal, "HOp", "HrsOp")):
            assert np.abs(rule.H - g[hk].reshape(rule.H.shape)).max() < 1e-13
            assert np.abs(rule.Hrs - g[dk]).max() < 1e-12 * max(np.abs(g[dk]).max(), 1.0)


@pytest.mark.parametrize("dim,ngl", CASES)
def test_elem_matrices_vs_reference(dim, ngl):
    g = golden(f"elem_{dim}d_ngl{ngl}.npz")
    tab = sem.SpectralTables(ngl, dim)
    for ci, X in enumerate(g["cells"]):
        K, Rw, Rd = sem.elem_kle(tab, X)
        S, D, C, W = sem.elem_ops(tab, X)
        assert _rel(W, g[f"c{ci}_W"]) <= TOL
        for name, arr in (("K", K), ("Rw", Rw), ("Rd", Rd), ("SrT", S), ("Div", D), ("Curl", C)):
            if f"c{ci}_{name}" in g:
                ref = g[f"c{ci}_{name}"]
                assert _rel(arr, ref) <= TOL, (name, ci)
            else:
                i, j, ref = g[f"c{ci}_{name}_i"], g[f"c{ci}_{name}_j"], g[f"c{ci}_{name}_v"]
                scale = max(np.abs(ref).max(), 1e-300)
                assert np.abs(arr[i, j] - ref).max() <= TOL * scale, (name, ci)
                assert np.abs(arr.sum(1) - g[f"c{ci}_{name}_rowsum"]).max() <= 1e-10 * scale * arr.shape[1] ** 0.5


def test_unit_cell_of_reference_tests():
    """K_e properties on the unit cell with the corner order of src/tests/test_element.py:270"""
    X = np.array([1, 1, 1, 0, 1, 1, 0, 0, 1, 1, 0, 1, 1, 1, 0, 1, 0, 0, 0, 0, 0, 0, 1, 0], float).reshape(8, 3)
    tab = sem.SpectralTables(2, 3)
    K, Rw, Rd = sem.elem_kle(tab, X)
    assert np.abs(K - K.T).max() < 1e-12 * np.abs(K).max()
    for c in range(3):                      # rigid translations are in the null space
        t = np.zeros(24)
        t[c::3] = 1.0
        assert np.abs(K @ t).max() < 1e-10 * np.abs(K).max()
    S, D, C, W = sem.elem_ops(tab, X)
    assert abs(abs(W.sum()) - 1.0) < 1e-13  # lumped mass sums to the (signed) cell volume
