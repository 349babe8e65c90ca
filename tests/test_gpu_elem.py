"""GPU parity: element kernels (k1-k3) through the C ABI against the golden
vectors produced by the reference's own Spectral class and against the oracle.
Tolerance: |cuda - reference| <= 1e-12 * max|reference| per element matrix
(north_star: values within 1e-12 relative)."""
import numpy as np
import pytest

from oracle import sem
from tests.util import golden

pytestmark = pytest.mark.gpu
TOL = 1e-12

CASES = [(2, 2), (2, 3), (2, 4), (2, 5), (2, 8), (3, 2), (3, 3), (3, 4), (3, 5), (3, 6)]


def _mesh_of_cells(dim, ngl, cells):
    from pynama_b200.device_mesh import DeviceMesh
    nn = ngl ** dim
    ne = len(cells)
    elem_nodes = np.arange(ne * nn, dtype=np.int32).reshape(ne, nn)   # disjoint elements
    return DeviceMesh(dim, ngl, np.asarray(cells), elem_nodes, ne * nn)


def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("dim,ngl", CASES)
def test_elem_matrices_vs_reference_golden(dim, ngl):
    g = golden(f"elem_{dim}d_ngl{ngl}.npz")
    cells = g["cells"]
    dm = _mesh_of_cells(dim, ngl, cells)
    K, Rw, Rd = dm.elem_kle()
    S, D, C, W = dm.elem_ops()
    got = {"K": K, "Rw": Rw, "Rd": Rd, "SrT": S, "Div": D, "Curl": C}
    for ci in range(len(cells)):
        assert _rel(W[ci], g[f"c{ci}_W"]) <= TOL
        for name, arr in got.items():
            if f"c{ci}_{name}" in g:
                assert _rel(arr[ci], g[f"c{ci}_{name}"]) <= TOL, (name, ci)
            else:   # large orders: seeded sample of entries + row/column sums
                i, j = g[f"c{ci}_{name}_i"], g[f"c{ci}_{name}_j"]
                ref = g[f"c{ci}_{name}_v"]
                scale = max(np.abs(ref).max(), 1e-300)
                assert np.abs(arr[ci][i, j] - ref).max() <= TOL * scale, (name, ci)
                assert np.abs(arr[ci].sum(1) - g[f"c{ci}_{name}_rowsum"]).max() <= 1e-10 * scale * arr[ci].shape[1] ** 0.5
                assert np.abs(arr[ci].sum(0) - g[f"c{ci}_{name}_colsum"]).max() <= 1e-10 * scale * arr[ci].shape[0] ** 0.5


@pytest.mark.parametrize("dim,ngl,nel", [(2, 3, 40), (2, 6, 9), (3, 3, 30), (3, 4, 12)])
def test_elem_matrices_vs_oracle_random_cells(dim, ngl, nel):
    """non-affine (perturbed-corner) cells, every element different; also checks that
    element batches and the zero structure of the collocated operators are right"""
    from pynama_b200.domain.box_mesh import CORNERS
    rng = np.random.default_rng(100 * dim + ngl)
    base = CORNERS[dim].astype(float)
    cells = np.array([base * rng.uniform(0.2, 1.5, dim) + rng.uniform(-1, 1, dim) +
                      0.12 * rng.uniform(-1, 1, base.shape) for _ in range(nel)])
    tab = sem.SpectralTables(ngl, dim)
    dm = _mesh_of_cells(dim, ngl, cells)
    K, Rw, Rd = dm.elem_kle()
    S, D, C, W = dm.elem_ops()
    K2, _, _ = dm.elem_kle(e0=nel // 2, ne=2)            # a sub-range gives the same blocks
    assert np.array_equal(K2, K[nel // 2:nel // 2 + 2])
    for e in range(nel):
        oK, oRw, oRd = sem.elem_kle(tab, cells[e])
        oS, oD, oC, oW = sem.elem_ops(tab, cells[e])
        for name, a, b in (("K", K[e], oK), ("Rw", Rw[e], oRw), ("Rd", Rd[e], oRd), ("SrT", S[e], oS),
                           ("Div", D[e], oD), ("Curl", C[e], oC), ("W", W[e], oW)):
            assert _rel(a, b) <= TOL, (name, e)
        # explicit zeros of the reference are exact zeros here too
        for a, b in ((S[e], oS), (D[e], oD), (C[e], oC)):
            assert np.array_equal(a == 0.0, b == 0.0)
    # K_e is symmetric and annihilates rigid translations
    assert np.abs(K - K.transpose(0, 2, 1)).max() <= 1e-11 * np.abs(K).max()
    for c in range(dim):
        t = np.zeros(ngl ** dim * dim)
        t[c::dim] = 1.0
        assert np.abs(K @ t).max() <= 1e-9 * np.abs(K).max()
