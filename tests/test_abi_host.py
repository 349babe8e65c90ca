"""CPU: the C-ABI library loads and exports every symbol include/pynama_b200.h
declares; host-only helpers (1-D rules, Lagrange tables, local order) and the
host build of the element arithmetic (tests/hostcheck) agree with the golden
vectors of the reference.  No compute entry point is called (no GPU here)."""
import ctypes
import os
import re

import numpy as np
import pytest

from pynama_b200 import _lib
from tests.util import golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pynama_b200.h")
HOSTCHECK = os.path.join(ROOT, "tests", "hostcheck", "libhostcheck.so")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pyn_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 55
    cdll = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in names if not hasattr(cdll, n)]
    assert not missing, f"declared in the header but not exported: {missing}"
    # the ctypes prototypes cover the header and nothing else
    assert sorted(_lib.EXPORTS) == names


def test_version_and_error_plumbing():
    l = _lib.lib()
    assert l.pyn_version() >= 100
    x = np.zeros(4)
    with pytest.raises(_lib.PynamaError):
        _lib.call("pyn_rule_1d", 7, 4, _lib.ptr(x), _lib.ptr(x))     # unknown rule kind
    assert l.pyn_last_error()


def test_host_rules_vs_reference():
    g = golden("rules_1d.npz")
    for n in range(2, 9):
        x, w = _lib.rule_1d("gauss", n)
        assert np.abs(x - g[f"gauss{n}_x"]).max() < 1e-14 and np.abs(w - g[f"gauss{n}_w"]).max() < 1e-14
        x, w = _lib.rule_1d("lobatto", n)
        assert np.abs(x - g[f"lobatto{n}_x"]).max() < 1e-14 and np.abs(w - g[f"lobatto{n}_w"]).max() < 1e-14


@pytest.mark.parametrize("dim,ngl", [(2, 2), (2, 3), (2, 5), (2, 8), (3, 2), (3, 3), (3, 4), (3, 6)])
def test_host_local_order_vs_reference(dim, ngl):
    g = golden(f"elem_{dim}d_ngl{ngl}.npz")
    nodes, _ = _lib.rule_1d("lobatto", ngl)
    t = _lib.local_order(dim, ngl)
    assert np.abs(nodes[t] - g["node_rst"]).max() < 1e-14


def _hc():
    if not os.path.exists(HOSTCHECK):
        pytest.skip("tests/hostcheck/libhostcheck.so not built (run __graft_entry__.build())")
    return ctypes.CDLL(HOSTCHECK)


@pytest.mark.parametrize("dim,ngl", [(2, 2), (2, 3), (2, 4), (2, 5), (3, 2), (3, 3), (3, 4)])
def test_element_arithmetic_host_build_vs_reference(dim, ngl):
    """the same inline functions the CUDA kernels call, run in CPU loops"""
    hc = _hc()
    g = golden(f"elem_{dim}d_ngl{ngl}.npz")
    nn, dw, ds = ngl ** dim, (1 if dim == 2 else 3), (3 if dim == 2 else 6)
    vp = ctypes.c_void_p
    for ci, X in enumerate(g["cells"]):
        X = np.ascontiguousarray(X, dtype=np.float64)
        K, Rw, Rd = np.zeros((nn * dim, nn * dim)), np.zeros((nn * dim, nn * dw)), np.zeros((nn * dim, nn))
        S, D = np.zeros((nn * ds, nn * dim)), np.zeros((nn * dim, nn * ds))
        C, W = np.zeros((nn * dw, nn * dim)), np.zeros(nn)
        assert hc.hc_kle(dim, ngl, X.ctypes.data_as(vp), K.ctypes.data_as(vp), Rw.ctypes.data_as(vp),
                         Rd.ctypes.data_as(vp)) == 0
        assert hc.hc_ops(dim, ngl, X.ctypes.data_as(vp), S.ctypes.data_as(vp), D.ctypes.data_as(vp),
                         C.ctypes.data_as(vp), W.ctypes.data_as(vp)) == 0
        for name, arr in (("K", K), ("Rw", Rw), ("Rd", Rd), ("SrT", S), ("Div", D), ("Curl", C), ("W", W)):
            ref = g[f"c{ci}_{name}"]
            err = np.abs(arr - ref).max() / np.abs(ref).max()
            assert err <= 1e-12, (name, ci, err)


def _classes(mesh_corners, dim, tol):
    hc = _hc()
    hc.hc_congruence_classes.restype = ctypes.c_longlong
    hc.hc_congruence_classes.argtypes = [ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_double,
                                         ctypes.c_void_p, ctypes.c_void_p]
    cx = np.ascontiguousarray(mesh_corners, dtype=np.float64).reshape(-1, 2 ** dim, dim)
    cls, rep = np.zeros(len(cx), np.int32), np.zeros(len(cx), np.int32)
    n = hc.hc_congruence_classes(dim, len(cx), cx.ctypes.data, tol, cls.ctypes.data, rep.ctypes.data)
    return n, cls, rep[:n]


def test_congruence_classes_host():
    """pynama_b200/csrc/classes.cuh (the host code behind pyn_mesh_set_congruence) on the box
    generator's meshes: uniform = one class, stretched = one per element size, displaced corners
    = never merged, tol <= 0 = off; representatives are the lowest elements of their classes"""
    from pynama_b200.domain.box_mesh import BoxMesh
    for nelem, lo, up in (([5, 4], [0, 0], [1, 1]), ([4, 3, 5], [0, 0, 0], [3, 1, 1]), ([16, 16, 16], [0] * 3, [1] * 3)):
        m = BoxMesh(nelem, lo, up, 3)
        n, cls, rep = _classes(m.corner_xyz, m.dim, 1e-13)
        assert n == 1 and not cls.any() and rep.tolist() == [0]
        n, cls, rep = _classes(m.corner_xyz, m.dim, 0.0)
        assert n == m.nel and np.array_equal(cls, np.arange(m.nel)) and np.array_equal(rep, np.arange(m.nel))
    # every cell of these boxes touches a displaced interior corner
    m = BoxMesh([3, 3, 2], [0, 0, 0], [1, 1, 1], 3, perturb=0.15)
    assert _classes(m.corner_xyz, 3, 1e-13)[0] == m.nel
    # cells away from the displaced corners keep a common class: 6x6 with only the middle displaced
    m = BoxMesh([6, 6], [0, 0], [1, 1], 2)
    cx = m.corner_xyz.copy().reshape(m.nel, 4, 2)
    hit = np.all(np.isclose(cx, [0.5, 0.5]), axis=2)
    cx[hit] += [0.01, -0.02]
    n, cls, rep = _classes(cx, 2, 1e-13)
    assert n == 5 and (cls == 0).sum() == m.nel - 4
    # two element sizes along x
    m = BoxMesh([4, 2], [0.0, 0.0], [1.0, 1.0], 3)
    cx = m.corner_xyz.copy().reshape(m.nel, 4, 2)
    cx[..., 0] = np.where(cx[..., 0] > 0.5, 0.5 + 2.0 * (cx[..., 0] - 0.5), cx[..., 0])
    n, cls, rep = _classes(cx, 2, 1e-13)
    assert n == 2
    for k in range(n):
        assert rep[k] == np.nonzero(cls == k)[0].min()
    # translation and tiny noise do not split a class, a visible change does
    m = BoxMesh([3, 3], [10.0, -7.0], [11.0, -6.0], 2)
    rng = np.random.default_rng(1)
    noisy = m.corner_xyz + 1e-15 * rng.uniform(-1, 1, m.corner_xyz.shape)
    assert _classes(noisy, 2, 1e-13)[0] == 1
    assert _classes(m.corner_xyz + 1e-9 * rng.uniform(-1, 1, m.corner_xyz.shape), 2, 1e-13)[0] == m.nel
    # degenerate input: no merging, no crash
    assert _classes(np.zeros((4, 4, 2)), 2, 1e-13)[0] == 4
    bad = m.corner_xyz.copy()
    bad[0, 0, 0] = np.nan
    assert _classes(bad, 2, 1e-13)[0] == m.nel


# ---- host arithmetic of the element-block Schwarz preconditioner (pynama_b200/csrc/asm_host.cuh) ----
def test_asm_dense_inverse_host():
    hc = _hc()
    hc.hc_asm_dense_inverse.restype = ctypes.c_int
    rng = np.random.default_rng(3)
    for n in (1, 2, 7, 48, 192):
        q = rng.standard_normal((n, n))
        spd = q @ q.T + n * np.eye(n)
        # a matrix that needs row exchanges: zero leading entry
        gen = rng.standard_normal((n, n)) + 0.1 * np.eye(n)
        gen[0, 0] = 0.0 if n > 1 else 2.0
        for a, sym in ((spd, 1), (gen, 0)):
            inv = np.ascontiguousarray(a.copy())
            assert hc.hc_asm_dense_inverse(n, inv.ctypes.data_as(ctypes.c_void_p), sym) == 0
            assert np.abs(inv @ a - np.eye(n)).max() < 1e-9 * np.linalg.cond(a)
            np.testing.assert_allclose(inv, np.linalg.inv(a), rtol=0, atol=1e-10 * np.abs(np.linalg.inv(a)).max()
                                       * np.linalg.cond(a))
            if sym:
                assert np.array_equal(inv, inv.T)          # exactly symmetric: CG needs it
    # K rows of constrained dofs are identity rows: the inverse keeps them
    a = np.eye(5)
    a[1:4, 1:4] = [[4, 1, 0], [1, 3, 1], [0, 1, 2]]
    inv = a.copy()
    assert hc.hc_asm_dense_inverse(5, inv.ctypes.data_as(ctypes.c_void_p), 1) == 0
    assert inv[0, 0] == 1.0 and inv[4, 4] == 1.0 and not inv[0, 1:].any() and not inv[4, :4].any()
    # singular, zero, non-finite: refused
    for bad in (np.zeros((3, 3)), np.ones((3, 3)) * np.nan, np.array([[1.0, 2.0], [2.0, 4.0]]) * 0 + [[0, 0], [0, 1]]):
        b = np.ascontiguousarray(bad, dtype=np.float64)
        assert hc.hc_asm_dense_inverse(b.shape[0], b.ctypes.data_as(ctypes.c_void_p), 0) == -1


def _block_hash(hc, a, perm=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    out = (ctypes.c_ulonglong * 2)()
    p = None
    if perm is not None:
        perm = np.ascontiguousarray(perm, dtype=np.int64)
        p = perm.ctypes.data_as(ctypes.c_void_p)
    hc.hc_asm_block_hash(a.shape[0], a.ctypes.data_as(ctypes.c_void_p), p, out)
    return out[0], out[1]


def test_asm_block_hash_and_grouping_host():
    """the device hashes an element block with one warp per row node and atomic integer adds, so the
    hash must not depend on the visiting order, must see the position of an entry and its last bit"""
    hc = _hc()
    hc.hc_asm_group_by_hash.restype = ctypes.c_longlong
    rng = np.random.default_rng(5)
    ld = 24
    a = rng.standard_normal((ld, ld))
    a[rng.uniform(size=a.shape) < 0.3] = 0.0
    h = _block_hash(hc, a)
    assert h == _block_hash(hc, a, rng.permutation(ld * ld))
    b = a.copy()
    i, j = np.argwhere(a != 0)[5]
    b[i, j] = np.nextafter(b[i, j], np.inf)
    assert _block_hash(hc, b) != h                                  # one ulp
    assert _block_hash(hc, a.T.copy()) != h                         # same values elsewhere
    z = a.copy()
    z[a == 0] = -0.0
    assert _block_hash(hc, z) == h                                  # signed zeros are structural zeros
    h0, h1 = h
    assert h0 != h1
    # grouping: classes numbered by lowest member, representatives = lowest members
    blocks = [a, b, a.T.copy(), a, b, a, np.zeros((ld, ld))]
    hh = np.array([_block_hash(hc, x) for x in blocks], dtype=np.uint64)
    cls = np.zeros(len(blocks), np.int32)
    rep = np.zeros(len(blocks), np.int32)
    n = hc.hc_asm_group_by_hash(len(blocks), hh.ctypes.data_as(ctypes.c_void_p), cls.ctypes.data_as(ctypes.c_void_p),
                                rep.ctypes.data_as(ctypes.c_void_p))
    assert n == 4 and cls.tolist() == [0, 1, 2, 0, 1, 0, 3] and rep[:4].tolist() == [0, 1, 2, 6]
    # agreeing in one of the two hashes is not enough
    hh2 = np.array([[1, 2], [1, 3], [1, 2]], dtype=np.uint64)
    n = hc.hc_asm_group_by_hash(3, hh2.ctypes.data_as(ctypes.c_void_p), cls.ctypes.data_as(ctypes.c_void_p),
                                rep.ctypes.data_as(ctypes.c_void_p))
    assert n == 2 and cls[:3].tolist() == [0, 1, 0]


def test_asm_tiles_host():
    hc = _hc()
    hc.hc_asm_make_tiles.restype = ctypes.c_longlong
    rng = np.random.default_rng(9)
    for nel, nclass, te in ((1, 1, 16), (64, 27, 16), (4096, 27, 16), (37, 37, 16), (100, 3, 4)):
        cls = rng.integers(0, nclass, nel).astype(np.int32)
        cls[:min(nel, nclass)] = np.arange(min(nel, nclass))        # every class used
        tile_elems = np.full((nel + nclass) * te, -7, np.int32)
        tile_class = np.full(nel + nclass, -7, np.int32)
        nt = hc.hc_asm_make_tiles(nel, cls.ctypes.data_as(ctypes.c_void_p), nclass, te,
                                  tile_elems.ctypes.data_as(ctypes.c_void_p),
                                  tile_class.ctypes.data_as(ctypes.c_void_p))
        want = sum(-(-int((cls == c).sum()) // te) for c in range(nclass))
        assert nt == want
        te_ = tile_elems[:nt * te].reshape(nt, te)
        real = te_[te_ >= 0]
        assert np.array_equal(np.sort(real), np.arange(nel))        # every element exactly once
        assert set(te_[te_ < 0].tolist()) <= {-1}
        for t in range(nt):
            m = te_[t][te_[t] >= 0]
            assert len(m) > 0 and np.all(cls[m] == tile_class[t]) and np.all(np.diff(m) > 0)
            assert np.all(te_[t][len(m):] == -1)                    # padding only at the end
        assert np.all(np.diff(tile_class[:nt]) >= 0)


def test_asm_coarse_tables_and_colouring_host():
    """the coarse vertices, their adjacency and the distance-2 colouring the Galerkin probing relies
    on (no vertex sees two vertices of one colour), on boxes and on an element-scrambled mesh"""
    from pynama_b200.domain.box_mesh import BoxMesh
    hc = _hc()
    hc.hc_asm_coarse_tables.restype = ctypes.c_longlong
    rng = np.random.default_rng(13)
    for nelem, ngl, scramble in (([4, 3], 3, False), ([3, 3, 3], 3, False), ([5, 4, 3], 4, True), ([7, 6], 5, True)):
        m = BoxMesh(nelem, [0.0] * len(nelem), [1.0] * len(nelem), ngl)
        en = np.ascontiguousarray(m.elem_nodes, dtype=np.int32)
        if scramble:                                     # element order of an unstructured mesh
            en = np.ascontiguousarray(en[rng.permutation(len(en))])
        dim, nel, nn = m.dim, len(en), en.shape[1]
        ncorner = 2 ** dim
        vert_node = np.zeros(m.nnodes, np.int32)
        elem_verts = np.zeros(nel * ncorner, np.int32)
        color = np.zeros(m.nnodes, np.int32)
        adj_ptr = np.zeros(m.nnodes + 1, np.int64)
        adj = np.zeros(m.nnodes * 125, np.int32)
        info = np.zeros(1, np.int32)
        phi = np.zeros(nn * ncorner)
        vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)          # noqa: E731
        nv = hc.hc_asm_coarse_tables(dim, ngl, nel, m.nnodes, vp(en), vp(vert_node), vp(elem_verts), vp(color),
                                     vp(adj_ptr), vp(adj), vp(info), vp(phi))
        assert nv == np.prod(np.array(nelem) + 1)
        vert_node, color, adj_ptr = vert_node[:nv], color[:nv], adj_ptr[:nv + 1]
        assert np.all(np.diff(vert_node) > 0)
        phi = phi.reshape(nn, ncorner)
        np.testing.assert_allclose(phi.sum(1), 1.0, atol=1e-13)      # partition of unity
        assert ((phi == 1.0).sum(0) == 1).all()                      # each corner function is 1 at its own node
        # P reproduces the coordinates (multilinear interpolation from the corners is exact on boxes)
        ev = elem_verts.reshape(nel, ncorner)
        for e in (0, nel // 2, nel - 1):
            np.testing.assert_allclose(phi @ m.coords[vert_node[ev[e]]], m.coords[en[e]], atol=1e-13)
        nbrs = [set(adj[adj_ptr[v]:adj_ptr[v + 1]].tolist()) for v in range(nv)]
        for v in range(nv):
            assert v in nbrs[v] and all(v in nbrs[w] for w in nbrs[v])          # symmetric, reflexive
            cols = color[sorted(nbrs[v])]
            assert len(set(cols.tolist())) == len(cols)                          # the probing property
        assert info[0] == color.max() + 1 and info[0] <= 5 ** dim


def test_asm_chebyshev_coefficients_host():
    """k steps with the coefficients reproduce the Chebyshev polynomial: on a diagonal matrix with
    spectrum inside [lmin, lmax] the error polynomial stays below 2 / (T_k(sigma)) everywhere"""
    hc = _hc()
    lmin, lmax, k = 0.1, 1.1, 5
    c1, c2 = np.zeros(k + 1), np.zeros(k + 1)
    hc.hc_asm_cheb_coefficients(k, ctypes.c_double(lmin), ctypes.c_double(lmax), c1.ctypes.data_as(ctypes.c_void_p),
                                c2.ctypes.data_as(ctypes.c_void_p))
    lam = np.linspace(lmin, lmax, 101)
    b = np.ones_like(lam)
    x, r, d = np.zeros_like(lam), b.copy(), c2[0] * b
    for s in range(1, k + 1):
        x += d
        r -= lam * d
        d = c1[s] * d + c2[s] * r
    theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)
    bound = 1.0 / np.cosh(k * np.arccosh(theta / delta))
    assert np.abs(1.0 - lam * x).max() <= bound * (1 + 1e-10)        # residual polynomial r_k = T_k(..)/T_k(sigma)
    assert np.allclose(r, 1.0 - lam * x)
