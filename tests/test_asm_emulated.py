"""The CUDA kernels of the element-block additive Schwarz preconditioner (-pc_type asm,
pynama_b200/csrc/asm_kernels.cuh) run on CPU threads through tests/hostcheck/cuda_emu.h and
compared with a numpy statement of the same preconditioner on the oracle's K.  There is no GPU in
the build container: this is how the kernels' indexing, tiling, padding and reduction logic is
checked before they reach one.  Operators: the oracle's free-slip KLE matrix,
converted to the product's row-plane BSR layout (rows padded to a multiple of 4 blocks)."""
import ctypes
import os

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import timestep
from pynama_b200.domain.box_mesh import BoxMesh
from tests.study_precond import element_schwarz

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU = os.path.join(ROOT, "tests", "hostcheck", "libemu_asm.so")


def _emu():
    if not os.path.exists(EMU):
        pytest.skip("tests/hostcheck/libemu_asm.so not built (run __graft_entry__.build())")
    lib = ctypes.CDLL(EMU)
    lib.emu_asm_apply.restype = ctypes.c_int
    return lib


def row_plane_bsr(K, B, pad=4, drop_zero_blocks=True):
    """scipy matrix -> (rowptr, colidx, val) in the layout of pyn_mat (mesh.cuh): block columns
    ascending, rows padded to a multiple of `pad` blocks by zero blocks repeating the last column,
    val[p0*B*B + q*nb + u]"""
    K = K.tocsr().copy()
    if drop_zero_blocks:
        K.eliminate_zeros()
    M = sp.bsr_matrix(K, blocksize=(B, B))
    M.sort_indices()
    nrows = M.shape[0] // B
    rowptr = np.zeros(nrows + 1, np.int64)
    cols, vals = [], []
    for r in range(nrows):
        lo, hi = M.indptr[r], M.indptr[r + 1]
        c = M.indices[lo:hi].astype(np.int32)
        d = M.data[lo:hi]                              # [nb][B][B]
        npad = (-len(c)) % pad if len(c) else 0
        if npad:
            c = np.concatenate([c, np.full(npad, c[-1], np.int32)])
            d = np.concatenate([d, np.zeros((npad, B, B))])
        nb = len(c)
        rowptr[r + 1] = rowptr[r] + nb
        cols.append(c)
        vals.append(d.reshape(nb, B * B).T.ravel())    # plane q = i*B+j holds nb doubles
    return rowptr, np.concatenate(cols), np.concatenate(vals)


def node_to_elements(elem_nodes, nnodes):
    nel, nn = elem_nodes.shape
    pairs = [[] for _ in range(nnodes)]
    for e in range(nel):
        for a in range(nn):
            pairs[elem_nodes[e, a]].append((e, a))
    ptr = np.zeros(nnodes + 1, np.int64)
    ptr[1:] = np.cumsum([len(p) for p in pairs])
    flat = np.array([x for p in pairs for x in p], dtype=np.int32).reshape(-1, 2)
    return ptr, flat


def _vp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class _LocalMesh:
    """what one rank of a row-partitioned run holds (pynama_b200/parallel.py)"""

    def __init__(self, mesh, part):
        self.dim, self.ngl = mesh.dim, mesh.ngl
        self.elem_nodes, self.nnodes, self.n_own = part.elem_nodes, part.n_loc, part.n_own


def run_emulated(mesh, K, B, r, grid=3, napply=1, want_inv=False, drop_zero_blocks=True, coarse_steps=0,
                 want_ac=False, info=None, n_own=None, valence_global=None, nodecls=None, quantum_rel=0.0):
    lib = _emu()
    nel, nn = mesh.elem_nodes.shape
    nnodes = mesh.nnodes
    en = np.ascontiguousarray(mesh.elem_nodes, dtype=np.int32)
    rowptr, colidx, val = row_plane_bsr(K, B, drop_zero_blocks=drop_zero_blocks)
    ptr, n2e = node_to_elements(en, nnodes)
    z = np.full(nnodes * B, np.nan)
    dots = np.zeros(2)
    nclass = ctypes.c_longlong()
    cls = np.zeros(nel, np.int32)
    ld = nn * B
    inv = np.zeros((nel, ld, ld)) if want_inv else None
    r = np.ascontiguousarray(r, dtype=np.float64)
    cinfo = np.zeros(2, np.int64)
    ac = None
    if want_ac:
        nv_max = nel * 2 ** mesh.dim
        ac = np.zeros((nv_max * B) ** 2)
    rc = lib.emu_asm_apply(B, nn, nel, nnodes, _vp(en), _vp(rowptr), _vp(colidx), _vp(val), _vp(ptr), _vp(n2e), _vp(r),
                           grid, napply, _vp(z), _vp(dots), ctypes.byref(nclass), _vp(cls),
                           _vp(inv) if want_inv else None, mesh.dim, mesh.ngl, coarse_steps, _vp(cinfo),
                           _vp(ac) if want_ac else None, ctypes.c_longlong(nnodes if n_own is None else n_own),
                           None if valence_global is None else _vp(np.ascontiguousarray(valence_global, dtype=np.float64)),
                           None if nodecls is None else _vp(np.ascontiguousarray(nodecls, dtype=np.uint8)), ctypes.c_double(quantum_rel))
    assert rc == 0, f"emulated pipeline failed an internal check ({rc})"
    if info is not None:
        info["nv"], info["ncolors"] = int(cinfo[0]), int(cinfo[1])
        if want_ac:
            nc = int(cinfo[0]) * B
            info["ac"] = ac[:nc * nc].reshape(nc, nc)
    return z, dots, nclass.value, cls, (inv[:nclass.value] if want_inv else None)


def kle_operator(dim, nelem, ngl):
    mesh = BoxMesh(nelem, [0.0] * dim, [1.0] * dim, ngl)
    prob = timestep.OracleProblem(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes(),
                                  rho=0.5, mu=0.01, solver="direct")
    return mesh, prob.K.tocsr()


def element_blocks(K, elem_nodes, B):
    nel, nn = elem_nodes.shape
    dofs = (elem_nodes[:, :, None] * B + np.arange(B)[None, None, :]).reshape(nel, nn * B)
    return np.stack([K[dofs[e]][:, dofs[e]].toarray() for e in range(nel)])


CASES = [
    # dim, elements, ngl          ld    what the shape exercises
    (3, [3, 3, 2], 3),          # 81    partial panel, partial last k-chunk, 18 elements over several CTAs
    (3, [2, 2, 2], 4),          # 192   exactly one full panel of 192 rows, 12 full k-chunks (the bench shape)
    (3, [2, 1, 1], 5),          # 375   two panels (192 + 183), 24 k-chunks with a partial one
    (2, [5, 4], 3),             # 18    2x2 blocks, rows far below one panel, 20 elements: tiles of several classes
    (2, [3, 2], 4),             # 32
]


@pytest.mark.parametrize("dim,nelem,ngl", CASES)
def test_emulated_kernels_match_numpy_schwarz(dim, nelem, ngl):
    mesh, K = kle_operator(dim, nelem, ngl)
    B = dim
    nel = mesh.elem_nodes.shape[0]
    rng = np.random.default_rng(11)
    r = rng.standard_normal(K.shape[0])
    z, dots, nclass, cls, inv = run_emulated(mesh, K, B, r, grid=3, want_inv=True)
    ref_apply, _ = element_schwarz(K, mesh.elem_nodes, B, weighted=True)
    zref = ref_apply(r)
    scale = np.abs(zref).max()
    assert np.abs(z - zref).max() < 1e-11 * scale
    assert abs(dots[0] - r @ zref) < 1e-10 * abs(r @ zref)
    assert abs(dots[1] - zref @ zref) < 1e-10 * (zref @ zref)
    assert dots[0] > 0                                    # SPD
    # classes: exactly the sets of bitwise equal element blocks, numbered by lowest member
    blocks = element_blocks(K, mesh.elem_nodes, B)
    assert 1 <= nclass <= nel
    for e in range(nel):
        for f in range(e):
            assert (cls[e] == cls[f]) == np.array_equal(blocks[e], blocks[f])
    first = [int(np.nonzero(cls == c)[0].min()) for c in range(nclass)]
    assert first == sorted(first)
    # the inverses the kernels were given
    for c in range(nclass):
        a = blocks[first[c]]
        assert np.abs(inv[c] @ a - np.eye(a.shape[0])).max() < 1e-8
        assert np.array_equal(inv[c], inv[c].T)


def test_emulated_apply_is_symmetric_reproducible_and_grid_independent():
    mesh, K = kle_operator(3, [3, 2, 2], 3)
    rng = np.random.default_rng(2)
    n = K.shape[0]
    r1, r2 = rng.standard_normal(n), rng.standard_normal(n)
    z1, d1, _, _, _ = run_emulated(mesh, K, 3, r1, grid=2)
    z2, _, _, _, _ = run_emulated(mesh, K, 3, r2, grid=2)
    assert abs(r2 @ z1 - r1 @ z2) < 1e-12 * abs(r2 @ z1)           # M^-1 symmetric: CG may use it
    # bitwise the same for any grid size and on repeated applies (fixed summation orders)
    for grid, napply in ((1, 1), (5, 2)):
        z, d, _, _, _ = run_emulated(mesh, K, 3, r1, grid=grid, napply=napply)
        assert np.array_equal(z, z1)
        assert d[1] == pytest.approx(d1[1], rel=1e-14)              # the dots fold per-CTA partials
    # explicit zero blocks stored in the matrix (nothing dropped) change nothing
    z, _, _, _, _ = run_emulated(mesh, K, 3, r1, grid=2, drop_zero_blocks=False)
    assert np.array_equal(z, z1)


def test_emulated_preconditioner_cuts_cg_iterations():
    """what the option is for: fewer CG iterations than node-block Jacobi on the bench operator's
    small sibling (tests/study_precond.py has the table for 4^3 and 8^3 elements)"""
    import scipy.sparse.linalg as spl
    from tests.study_precond import node_block_jacobi
    mesh, K = kle_operator(3, [3, 3, 3], 3)
    n = K.shape[0]
    rng = np.random.default_rng(4)
    b = K @ rng.standard_normal(n)

    def count(M):
        its = [0]

        def cb(_):
            its[0] += 1
        x, info = spl.cg(K, b, rtol=1e-10, atol=0.0, maxiter=2000, M=spl.LinearOperator((n, n), matvec=M), callback=cb)
        assert info == 0 and np.linalg.norm(b - K @ x) < 1e-8 * np.linalg.norm(b)
        return its[0]

    its_bj = count(node_block_jacobi(K, 3))
    # (through the numpy statement the emulated kernels were shown equal to above: an emulated apply
    # repeats the whole set-up, far too slow inside a Krylov loop)
    ref_apply = element_schwarz(K, mesh.elem_nodes, 3, weighted=True)[0]
    z = run_emulated(mesh, K, 3, b, grid=2)[0]
    assert np.abs(z - ref_apply(b)).max() < 1e-11 * np.abs(z).max()
    its_asm = count(ref_apply)
    assert its_asm < 0.7 * its_bj, (its_asm, its_bj)


# ---- second level: Q1 coarse space on the element corners ------------------------------------------
def coarse_reference(mesh, K, B):
    """numpy statement: P = multilinear interpolation from the element corners, constrained fine
    dofs (identity rows of K) left out, Ac = P^T K P"""
    X = mesh.coords
    nel, nn = mesh.elem_nodes.shape
    dim = mesh.dim
    K = K.tocsr()
    offdiag = K - sp.diags(K.diagonal())
    keep = (np.asarray(abs(offdiag).sum(1)).ravel() > 0).astype(float)
    verts = {}
    rows, cols, vals = {}, None, None
    for e in range(nel):
        en = mesh.elem_nodes[e]
        lo, hi = X[en].min(0), X[en].max(0)
        xi = (X[en] - lo) / (hi - lo)
        corner = np.all((np.abs(xi) < 1e-12) | (np.abs(xi - 1) < 1e-12), axis=1)
        for c, tc in zip(en[corner], xi[corner]):
            w = np.prod(np.where(tc > 0.5, xi, 1.0 - xi), axis=1)
            for a, wa in zip(en, w):
                if wa > 1e-13:
                    rows[(int(a), int(c))] = wa
    vert_nodes = sorted({c for _, c in rows})
    vid = {c: i for i, c in enumerate(vert_nodes)}
    rr = [a for a, _ in rows]
    cc = [vid[c] for _, c in rows]
    P = sp.csr_matrix((list(rows.values()), (rr, cc)), shape=(len(X), len(vert_nodes)))
    P = sp.diags(keep) @ sp.kron(P, sp.identity(B)).tocsr()
    return P.tocsr(), (P.T @ K @ P).toarray(), keep


def chebyshev_reference(Ac, k):
    d = np.diag(Ac).copy()
    dinv = np.where(d > 1e-12 * d.max(), 1.0 / np.where(d == 0, 1, d), 0.0)
    lmax = 1.1 * np.abs(np.linalg.eigvals(dinv[:, None] * Ac)).max()
    lmin = 0.1 * lmax
    theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)

    def apply(b):
        x = np.zeros_like(b)
        r = b.copy()
        sigma = theta / delta
        rho = 1.0 / sigma
        dv = dinv * r / theta
        for _ in range(k):
            x = x + dv
            r = r - Ac @ dv
            rho_new = 1.0 / (2.0 * sigma - rho)
            dv = rho_new * rho * dv + 2.0 * rho_new / delta * (dinv * r)
            rho = rho_new
        return x
    return apply


@pytest.mark.parametrize("dim,nelem,ngl", [(3, [3, 2, 2], 3), (3, [2, 1, 1], 4), (2, [5, 4], 3), (2, [3, 3], 5)])
def test_emulated_coarse_level_matches_numpy(dim, nelem, ngl):
    mesh, K = kle_operator(dim, nelem, ngl)
    B = dim
    rng = np.random.default_rng(21)
    r = rng.standard_normal(K.shape[0])
    steps = 6
    info = {}
    z2, dots, _, _, _ = run_emulated(mesh, K, B, r, grid=2, coarse_steps=steps, want_ac=True, info=info)
    z1 = element_schwarz(K, mesh.elem_nodes, B, weighted=True)[0](r)       # the one-level part (checked above)
    P, Ac, keep = coarse_reference(mesh, K, B)
    assert info["nv"] * B == Ac.shape[0] == np.prod(np.array(nelem) + 1) * B
    # the probed Galerkin operator: one product with K per colour and component
    assert 2 ** dim <= info["ncolors"] <= 5 ** dim
    assert np.abs(info["ac"] - Ac).max() < 1e-11 * np.abs(Ac).max()
    assert np.array_equal(info["ac"], info["ac"].T)
    # z = one-level part + P cheb_k(Ac) P^T r; the power iteration's lmax differs from the exact one
    # in the reference by a few per cent, which moves the polynomial slightly: compare loosely, then
    # exactly through the properties that matter
    zc = P @ chebyshev_reference(Ac, steps)(P.T @ r)
    assert np.abs((z2 - z1) - zc).max() < 0.05 * np.abs(zc).max()
    assert np.abs((z2 - z1)[keep == 0]).max() < 1e-12 * np.abs(z2).max()   # constrained dofs: no coarse correction
    assert abs(dots[0] - r @ z2) < 1e-10 * abs(r @ z2) and abs(dots[1] - z2 @ z2) < 1e-10 * (z2 @ z2)
    assert r @ z2 > 0
    if dim == 2 and ngl == 3:                                      # still symmetric
        r2 = rng.standard_normal(K.shape[0])
        y2, _, _, _, _ = run_emulated(mesh, K, B, r2, grid=1, coarse_steps=steps)
        assert abs(r2 @ z2 - r @ y2) < 1e-11 * abs(r2 @ z2)


def test_oracle_asm_is_the_same_preconditioner_and_the_same_time_step():
    """oracle/timestep.py's pc="asm" is the checker of the GPU tests: it must be the operator the
    emulated kernels apply, and a time step through it must land on the direct-solve fields"""
    mesh = BoxMesh([3, 3, 3], [0.0] * 3, [1.0] * 3, 3)
    args = (3, 3, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes())
    direct = timestep.OracleProblem(*args, rho=0.5, mu=0.01, solver="direct")
    asm = timestep.OracleProblem(*args, rho=0.5, mu=0.01, solver="cg", pc="asm", rtol=1e-12)
    bj = timestep.OracleProblem(*args, rho=0.5, mu=0.01, solver="cg", pc="bjacobi", rtol=1e-12)
    r = np.random.default_rng(6).standard_normal(asm.K.shape[0])
    z, _, _, _, _ = run_emulated(mesh, asm.K, 3, r, grid=2)
    zo = asm._M.matvec(r)
    assert np.abs(z - zo).max() < 1e-11 * np.abs(zo).max()
    # the two-level statement (pc="asm+coarse" = -pc_asm_coarse 5): same operator as the emulated kernels up
    # to the power iteration's estimate of lambda_max (a few per cent of the coarse part)
    two = timestep.OracleProblem(*args, rho=0.5, mu=0.01, solver="cg", pc="asm+coarse", rtol=1e-12)
    z2 = run_emulated(mesh, asm.K, 3, r, grid=1, coarse_steps=5)[0]
    zo2 = two._M.matvec(r)
    assert np.abs((z2 - z) - (zo2 - zo)).max() < 0.05 * np.abs(zo2 - zo).max()
    for p in (direct, asm, bj):
        p.step(0.0, 2e-3)
    for name in ("vort", "vel"):
        want = getattr(direct, name)
        assert np.abs(getattr(asm, name) - want).max() < 1e-9 * np.abs(want).max()
    assert asm.iterations < 0.75 * bj.iterations, (asm.iterations, bj.iterations)


# ---- row-partitioned operator: one rank's share, ghost rows absent ----------------------------------
# (blocks at least 3 elements thick: the first layer of an upper block is incomplete and needs an
# interior twin behind it; thinner blocks are refused by the set-up, second test)
@pytest.mark.parametrize("dim,nelem,ngl,size", [(3, [4, 4, 6], 3, 2), (3, [6, 6, 6], 3, 8), (2, [8, 6], 4, 4)])
def test_emulated_partitioned_apply_matches_the_global_one(dim, nelem, ngl, size):
    """every rank: its elements (block + halo layer), the owned rows of K only (local numbering), r with
    refreshed ghost entries; the blocks of elements with ghost nodes come from complete elements with the
    same signature.  The owned entries of z must equal the single-GPU preconditioner's."""
    from pynama_b200.parallel import Partition
    from tests.util import fs_classes
    mesh, K = kle_operator(dim, nelem, ngl)
    B = dim
    n = K.shape[0]
    r = np.random.default_rng(31).standard_normal(n)
    zref = element_schwarz(K, mesh.elem_nodes, B, weighted=True)[0](r)
    valence = np.bincount(mesh.elem_nodes.ravel(), minlength=mesh.nnodes).astype(float)
    cls = fs_classes(mesh).reshape(mesh.nnodes, B)
    pack = np.zeros(mesh.nnodes, np.uint8)
    for i in range(B):
        pack |= (cls[:, i].astype(np.uint8) & 3) << (2 * i)
    K = K.tocsr()
    total_rz = 0.0
    ranks = range(size) if size <= 4 else (0, 3, size - 1)       # (8 ranks: a corner, a mixed and the last block)
    for rank in ranks:
        part = Partition(mesh, rank, size)
        l2g = part.l2g
        gd = (l2g[:, None] * B + np.arange(B)[None, :]).ravel()            # local dof -> global dof
        Kloc = K[gd][:, gd].tolil()
        Kloc[part.n_own * B:, :] = 0.0                                      # ghost rows are not assembled here
        Kloc = Kloc.tocsr()
        lm = _LocalMesh(mesh, part)
        z, dots, nclass, _, _ = run_emulated(lm, Kloc, B, r[gd], grid=2, n_own=part.n_own,
                                             valence_global=valence[l2g], nodecls=pack[l2g])
        own = gd[:part.n_own * B]
        # (the oracle's scipy sums leave last-bit differences between translated blocks, so a borrowed
        # inverse is the twin's, not the element's own: equal to rounding times the block's condition)
        assert np.abs(z[:part.n_own * B] - zref[own]).max() < 1e-9 * np.abs(zref).max(), rank
        total_rz += dots[0]
    if len(ranks) == size:
        assert abs(total_rz - r @ zref) < 1e-9 * abs(r @ zref)              # what the allreduce sums


def test_emulated_partitioned_setup_refuses_blocks_without_a_twin():
    from pynama_b200.parallel import Partition
    from tests.util import fs_classes
    mesh, K = kle_operator(3, [4, 4, 4], 3)                    # z-slabs two elements thick
    part = Partition(mesh, 1, 2)
    gd = (part.l2g[:, None] * 3 + np.arange(3)[None, :]).ravel()
    Kloc = K.tocsr()[gd][:, gd].tolil()
    Kloc[part.n_own * 3:, :] = 0.0
    valence = np.bincount(mesh.elem_nodes.ravel(), minlength=mesh.nnodes).astype(float)
    with pytest.raises(AssertionError, match="-10"):
        run_emulated(_LocalMesh(mesh, part), Kloc.tocsr(), 3, np.ones(len(gd)), grid=1, n_own=part.n_own,
                     valence_global=valence[part.l2g], nodecls=np.zeros(part.n_loc, np.uint8))


def test_emulated_rounded_grouping_merges_blocks_equal_up_to_rounding():
    """the set-up's second attempt (when bitwise grouping leaves too many classes): the oracle's scipy
    sums leave last-bit differences between translated blocks, which values rounded to 1e-7 of max |K|
    ignore -- a uniform box then has its 3^dim position classes and the preconditioner moves by
    rounding only"""
    mesh, K = kle_operator(2, [6, 5], 4)
    r = np.random.default_rng(41).standard_normal(K.shape[0])
    z0, _, n0, _, _ = run_emulated(mesh, K, 2, r, grid=2)
    z1, _, n1, cls, _ = run_emulated(mesh, K, 2, r, grid=2, quantum_rel=1e-7)
    assert n1 == 9 and n1 <= n0
    pos = np.array([[min(i, 1) + (i == 5), min(j, 1) + (j == 4)] for i, j in mesh.cell_ijk[:, :2]])  # first/mid/last
    for c in range(n1):
        assert len({tuple(p) for p in pos[cls == c]}) == 1              # one position class per hash class
    assert np.abs(z1 - z0).max() < 1e-9 * np.abs(z0).max()


def test_emulated_apply_on_an_unsymmetric_operator():
    """GMRES / BiCGStab may hand the preconditioner an unsymmetric operator: the inverses are then
    not symmetrised and are stored transposed (the product kernel reads row k for column k)"""
    mesh, K = kle_operator(2, [4, 3], 4)
    rng = np.random.default_rng(51)
    K2 = (K @ sp.diags(rng.uniform(0.5, 2.0, K.shape[0]))).tocsr()
    r = rng.standard_normal(K.shape[0])
    z, _, _, _, inv = run_emulated(mesh, K2, 2, r, grid=2, want_inv=True)
    zref = element_schwarz(K2, mesh.elem_nodes, 2, weighted=True)[0](r)
    assert np.abs(z - zref).max() < 1e-10 * np.abs(zref).max()
    assert not np.allclose(inv[0], inv[0].T)
