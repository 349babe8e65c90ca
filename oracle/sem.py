"""Oracle: reference-element tables and per-element matrices (numpy, fp64).

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Restates

* 1-D quadrature rules            src/domain/elements/utilities.py:43-92
* 1-D Lagrange basis/derivative   src/domain/elements/element.py:13-44
* "spectral order" of the nodes   src/domain/elements/spectral.py:230-441
* table set-up                    src/domain/elements/spectral.py:42-90
* K_e, Rw_e, Rd_e                 src/domain/elements/spectral.py:92-160
* SrT_e, DivSrT_e, Curl_e, W_e    src/domain/elements/spectral.py:162-228

The reference loops over quadrature points in Python; here every loop over
points / nodes is an einsum, the math is the same.  Pinned by
tests/golden/elem_*.npz (generated from the reference's Spectral class).
"""
import numpy as np

ALPHA_W = 1e2      # spectral.py:96
ALPHA_D = 1e3      # spectral.py:97


# ----------------------------------------------------------------------------
# 1-D rules
# ----------------------------------------------------------------------------
def gauss_legendre(n):
    """Golub-Welsch nodes/weights, symmetrised (utilities.py:43-61)."""
    k = np.arange(1, n)
    off = 0.5 / np.sqrt(1.0 - (2.0 * k) ** -2.0)
    jac = np.diag(off, 1) + np.diag(off, -1)
    lam, vec = np.linalg.eig(jac)
    order = np.argsort(lam)
    x = lam[order]
    w = 2.0 * vec[0, order] ** 2
    return (x - x[::-1]) / 2.0, (w + w[::-1]) / 2.0


def gauss_lobatto(n):
    """Gauss-Lobatto-Legendre nodes/weights by Newton iteration from the
    Chebyshev guess, symmetrised (utilities.py:63-92)."""
    x = np.cos(np.linspace(0.0, np.pi, n))
    leg = np.zeros((n, n))
    prev = 2.0
    while np.max(np.abs(x - prev)) > 1e-15:
        prev = x
        leg[:, 0] = 1.0
        leg[:, 1] = x
        for k in range(2, n):
            leg[:, k] = ((2 * k - 1) * x * leg[:, k - 1] - (k - 1) * leg[:, k - 2]) / k
        x = prev - (x * leg[:, n - 1] - leg[:, n - 2]) / (n * leg[:, n - 1])
    w = 2.0 / ((n - 1) * n * leg[:, n - 1] ** 2)
    return (x[::-1] - x) / 2.0, (w[::-1] + w) / 2.0


def lagrange_1d(nodes, pts):
    """h[p, j] = l_j(pts[p]),  dh[p, j] = l_j'(pts[p])  (element.py:13-44).

    Same product formulas as the reference (so that a point sitting on a node
    gives exact 0/1), written per point with boolean masks instead of the
    reference's filled matrices.
    """
    nodes = np.asarray(nodes, dtype=np.float64)
    pts = np.asarray(pts, dtype=np.float64)
    n = len(nodes)
    h = np.zeros((len(pts), n))
    dh = np.zeros((len(pts), n))
    offdiag = ~np.eye(n, dtype=bool)
    # den[j] = prod_{k != j} (x_j - x_k), accumulated in k order like np.prod
    diffn = np.where(offdiag, -nodes[None, :] + nodes[:, None], 1.0)
    den = np.prod(diffn, axis=1)
    for p, xp in enumerate(pts):
        d = np.where(offdiag, xp - nodes[None, :] * np.ones((n, 1)), 1.0)
        h[p] = np.prod(d, axis=1) / den
        acc = np.zeros(n)
        for m in range(n):
            dm = d.copy()
            dm[:, m] = offdiag[:, m]          # drop factor m; row m contributes 0
            acc = acc + np.prod(dm, axis=1)
        dh[p] = acc / den
    return h, dh


# ----------------------------------------------------------------------------
# Local ("spectral") ordering of an n^dim tensor grid: vertices, edge
# interiors, face interiors, body.  Returned as tensor indices (a1, a2[, a3])
# per local position, a_d indexing the ascending 1-D points of reference
# direction d (r, s, t).
# ----------------------------------------------------------------------------
def _up(n):
    return list(range(1, n - 1))


def _dn(n):
    return list(range(n - 2, 0, -1))


def local_order_2d(n):
    """spectral.py:230-252 (the `Ind` table and its inverse permutation)."""
    if n == 1:
        return np.zeros((1, 2), dtype=np.int64)
    hi = n - 1
    seq = [(hi, hi), (0, hi), (0, 0), (hi, 0)]
    seq += [(t, hi) for t in _dn(n)]
    seq += [(0, t) for t in _dn(n)]
    seq += [(t, 0) for t in _up(n)]
    seq += [(hi, t) for t in _up(n)]
    seq += [(a, b) for a in _up(n) for b in _dn(n)]
    return np.array(seq, dtype=np.int64)


def local_order_3d(n):
    """spectral.py:356-441 (`getSpectralOrder`): 8 vertices, 12 edges, 6 faces,
    body.  Directions of travel along every edge / face were read off the
    reference's index table and are checked against it for ngl = 2..7 by
    tests/golden/make_golden.py."""
    if n == 1:
        return np.zeros((1, 3), dtype=np.int64)
    hi = n - 1
    up, dn = _up(n), _dn(n)
    seq = [(0, 0, 0), (0, hi, 0), (hi, hi, 0), (hi, 0, 0),
           (0, 0, hi), (hi, 0, hi), (hi, hi, hi), (0, hi, hi)]
    # edges
    seq += [(0, t, 0) for t in up]
    seq += [(t, hi, 0) for t in up]
    seq += [(hi, t, 0) for t in dn]
    seq += [(t, 0, 0) for t in dn]
    seq += [(t, 0, hi) for t in up]
    seq += [(hi, t, hi) for t in up]
    seq += [(t, hi, hi) for t in dn]
    seq += [(0, t, hi) for t in dn]
    seq += [(hi, 0, t) for t in up]
    seq += [(0, 0, t) for t in dn]
    seq += [(0, hi, t) for t in up]
    seq += [(hi, hi, t) for t in dn]
    # faces: slow index first, fast index second
    seq += [(x, y, 0) for y in dn for x in up]
    seq += [(x, y, hi) for y in up for x in dn]
    seq += [(x, 0, z) for x in dn for z in up]
    seq += [(x, hi, z) for x in up for z in dn]
    seq += [(hi, y, z) for z in dn for y in dn]
    seq += [(0, y, z) for z in up for y in up]
    # body
    seq += [(x, y, z) for z in dn for y in dn for x in up]
    return np.array(seq, dtype=np.int64)


def local_order(n, dim):
    return local_order_2d(n) if dim == 2 else local_order_3d(n)


# ----------------------------------------------------------------------------
# Table sets
# ----------------------------------------------------------------------------
class Rule:
    """One quadrature rule of the element: basis values/derivatives of the
    element basis (H, Hrs) and of the corner (multi-linear geometry) basis
    (Hc, Hrsc) at the rule's points, all in local order.

      H    [g, a]        Hrs  [g, d, a]
      Hc   [g, c]        Hrsc [g, d, c]       w [g]     pts [g, dim]
    """

    def __init__(self, nodes1d, pts1d, w1d, dim):
        cn1d = np.array([-1.0, 1.0])
        h, dh = lagrange_1d(nodes1d, pts1d)
        hc, dhc = lagrange_1d(cn1d, pts1d)
        n, npts = len(nodes1d), len(pts1d)
        node_t = local_order(n, dim)            # [a, dim]
        corn_t = local_order(2, dim)            # [c, dim]
        gp_t = local_order(npts, dim)           # [g, dim]
        self.h1d, self.dh1d, self.hc1d, self.dhc1d = h, dh, hc, dhc
        self.pts1d, self.w1d = np.asarray(pts1d), np.asarray(w1d)
        self.gp_t, self.node_t, self.corn_t = gp_t, node_t, corn_t
        self.H, self.Hrs = self._tensor(h, dh, gp_t, node_t, dim)
        self.Hc, self.Hrsc = self._tensor(hc, dhc, gp_t, corn_t, dim)
        # weight = w[c1]*w[c2](*w[c3]) multiplied left to right (utilities.py:15-41)
        w = self.w1d[gp_t[:, 0]].copy()
        for d in range(1, dim):
            w = w * self.w1d[gp_t[:, d]]
        self.w = w
        self.pts = self.pts1d[gp_t]

    @staticmethod
    def _tensor(h, dh, gp_t, nd_t, dim):
        """Tensor-product basis at tensor points; factors multiplied in
        direction order like the reference's reduce(mul, ...)
        (spectral.py:283-303, 322-345)."""
        ng, na = len(gp_t), len(nd_t)
        fac = [h[gp_t[:, d]][:, nd_t[:, d]] for d in range(dim)]      # [g, a]
        dfac = [dh[gp_t[:, d]][:, nd_t[:, d]] for d in range(dim)]
        H = fac[0].copy()
        for d in range(1, dim):
            H = H * fac[d]
        Hrs = np.zeros((ng, dim, na))
        for dd in range(dim):
            t = (dfac[0] if dd == 0 else fac[0]).copy()
            for d in range(1, dim):
                t = t * (dfac[d] if d == dd else fac[d])
            Hrs[:, dd, :] = t
        return H, Hrs


class SpectralTables:
    """The six table sets of spectral.py:42-90: {element basis, corner basis}
    x {full, reduced, nodal} rules, plus the curl/div index tables of
    spectral.py:26-33."""

    def __init__(self, ngl, dim):
        assert dim in (2, 3) and ngl >= 2
        self.ngl, self.dim = ngl, dim
        self.dim_w = 1 if dim == 2 else 3
        self.dim_s = 3 if dim == 2 else 6
        self.nn = ngl ** dim
        nodes1d, w_nodal = gauss_lobatto(ngl)
        if ngl <= 3:                                       # spectral.py:44-45,69-70
            full1d, w_full = gauss_legendre(ngl)
        else:
            full1d, w_full = gauss_lobatto(ngl)
        red1d, w_red = gauss_legendre(ngl - 1)             # spectral.py:46,71
        self.nodes1d = nodes1d
        self.full = Rule(nodes1d, full1d, w_full, dim)
        self.red = Rule(nodes1d, red1d, w_red, dim)
        self.nodal = Rule(nodes1d, nodes1d, w_nodal, dim)
        self.node_t = self.full.node_t
        if dim == 2:
            # (row, column offset, derivative) with sign (-1)^i
            self.ind_wcurl = [(0, 0, 1), (1, 0, 0)]
            self.ind_curl = [(0, 1, 0), (0, 0, 1)]
            self.ind_bdiv = [(0, 1), (1, 2)]
        else:
            self.ind_wcurl = [(0, 2, 1), (0, 1, 2), (1, 0, 2), (1, 2, 0), (2, 1, 0), (2, 0, 1)]
            self.ind_curl = list(self.ind_wcurl)
            self.ind_bdiv = [(0, 1, 5), (1, 2, 3), (5, 3, 4)]

    # -- coefficient tensors: operator row r, component c  <-  d/dx_k --------
    def curl_coef(self, table):
        """C[r, c, k]: row r of the curl gets sign * d/dx_k of component c."""
        nrow = self.dim_w if table is self.ind_curl else self.dim
        ncol = self.dim if table is self.ind_curl else self.dim_w
        C = np.zeros((nrow, ncol, self.dim))
        for i, (r, c, k) in enumerate(table):
            C[r, c, k] = (-1.0) ** i
        return C

    def srt_coef(self):
        """S[s, c, k] of B_srt after the 0.5 scaling (spectral.py:199-217)."""
        dim = self.dim
        S = np.zeros((self.dim_s, dim, dim))
        for x in range(dim):
            S[2 * x, x, x] = 2.0
            for i in range(dim):
                if i != x:
                    S[(x + i) if (x + i) != 2 else 5, i, x] = 1.0
        return 0.5 * S

    def divs_coef(self):
        """D[i, s, k]: row i of div(T) gets d/dx_k of packed component s
        (spectral.py:203-204)."""
        dim = self.dim
        D = np.zeros((dim, self.dim_s, dim))
        for x in range(dim):
            for i in range(dim):
                D[i, self.ind_bdiv[x][i], x] = 1.0
        return D


def _geometry(rule, X):
    """J = Hrsc.X, its inverse and determinant, and Hxy = J^-1 Hrs at every
    point of the rule (spectral.py:120-123, 141-145, 187-191)."""
    J = np.einsum('gdc,ce->gde', rule.Hrsc, X)
    Jinv = np.linalg.inv(J)
    det = np.linalg.det(J)
    Hxy = np.einsum('gde,gea->gda', Jinv, rule.Hrs)
    return det, Hxy


def elem_kle(tab, X):
    """(K_e, Rw_e, Rd_e) of spectral.py:92-160 for one element.

    X: corner coordinates [2^dim, dim] in the reference's corner order.
    Row/column dof = node*ncomp + comp (interleaved).
    """
    dim, dim_w, nn = tab.dim, tab.dim_w, tab.nn
    X = np.asarray(X, dtype=np.float64).reshape(2 ** dim, dim)
    K = np.zeros((nn, dim, nn, dim))
    Rw = np.zeros((nn, dim, nn, dim_w))
    Rd = np.zeros((nn, dim, nn))
    Cw = tab.curl_coef(tab.ind_wcurl)      # [dim, dim_w, k]
    Cv = tab.curl_coef(tab.ind_curl)       # [dim_w, dim, k]

    # full rule: gradient term, vorticity-curl coupling, -H^T grad   (:116-136)
    r = tab.full
    det, G = _geometry(r, X)
    wd = r.w * det
    lap = np.einsum('g,gka,gkb->ab', wd, G, G)
    for c in range(dim):
        K[:, c, :, c] += lap
    Rw += np.einsum('g,ga,rck,gkb->arbc', wd, r.H, Cw, G)
    Rd -= np.einsum('g,ga,grb->arb', wd, r.H, G)

    # reduced rule: divergence / curl penalties                     (:138-159)
    r = tab.red
    det, G = _geometry(r, X)
    wd = r.w * det
    Bdiv = G.transpose(0, 2, 1)                                  # [g, a, c] = dN_a/dx_c
    Bcurl = np.einsum('rck,gka->grac', Cv, G)                    # [g, r, a, c]
    K += ALPHA_D * np.einsum('g,gac,gbe->acbe', wd, Bdiv, Bdiv)
    K += ALPHA_W * np.einsum('g,grac,grbe->acbe', wd, Bcurl, Bcurl)
    Rw += ALPHA_W * np.einsum('g,grac,gb->acbr', wd, Bcurl, r.H)
    Rd += ALPHA_D * np.einsum('g,gac,gb->acb', wd, Bdiv, r.H)
    n = nn
    return (K.reshape(n * dim, n * dim), Rw.reshape(n * dim, n * dim_w),
            Rd.reshape(n * dim, n))


def elem_ops(tab, X):
    """(SrT_e, DivSrT_e, Curl_e, W_e) of spectral.py:162-228 (nodal rule)."""
    dim, dim_w, dim_s, nn = tab.dim, tab.dim_w, tab.dim_s, tab.nn
    X = np.asarray(X, dtype=np.float64).reshape(2 ** dim, dim)
    r = tab.nodal
    det, G = _geometry(r, X)
    wd = r.w * det
    S, D, Cv = tab.srt_coef(), tab.divs_coef(), tab.curl_coef(tab.ind_curl)
    SrT = np.einsum('g,ga,sck,gkb->asbc', wd, r.H, S, G)
    Div = np.einsum('g,ga,isk,gkb->aibs', wd, r.H, D, G)
    Curl = np.einsum('g,ga,rck,gkb->arbc', wd, r.H, Cv, G)
    W = np.einsum('g,ga,gb->a', wd, r.H, r.H)
    return (SrT.reshape(nn * dim_s, nn * dim), Div.reshape(nn * dim, nn * dim_s),
            Curl.reshape(nn * dim_w, nn * dim), W)


def node_coords(tab, X):
    """Physical position of every element node in local order: corner basis at
    the nodal points times the corner coordinates (dmplex.py:83-85)."""
    X = np.asarray(X, dtype=np.float64).reshape(2 ** tab.dim, tab.dim)
    return tab.nodal.Hc @ X
