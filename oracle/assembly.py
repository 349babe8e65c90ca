"""Oracle: scatter-add assembly of the KLE matrices and operators (scipy CSR).

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Restates, cell by cell and in
the reference's insertion order,

* MatFS.buildFS          src/matrices/mat_fs.py:150-192
* MatNS.buildNS          src/matrices/mat_ns.py:47-145
* Operators.setValues    src/matrices/mat_fs.py:238-251
* Operators.assembleAll  src/matrices/mat_fs.py:253-271

Inputs are plain tables (corner coordinates, element->node table, per-dof
classes), i.e. what the reference's Domain hands to the loop.  PETSc AIJ keeps
every inserted entry, zeros included (MAT_IGNORE_ZERO_ENTRIES is never set), so
the pattern of a matrix is the union of the inserted index pairs; scipy's
COO->CSR conversion has the same semantics (duplicates summed, explicit zeros
kept).

dof classes (velocity dofs, node*dim+c): 0 free, 1 Dirichlet (free-slip BC
nodes, mat_fs.py:158-172), 2 no-slip normal, 3 no-slip tangential
(mat_ns.py:59-75).
"""
import numpy as np
import scipy.sparse as sp

from . import sem

FREE, DIR, NORMAL, TANG = 0, 1, 2, 3


class _Coo:
    def __init__(self, shape):
        self.shape = shape
        self.r, self.c, self.v = [], [], []

    def add(self, rows, cols, block):
        """Mat.setValues(rows, cols, block, addv=True)"""
        rows = np.asarray(rows, dtype=np.int64)
        cols = np.asarray(cols, dtype=np.int64)
        if rows.size == 0 or cols.size == 0:
            return
        block = np.asarray(block, dtype=np.float64).reshape(rows.size, cols.size)
        self.r.append(np.repeat(rows, cols.size))
        self.c.append(np.tile(cols, rows.size))
        self.v.append(block.ravel())

    def add_diag(self, idx, value):
        idx = np.asarray(sorted(idx), dtype=np.int64)
        if idx.size:
            self.r.append(idx)
            self.c.append(idx)
            self.v.append(np.full(idx.size, float(value)))

    def tocsr(self):
        if not self.r:
            return sp.csr_matrix(self.shape)
        m = sp.coo_matrix((np.concatenate(self.v), (np.concatenate(self.r), np.concatenate(self.c))),
                          shape=self.shape)
        m = m.tocsr()            # sums duplicates, keeps explicit zeros
        m.sort_indices()
        return m


class ElementCache:
    """The reference recomputes identical matrices for congruent cells
    (SURVEY.md 3.1); the oracle may reuse them (reuse=True) -- translation does
    not change J.  reuse=False is the reference's true cost."""

    def __init__(self, tab, fn, reuse=True):
        self.tab, self.fn, self.reuse, self.store = tab, fn, reuse, {}

    def get(self, X):
        if not self.reuse:
            return self.fn(self.tab, X)
        key = np.round((X - X[0]) * 2.0 ** 40).astype(np.int64).tobytes()
        if key not in self.store:
            self.store[key] = self.fn(self.tab, X)
        return self.store[key]


def assemble_kle(tab, corner_xyz, elem_nodes, nnodes, cls, mode="FS", reuse=True):
    """Returns dict name -> csr for K, Krhs, Rw, Rd (+ Kfs, Krhsfs, Rwfs, Rdfs in NS mode)."""
    dim, dim_w = tab.dim, tab.dim_w
    n, nw = nnodes * dim, nnodes * dim_w
    cls = np.asarray(cls, dtype=np.uint8)
    names = ["K", "Krhs", "Rw", "Rd"] + (["Kfs", "Krhsfs", "Rwfs", "Rdfs"] if mode == "NS" else [])
    shapes = {"K": (n, n), "Krhs": (n, n), "Rw": (n, nw), "Rd": (n, nnodes),
              "Kfs": (n, n), "Krhsfs": (n, n), "Rwfs": (n, nw), "Rdfs": (n, nnodes)}
    M = {k: _Coo(shapes[k]) for k in names}
    cache = ElementCache(tab, sem.elem_kle, reuse)
    comp = np.arange(dim)
    compw = np.arange(dim_w)
    ones, onesfs = set(), set()
    for e in range(len(elem_nodes)):
        nodes = np.asarray(elem_nodes[e], dtype=np.int64)
        locK, locRw, locRd = cache.get(np.asarray(corner_xyz[e], dtype=np.float64))
        gvel = (nodes[:, None] * dim + comp[None, :]).ravel()         # indicesVel
        gw = (nodes[:, None] * dim_w + compw[None, :]).ravel()        # indicesW
        c = cls[gvel]
        free = np.nonzero(c == FREE)[0]
        if mode == "FS":
            bc = np.nonzero(c == DIR)[0]
            if bc.size:                                                # mat_fs.py:174-177
                M["Krhs"].add(gvel[free], gvel[bc], -locK[np.ix_(free, bc)])
            M["K"].add(gvel[free], gvel[free], locK[np.ix_(free, free)])       # :179-180
            M["K"].add_diag(gvel[bc], 0.0)                                     # :182-183
            M["Rw"].add(gvel[free], gw, locRw[free, :])                        # :185-186
            ones.update(gvel[bc].tolist())
        else:
            tang = np.nonzero(c == TANG)[0]
            normal = np.nonzero(c == NORMAL)[0]
            both = np.nonzero((c == TANG) | (c == NORMAL))[0]
            if both.size:                                              # mat_ns.py:77-114
                M["Krhs"].add(gvel[free], gvel[both], -locK[np.ix_(free, both)])
                ones.update(gvel[both].tolist())
                M["Krhs"].add_diag(gvel[both], 0.0)
                M["Kfs"].add(gvel[tang], gvel[free], locK[np.ix_(tang, free)])
                M["Kfs"].add(gvel[free], gvel[tang], locK[np.ix_(free, tang)])
                M["Kfs"].add(gvel[tang], gvel[tang], locK[np.ix_(tang, tang)])
                onesfs.update(gvel[tang].tolist())
                M["Rwfs"].add(gvel[tang], gw, locRw[tang, :])
                M["Rdfs"].add(gvel[tang], nodes, locRd[tang, :])
                M["Krhsfs"].add(gvel[tang], gvel[normal], -locK[np.ix_(tang, normal)])
                M["Krhsfs"].add(gvel[free], gvel[normal], -locK[np.ix_(free, normal)])
                M["Krhsfs"].add_diag(gvel[normal], 0.0)
            M["K"].add(gvel[free], gvel[free], locK[np.ix_(free, free)])       # :116-117
            M["K"].add_diag(gvel[both], 0.0)                                   # :119-120
            M["Rw"].add(gvel[free], gw, locRw[free, :])                        # :122-123
            M["Rd"].add(gvel[free], nodes, locRd[free, :])                     # :125-127
    if mode == "FS":
        # every Dirichlet dof of the mesh gets the unit diagonal, touched by a cell or not
        ones = set(np.nonzero(cls == DIR)[0].tolist())                 # mat_fs.py:188-189
    M["K"].add_diag(ones, 1.0)                                         # setIndices2One, :115-120
    M["Krhs"].add_diag(ones, 1.0)
    out = {k: M[k].tocsr() for k in names}
    if mode == "NS":
        M["Kfs"].add_diag(onesfs, -1.0)                                # mat_ns.py:132-133
        out["Kfs"] = M["Kfs"].tocsr()
        # Krhsfs diagonal SET to 1 on the normal dofs (INSERT_VALUES, mat_ns.py:140-142)
        kr = out["Krhsfs"]
        for i in sorted(ones - onesfs):
            lo, hi = kr.indptr[i], kr.indptr[i + 1]
            pos = lo + np.searchsorted(kr.indices[lo:hi], i)
            assert pos < hi and kr.indices[pos] == i      # inserted as an explicit 0 above
            kr.data[pos] = 1.0
    return out


def assemble_ops(tab, corner_xyz, elem_nodes, nnodes, reuse=True):
    """Curl, SrT, DivSrT row-scaled by the reciprocal lumped mass, and the lumped
    mass per node (mat_fs.py:194-201, 238-271)."""
    dim, dim_w, dim_s = tab.dim, tab.dim_w, tab.dim_s
    Curl = _Coo((nnodes * dim_w, nnodes * dim))
    SrT = _Coo((nnodes * dim_s, nnodes * dim))
    Div = _Coo((nnodes * dim, nnodes * dim_s))
    W = np.zeros(nnodes)
    cache = ElementCache(tab, sem.elem_ops, reuse)
    for e in range(len(elem_nodes)):
        nodes = np.asarray(elem_nodes[e], dtype=np.int64)
        locS, locD, locC, locW = cache.get(np.asarray(corner_xyz[e], dtype=np.float64))
        gvel = (nodes[:, None] * dim + np.arange(dim)[None, :]).ravel()
        gw = (nodes[:, None] * dim_w + np.arange(dim_w)[None, :]).ravel()
        gs = (nodes[:, None] * dim_s + np.arange(dim_s)[None, :]).ravel()
        Curl.add(gw, gvel, locC)
        SrT.add(gs, gvel, locS)
        Div.add(gvel, gs, locD)
        np.add.at(W, nodes, locW)
    winv = 1.0 / W
    out = {}
    for name, coo, br in (("Curl", Curl, dim_w), ("SrT", SrT, dim_s), ("DivSrT", Div, dim)):
        m = coo.tocsr()
        scale = np.repeat(winv, br)
        m.data *= np.repeat(scale, np.diff(m.indptr))     # Mat.diagonalScale(L=w^-1)
        out[name] = m
    out["W"] = W
    return out


def elements_of_nodes(elem_nodes, nodes):
    """ascending list of the elements touching any of `nodes`"""
    en = np.asarray(elem_nodes)
    return np.flatnonzero(np.isin(en, np.asarray(nodes)).any(axis=1))


def assemble_rows(tab, corner_xyz, elem_nodes, nnodes, cls, nodes, mode="FS", ops=True, reuse=True):
    """The rows of the sampled `nodes` of every matrix, for meshes too large to assemble whole on
    the CPU: the cell loops of mat_fs.py:150-201 / mat_ns.py:47-145 restricted to the cells that
    touch a sampled node, in the same ascending cell order -- those rows receive exactly the
    contributions (and the summation order) of the full loop.  Returns (dict name -> csr holding
    only the sampled rows, all other rows empty; 'W' -> lumped mass at the nodes) and the cells used."""
    nodes = np.unique(np.asarray(nodes, dtype=np.int64))
    els = elements_of_nodes(elem_nodes, nodes)
    cx = np.asarray(corner_xyz).reshape(len(elem_nodes), -1, tab.dim)[els]
    en = np.asarray(elem_nodes)[els]
    out = dict(assemble_kle(tab, cx, en, nnodes, cls, mode, reuse=reuse))
    if ops:
        with np.errstate(divide="ignore", invalid="ignore"):      # nodes no sampled cell touches: W = 0
            out.update(assemble_ops(tab, cx, en, nnodes, reuse=reuse))

    def keep_rows(m, br):
        rows = (nodes[:, None] * br + np.arange(br)[None, :]).ravel()
        sel = np.zeros(m.shape[0], bool)
        sel[rows] = True
        m = m.tocsr()
        lens = np.diff(m.indptr)
        mask = np.repeat(sel, lens)
        ip = np.concatenate([[0], np.cumsum(np.where(sel, lens, 0))])
        return sp.csr_matrix((m.data[mask], m.indices[mask], ip), shape=m.shape)
    br = {"Curl": tab.dim_w, "SrT": tab.dim_s}
    res = {k: (m[nodes] if k == "W" else keep_rows(m, br.get(k, tab.dim))) for k, m in out.items()}
    return res, els
