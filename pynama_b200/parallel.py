"""Row partition of a spectral-element mesh over the GPUs of one box.

Replaces what PETSc does for the reference when it runs under ``mpiexec``
(``DMPlex.distribute()``, src/domain/dmplex.py:21; AIJ local-row ownership,
src/matrices/mat_fs.py:102-106; VecScatter ghost updates inside ``MatMult``):

* the box of elements is cut into ``px * py * pz`` contiguous blocks (z-slabs for
  2 ranks, 2x2x1 for 4, 2x2x2 for 8: smallest halo surface), rank = bx + px*(by + py*bz);
  an unstructured (Gmsh) mesh is cut by recursive coordinate bisection of the element centroids;
* a node is owned by the lowest rank whose block touches it;
* every rank holds **all elements that touch one of its owned nodes** (its block plus one
  layer of elements on the upper sides): owned rows are then assembled completely and
  locally -- owner-computes-rows with redundant halo elements, no assembly exchange;
* local numbering: owned nodes first (ascending global id), then ghost nodes grouped by owner
  rank (ascending global id inside a group), so every neighbour's ghost values are ONE
  contiguous run of the local vector and a halo receive needs no unpacking;
* global numbering is the 1-GPU numbering: results are directly comparable.

Everything here is replicated host logic (numpy): any rank can compute any other rank's
lists, so the send lists need no communication to set up.
"""
import numpy as np


def block_grid(size, nelem):
    """(px, py[, pz]) with px*py*pz == size, cutting the longest axes last-to-first:
    2 -> z-slabs, 4 -> 2x2 in (y, z) ... chosen to minimise the halo surface."""
    dim = len(nelem)
    p = [1] * dim
    n = int(size)
    f = 2
    factors = []
    while n > 1:
        while n % f == 0:
            factors.append(f)
            n //= f
        f += 1
    for f in sorted(factors, reverse=True):
        # give the factor to the axis with the most elements per block (ties: last axis first)
        per = [nelem[d] / p[d] for d in range(dim)]
        d = max(range(dim), key=lambda a: (per[a], a))
        p[d] *= f
    for d in range(dim):
        if p[d] > nelem[d]:
            raise ValueError(f"cannot cut {nelem[d]} elements along axis {d} into {p[d]} blocks")
    return tuple(p)


def _cuts(n, parts):
    """element index ranges [c[b], c[b+1]) of `parts` nearly equal blocks"""
    base, rem = divmod(n, parts)
    sizes = [base + (1 if b < rem else 0) for b in range(parts)]
    return np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)


def rcb(centroids, size):
    """recursive coordinate bisection of element centroids into `size` parts (unstructured meshes,
    SURVEY.md 8e-E1): cut the longest extent at the weighted median, recurse.  Returns the part of
    every element; parts are numbered left to right along the cuts."""
    c = np.asarray(centroids, dtype=np.float64)
    part = np.zeros(len(c), dtype=np.int64)

    def split(idx, first, nparts):
        if nparts == 1:
            part[idx] = first
            return
        nl = nparts // 2
        ext = np.ptp(c[idx], axis=0)
        axis = int(np.argmax(ext))
        order = idx[np.lexsort((idx, c[idx, axis]))]          # ties broken by element id: deterministic
        k = (len(order) * nl) // nparts
        split(order[:k], first, nl)
        split(order[k:], first + nl, nparts - nl)

    split(np.arange(len(c)), 0, int(size))
    return part


class Partition:
    """Local tables of one rank.  `mesh` is the global mesh (BoxMesh or GmshMesh, replicated)."""

    def __init__(self, mesh, rank, size, grid=None):
        self.rank, self.size = int(rank), int(size)
        self.dim = dim = mesh.dim
        if hasattr(mesh, "nelem"):                       # structured box: contiguous element blocks
            self.pgrid = tuple(grid) if grid is not None else block_grid(size, mesh.nelem)
            assert int(np.prod(self.pgrid)) == size
            self.cuts = [_cuts(mesh.nelem[d], self.pgrid[d]) for d in range(dim)]
            elem_rank = np.zeros(mesh.nel, dtype=np.int64)
            stride = 1
            for d in range(dim):
                blk = np.searchsorted(self.cuts[d][1:], mesh.cell_ijk[:, d], side="right")
                elem_rank += blk * stride
                stride *= self.pgrid[d]
        else:                                            # unstructured: recursive coordinate bisection
            self.pgrid = (size,)
            elem_rank = rcb(mesh.corner_xyz.mean(axis=1), size)
        self.elem_rank = elem_rank
        # a node belongs to the lowest rank whose elements touch it
        owner = np.full(mesh.nnodes, size, dtype=np.int64)
        np.minimum.at(owner, mesh.elem_nodes.ravel(), np.repeat(elem_rank, mesh.elem_nodes.shape[1]))
        self.node_owner = owner
        owned = np.nonzero(owner == self.rank)[0]                           # ascending global id
        # elements touching an owned node
        touch = (owner[mesh.elem_nodes] == self.rank).any(axis=1)
        self.elems = np.nonzero(touch)[0]                                   # ascending global id
        # owned nodes are numbered interior first, then the ones that share an element with a ghost node
        # (the rows of a product that read ghost columns), each group by ascending global id: the
        # device multiplies the interior rows while the ghost values are still on their way
        en_loc = mesh.elem_nodes[self.elems]
        with_ghost = (owner[en_loc] != self.rank).any(axis=1)
        bnd = np.zeros(mesh.nnodes, dtype=bool)
        bnd[np.unique(en_loc[with_ghost])] = True
        is_bnd = bnd[owned]
        self.owned = np.concatenate([owned[~is_bnd], owned[is_bnd]])
        self.n_interior = int((~is_bnd).sum())
        loc_nodes = np.unique(mesh.elem_nodes[self.elems])
        ghost = loc_nodes[owner[loc_nodes] != self.rank]
        order = np.lexsort((ghost, owner[ghost]))                           # by owner, then global id
        self.ghost = ghost[order]
        self.ghost_owner = owner[self.ghost]
        self.n_own, self.n_ghost = len(self.owned), len(self.ghost)
        self.l2g = np.concatenate([self.owned, self.ghost]).astype(np.int64)
        self.n_loc = len(self.l2g)
        g2l = np.full(mesh.nnodes, -1, dtype=np.int64)
        g2l[self.l2g] = np.arange(self.n_loc)
        self.g2l = g2l
        # local tables for the device mesh
        self.elem_nodes = np.ascontiguousarray(g2l[mesh.elem_nodes[self.elems]], dtype=np.int32)
        self.corner_xyz = np.ascontiguousarray(mesh.corner_xyz[self.elems])
        self.coords = mesh.coords[self.l2g]
        # receives: one contiguous run of the ghost section per owner rank
        self.neighbors_recv, self.recv_ptr = [], [0]
        for q in np.unique(self.ghost_owner):
            self.neighbors_recv.append(int(q))
            self.recv_ptr.append(int(np.searchsorted(self.ghost_owner, q, side="right")))
        self._mesh = mesh
        self._sends = None

    # -- sends need the other ranks' ghost lists (replicated computation, no communication)
    def _ghosts_of(self, q):
        mesh = self._mesh
        touch = (self.node_owner[mesh.elem_nodes] == q).any(axis=1)
        nodes = np.unique(mesh.elem_nodes[touch])
        return nodes[self.node_owner[nodes] != q]

    def sends(self):
        """[(rank q, local ids of my owned nodes that q holds as ghosts, ascending global id)]"""
        if self._sends is None:
            out = []
            for q in range(self.size):
                if q == self.rank:
                    continue
                g = self._ghosts_of(q)
                mine = g[self.node_owner[g] == self.rank]               # np.unique output: ascending
                if len(mine):
                    out.append((q, self.g2l[mine].astype(np.int32)))
            self._sends = out
        return self._sends

    def halo_tables(self):
        """flat tables for pyn_halo_create: neighbour ranks (union of send and receive
        partners, ascending), send_ptr/send_nodes, recv_ptr (node offsets into the ghost section)"""
        sends = dict(self.sends())
        recvs = {q: (self.recv_ptr[i], self.recv_ptr[i + 1]) for i, q in enumerate(self.neighbors_recv)}
        ranks = sorted(set(sends) | set(recvs))
        send_ptr, send_nodes, recv_off, recv_cnt = [0], [], [], []
        for q in ranks:
            s = sends.get(q, np.zeros(0, np.int32))
            send_nodes.append(s)
            send_ptr.append(send_ptr[-1] + len(s))
            a, b = recvs.get(q, (0, 0))
            recv_off.append(a)
            recv_cnt.append(b - a)
        return (np.asarray(ranks, dtype=np.int32), np.asarray(send_ptr, dtype=np.int64),
                np.concatenate(send_nodes).astype(np.int32) if send_nodes else np.zeros(0, np.int32),
                np.asarray(recv_off, dtype=np.int64), np.asarray(recv_cnt, dtype=np.int64))

    # -- index helpers -----------------------------------------------------------------------
    def local_dofs(self, global_dofs, ncomp):
        """global dof ids (node*ncomp+c) -> (positions kept, local dof ids); dofs of nodes this
        rank does not hold are dropped"""
        gd = np.asarray(global_dofs, dtype=np.int64).ravel()
        ln = self.g2l[gd // ncomp]
        keep = np.nonzero(ln >= 0)[0]
        return keep, ln[keep] * ncomp + gd[keep] % ncomp

    def owned_dofs_global(self, ncomp):
        return (self.owned[:, None] * ncomp + np.arange(ncomp)[None, :]).ravel()
