"""CPU: the row partition (pynama_b200/parallel.py) -- ownership, halo tables, and that
owner-computes-rows with redundant halo elements reproduces the global matrices row for row
(checked with the oracle assembly standing in for the device kernels)."""
import numpy as np
import pytest

from oracle import assembly, sem
from pynama_b200.domain.box_mesh import BoxMesh
from pynama_b200.parallel import Partition, block_grid
from tests.util import fs_classes


def test_block_grid():
    assert block_grid(1, [16, 16, 16]) == (1, 1, 1)
    assert block_grid(2, [16, 16, 16]) == (1, 1, 2)          # z-slabs
    assert sorted(block_grid(4, [16, 16, 16])) == [1, 2, 2]
    assert block_grid(8, [16, 16, 16]) == (2, 2, 2)
    assert block_grid(2, [8, 4]) == (2, 1)
    assert int(np.prod(block_grid(6, [6, 6, 6]))) == 6
    with pytest.raises(ValueError):
        block_grid(8, [1, 1, 2])


def _gmsh_twin(tmp_path, nelem, ngl):
    from pynama_b200.domain.gmsh_mesh import GmshMesh
    from tests.test_gmsh_mesh import box_as_msh
    path = str(tmp_path / "m.msh")
    box_as_msh(path, nelem)
    return GmshMesh(path, ngl)


def test_block_owner_rule_is_the_lowest_block_touching_the_node():
    """structured meshes: ownership from element ranks == the per-axis rule of the first version"""
    mesh = BoxMesh([5, 4, 3], [0] * 3, [1] * 3, 3)
    p = Partition(mesh, 0, 8)
    q = mesh.ngl - 1
    want = np.zeros(mesh.nnodes, dtype=np.int64)
    stride = 1
    for d in range(3):
        upper = p.cuts[d][1:] * q
        want += np.searchsorted(upper, mesh.node_ijk[:, d], side="left") * stride
        stride *= p.pgrid[d]
    assert np.array_equal(p.node_owner, want)


def test_rcb_balances_and_is_deterministic():
    from pynama_b200.parallel import rcb
    rng = np.random.default_rng(0)
    c = rng.uniform(0, 1, (1001, 3)) * [4.0, 1.0, 2.0]
    for size in (2, 3, 4, 8):
        part = rcb(c, size)
        counts = np.bincount(part, minlength=size)
        assert counts.max() - counts.min() <= 2, counts
        assert np.array_equal(part, rcb(c, size))
    assert np.all(c[rcb(c, 2) == 0][:, 0].max() <= c[rcb(c, 2) == 1][:, 0].min())      # first cut: longest axis


@pytest.mark.parametrize("nelem,ngl,size", [([4, 4], 3, 2), ([4, 3], 4, 4), ([3, 3, 4], 3, 2),
                                            ([4, 4, 4], 3, 8), ([3, 4, 5], 2, 3),
                                            ("gmsh", [5, 4], 3, 2), ("gmsh", [3, 3, 4], 3, 4)][:5] +
                         [(("gmsh", [5, 4]), 3, 2), (("gmsh", [3, 3, 4]), 3, 4), (("gmsh", [4, 3, 3]), 4, 8)])
def test_ownership_and_halo_tables(tmp_path, nelem, ngl, size):
    if isinstance(nelem, tuple):
        mesh = _gmsh_twin(tmp_path, nelem[1], ngl)
    else:
        mesh = BoxMesh(nelem, [0] * len(nelem), [1] * len(nelem), ngl)
    parts = [Partition(mesh, r, size) for r in range(size)]
    # every node owned exactly once; owned sets partition the nodes
    allown = np.concatenate([p.owned for p in parts])
    assert len(allown) == mesh.nnodes and len(np.unique(allown)) == mesh.nnodes
    # every element is held by at least its block owner; local tables are consistent
    held = np.zeros(mesh.nel, int)
    for p in parts:
        held[p.elems] += 1
        assert np.array_equal(p.l2g[p.elem_nodes], mesh.elem_nodes[p.elems])
        # owned nodes: interior first, then the ones that share an element with a ghost, each ascending
        ni = p.n_interior
        assert np.all(np.diff(p.owned[:ni]) > 0) and np.all(np.diff(p.owned[ni:]) > 0) and np.all(np.diff(p.elems) > 0)
        has_ghost = (p.elem_nodes >= p.n_own).any(axis=1)
        touched = np.zeros(p.n_loc, bool)
        touched[np.unique(p.elem_nodes[has_ghost])] = True
        assert not touched[:ni].any() and touched[ni:p.n_own].all()           # rows [0, ni) read no ghost column
        assert np.all(p.g2l[p.l2g] == np.arange(p.n_loc))
        # all nodes coupled to an owned node are local (columns of owned rows)
        touch = (p.node_owner[mesh.elem_nodes] == p.rank).any(axis=1)
        assert np.array_equal(np.nonzero(touch)[0], p.elems)
    assert held.min() >= 1
    # send list r -> q  ==  q's ghost run owned by r, same order
    for p in parts:
        ranks, send_ptr, send_nodes, recv_off, recv_cnt = p.halo_tables()
        for i, q in enumerate(ranks):
            sent_global = p.l2g[send_nodes[send_ptr[i]:send_ptr[i + 1]]]
            pq = parts[q]
            rq, _, _, roff, rcnt = pq.halo_tables()
            j = list(rq).index(p.rank)
            recv_global = pq.ghost[roff[j]:roff[j] + rcnt[j]]
            assert np.array_equal(sent_global, recv_global)
            assert np.all(p.node_owner[sent_global] == p.rank)


@pytest.mark.parametrize("nelem,ngl,size", [([4, 4], 3, 2), ([3, 3, 4], 3, 4), ([4, 4, 4], 2, 8)])
def test_owner_computes_rows_matches_global(nelem, ngl, size):
    dim = len(nelem)
    mesh = BoxMesh(nelem, [0] * dim, [1] * dim, ngl, perturb=0.1)
    tab = sem.SpectralTables(ngl, dim)
    cls = fs_classes(mesh)
    G = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    Gops = assembly.assemble_ops(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
    rng = np.random.default_rng(1)
    dw, ds = (1, 3) if dim == 2 else (3, 6)
    xv, xw, xs = rng.uniform(-1, 1, mesh.nnodes * dim), rng.uniform(-1, 1, mesh.nnodes * dw), \
        rng.uniform(-1, 1, mesh.nnodes * ds)
    yK, yRw = G["K"] @ xv, G["Rw"] @ xw
    yC, yS, yD = Gops["Curl"] @ xv, Gops["SrT"] @ xv, Gops["DivSrT"] @ xs
    for r in range(size):
        p = Partition(mesh, r, size)
        ldofs = lambda nc: (p.l2g[:, None] * nc + np.arange(nc)[None, :]).ravel()
        lcls = cls[ldofs(dim)]
        L = assembly.assemble_kle(tab, p.corner_xyz, p.elem_nodes, p.n_loc, lcls, "FS")
        Lops = assembly.assemble_ops(tab, p.corner_xyz, p.elem_nodes, p.n_loc)
        own = lambda nc: slice(0, p.n_own * nc)
        gown = lambda nc: p.owned_dofs_global(nc)
        # "halo exchange": local x = owned + ghost values of the global vector
        for name, loc, x, y, ncin, ncout in (("K", L["K"], xv, yK, dim, dim), ("Rw", L["Rw"], xw, yRw, dw, dim),
                                             ("Curl", Lops["Curl"], xv, yC, dim, dw),
                                             ("SrT", Lops["SrT"], xv, yS, dim, ds),
                                             ("DivSrT", Lops["DivSrT"], xs, yD, ds, dim)):
            got = (loc @ x[ldofs(ncin)])[own(ncout)]
            want = y[gown(ncout)]
            assert np.abs(got - want).max() <= 1e-12 * max(np.abs(want).max(), 1e-300), (name, r)
        # lumped mass of owned nodes is complete
        assert np.abs(Lops["W"][:p.n_own] - Gops["W"][p.owned]).max() <= 1e-14
