"""CPU: the Gmsh 4.1 reader / unstructured mesh tables against the structured
generator: a box mesh written to a .msh file with shuffled vertex numbering, shuffled cell order,
rotated and mirrored cells must give the same nodes, borders and -- through the oracle -- the same
KLE solution at the same coordinates."""
import os

import numpy as np
import pytest
import scipy.sparse.linalg as spl

from oracle import assembly, sem
from pynama_b200.domain.box_mesh import BoxMesh, SIDE_PLANE
from pynama_b200.domain.gmsh_mesh import GmshMesh, read_msh, write_msh

CLOSURE_TO_GMSH_HEX = [0, 3, 2, 1, 4, 5, 6, 7]


def box_as_msh(path, nelem, perturb=0.1, seed=1, scramble=True, version=4.1):
    dim = len(nelem)
    box = BoxMesh(nelem, [0.0] * dim, [1.0, 1.5, 0.75][:dim], 2, perturb=perturb)      # ngl=2: nodes = vertices
    rng = np.random.default_rng(seed)
    nv = box.nnodes
    perm = rng.permutation(nv) if scramble else np.arange(nv)              # new id of old vertex
    vxyz = np.zeros((nv, dim))
    vxyz[perm] = box.coords
    # cell vertex lists in closure order -> gmsh order
    loc = np.abs(box.local_offset)                                         # [4|8, dim] in {0,1}
    closure = np.zeros(2 ** dim, dtype=int)
    from pynama_b200.domain.box_mesh import CORNERS
    for c, bits in enumerate(CORNERS[dim]):
        closure[c] = np.nonzero((loc == bits).all(axis=1))[0][0]
    cells = box.elem_nodes[:, closure]
    if dim == 3:
        g = np.empty_like(cells)
        g[:, CLOSURE_TO_GMSH_HEX] = cells                                  # closure c is gmsh vertex map[c]
        cells = g
    cells = perm[cells]
    if scramble:
        cells = cells[rng.permutation(len(cells))]
        for e in range(len(cells)):
            r = rng.integers(0, 4)
            if dim == 2:
                cells[e] = np.roll(cells[e], r)
                if rng.random() < 0.3:
                    cells[e] = cells[e][[0, 3, 2, 1]]                      # mirrored (clockwise) cell
            else:
                cells[e] = np.concatenate([np.roll(cells[e][:4], r), np.roll(cells[e][4:], r)])
                if rng.random() < 0.3:
                    cells[e] = cells[e][[4, 5, 6, 7, 0, 1, 2, 3]]          # left-handed cell
    # boundary elements tagged by side name
    boundary, names = [], {}
    for k, side in enumerate(box.sides):
        names[(dim - 1, k + 1)] = side
        axis, hi = SIDE_PLANE[side]
        for e in range(box.nel):
            vs = box.elem_nodes[e][closure]
            on = [v for v in vs if box.node_ijk[v, axis] == (box.grid[axis] - 1 if hi else 0)]
            if len(on) == 2 ** (dim - 1):
                if dim == 3:                                               # order the quad around its perimeter
                    pts = box.coords[on]
                    ctr = pts.mean(axis=0)
                    ax = [a for a in range(3) if a != axis]
                    ang = np.arctan2(pts[:, ax[1]] - ctr[ax[1]], pts[:, ax[0]] - ctr[ax[0]])
                    on = [on[i] for i in np.argsort(ang)]
                boundary.append((k + 1, perm[on]))
    names[(dim, 1000)] = "Internal"
    write_msh(path, vxyz, cells, boundary, names, version=version)
    return box


@pytest.mark.parametrize("nelem,ngl", [([4, 3], 2), ([4, 3], 4), ([3, 2, 2], 3), ([2, 2, 3], 4)])
def test_scrambled_box_gives_the_box_tables(tmp_path, nelem, ngl):
    path = str(tmp_path / "box.msh")
    box2 = box_as_msh(path, nelem)
    dim = len(nelem)
    box = BoxMesh(nelem, [0.0] * dim, [1.0, 1.5, 0.75][:dim], ngl, perturb=0.1)
    assert np.allclose(box.vertex_grid, box2.vertex_grid)
    g = GmshMesh(path, ngl)
    assert (g.dim, g.nel, g.nnodes, g.nn) == (box.dim, box.nel, box.nnodes, box.nn)

    def key(x):
        return np.round(np.asarray(x) * 1e9).astype(np.int64)
    kb = {tuple(k): i for i, k in enumerate(key(box.coords))}
    to_box = np.array([kb[tuple(k)] for k in key(g.coords)])               # gmsh node -> box node
    assert len(set(to_box)) == box.nnodes
    # every element's nodes sit where the corner basis puts them, and Jacobians are positive
    tab = sem.SpectralTables(ngl, dim)
    for e in range(g.nel):
        assert np.abs(g.coords[g.elem_nodes[e]] - sem.node_coords(tab, g.corner_xyz[e])).max() < 1e-12
    for side in box.sides:
        assert set(to_box[g.getBorderNodes(side)]) == set(box.getBorderNodes(side)), side
    assert set(to_box[g.getBordersNodes()]) == set(box.getBordersNodes())
    # numbering: lexicographic in the coordinates, x fastest
    order = np.lexsort(key(g.coords).T)
    assert np.array_equal(order, np.arange(g.nnodes))


def test_same_kle_solution_on_both_meshes(tmp_path):
    """KAT-2-like solve (src/tests/test_solver.py:27-37) on the box and on its scrambled .msh twin"""
    nelem, ngl = [5, 4], 4
    path = str(tmp_path / "box.msh")
    box_as_msh(path, nelem)
    box = BoxMesh(nelem, [0.0, 0.0], [1.0, 1.5], ngl, perturb=0.1)
    g = GmshMesh(path, ngl)
    tab = sem.SpectralTables(ngl, 2)
    sol = {}
    for name, m in (("box", box), ("gmsh", g)):
        cls = np.zeros(m.nnodes * 2, np.uint8)
        b = m.getBordersNodes()
        cls[b * 2] = cls[b * 2 + 1] = 1
        mats = assembly.assemble_kle(tab, m.corner_xyz, m.elem_nodes, m.nnodes, cls, "FS")
        x_, y_ = 2 * np.pi * m.coords[:, 0], 2 * np.pi * m.coords[:, 1]
        vel = np.stack([np.cos(x_) * np.sin(y_), -np.sin(x_) * np.cos(y_)], 1).ravel()
        vort = -4 * np.pi * np.cos(x_) * np.cos(y_)
        vbc = np.where(cls == 1, vel, 0.0)
        sol[name] = (m.coords, spl.spsolve(mats["K"].tocsc(), mats["Rw"] @ vort + mats["Krhs"] @ vbc).reshape(-1, 2))
    kb = {tuple(k): i for i, k in enumerate(np.round(sol["box"][0] * 1e9).astype(np.int64))}
    idx = np.array([kb[tuple(k)] for k in np.round(sol["gmsh"][0] * 1e9).astype(np.int64)])
    assert np.abs(sol["gmsh"][1] - sol["box"][1][idx]).max() < 1e-10


def test_reader_rejects_other_formats(tmp_path):
    p = tmp_path / "bin.msh"
    p.write_text("$MeshFormat\n4.1 1 8\n$EndMeshFormat\n")                 # binary
    with pytest.raises(ValueError):
        read_msh(str(p))
    p.write_text("$MeshFormat\n1.0 0 8\n$EndMeshFormat\n")
    with pytest.raises(ValueError):
        read_msh(str(p))


@pytest.mark.parametrize("nelem,ngl", [([4, 3], 3), ([3, 2, 2], 3)])
def test_msh22_file_gives_the_same_mesh_as_msh41(tmp_path, nelem, ngl):
    """MSH 2.2 ASCII (PETSc 3.12's DMPlexCreateFromFile reads both): same scrambled mesh written in both formats"""
    p4, p2 = str(tmp_path / "a41.msh"), str(tmp_path / "a22.msh")
    box_as_msh(p4, nelem, version=4.1)
    box_as_msh(p2, nelem, version=2.2)
    assert open(p2).read().startswith("$MeshFormat\n2.2 0 8")
    a, b = GmshMesh(p4, ngl), GmshMesh(p2, ngl)
    assert (a.dim, a.nel, a.nnodes) == (b.dim, b.nel, b.nnodes)
    assert np.array_equal(a.elem_nodes, b.elem_nodes) and np.abs(a.corner_xyz - b.corner_xyz).max() == 0.0
    assert np.abs(a.coords - b.coords).max() == 0.0 and a.getBordersNames() == b.getBordersNames()
    for s_ in a.getBordersNames():
        assert np.array_equal(a.getBorderNodes(s_), b.getBorderNodes(s_)) and len(a.getBorderNodes(s_)) > 0
    assert np.array_equal(a.getBordersNodes(), b.getBordersNodes())


def test_positional_border_rule(tmp_path):
    """without physical names a border name means face set index+1 (dmplex.py:167-170)"""
    path = str(tmp_path / "b.msh")
    vx = np.array([[0, 0], [1, 0], [2, 0], [0, 1], [1, 1], [2, 1]], dtype=float)
    cells = [[0, 1, 4, 3], [1, 2, 5, 4]]
    boundary = [(1, [0, 1]), (1, [1, 2]), (2, [2, 5]), (3, [5, 4]), (3, [4, 3]), (4, [3, 0])]
    write_msh(path, vx, cells, boundary)
    g = GmshMesh(path, 3)
    assert g.nnodes == 15 and g.nel == 2
    assert len(g.getBorderNodes("down")) == 5 and len(g.getBorderNodes("right")) == 3
    assert len(g.getBorderNodes("up")) == 5 and len(g.getBorderNodes("left")) == 3
    assert len(g.getBordersNodes()) == 12


REF_MSH = os.path.join(os.environ.get("PYNAMA_REFERENCE", "/root/reference"), "src", "tests", "test.msh")


@pytest.mark.skipif(not os.path.isfile(REF_MSH), reason="reference tree not present")
def test_reference_gmsh_fixture_counts():
    """the reference's own Gmsh known answers (src/tests/test_domain.py:39-48, :62-69: 'this number is from Gmsh')
    on its own mesh file, read where it lies: 33 cells, 153 nodes at ngl = 3, 1688 at ngl = 8; the file's single
    physical group of boundary edges has tag 1, which the reference's positional rule calls "down"
    (dmplex.py:164-170), and no extra border is invented for it"""
    from pynama_b200.domain.gmsh_mesh import GmshMesh
    for ngl, nnodes in ((3, 153), (8, 1688)):
        m = GmshMesh(REF_MSH, ngl)
        assert m.dim == 2 and m.nel == 33 and m.nnodes == nnodes
        assert m.getBordersNames() == ["down", "right", "up", "left"]
        assert len(m.getBorderNodes("down")) == len(m.getBordersNodes()) > 0
        assert all(len(m.getBorderNodes(s)) == 0 for s in ("right", "up", "left"))
    m = GmshMesh(REF_MSH, 2)
    assert m.nnodes == 44                      # the $Nodes header of the file
