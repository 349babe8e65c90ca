"""Generator of the mesh of BASELINE.json configs[3]: 3-D uniform flow past a cylinder on an
unstructured hexahedral mesh, written as a Gmsh 4.1 file for the reference's way in, ``GmshDom``
(src/domain/dmplex.py:390-395).

The reference tree holds no 3-D mesh (src/tests/test.msh is 2-D; the cases that name .msh files,
src/cases/ibm-static.yaml:13, point at absent ones) and its no-slip matrices split a wall dof into
normal / tangential by COORDINATE axis (src/matrices/mat_ns.py:59-75, boundary.py:151-162), which a
curved wall cannot satisfy.  SURVEY.md 8d-D2 therefore allows "the same box with the cylinder
approximated by no-slip blocked cells": the channel [-4, 12] x [-4, 4] x [0, 2] is cut into cubes of
edge h and the cubes whose centre lies inside the cylinder of diameter 1 around the z axis are
removed.  What is left is NOT a box: the file lists vertices and cells in scrambled order, goes
through the unstructured reader (gmsh_mesh.py) and is cut by recursive coordinate bisection on
several GPUs, like any Gmsh mesh.

Borders: the six channel sides keep the reference's face-set ids 1..6 (dmplex.py:26-29: back, front,
down, up, right, left); the staircase wall of the cylinder is split by the axis of its face normals
into two more physical groups, ``cylinder-x`` (id 7) and ``cylinder-y`` (id 8), so that the
reference's per-side normal/tangential rule applies to them unchanged (boundary_conditions.py reads
the normal axis of a side that is not one of the six box names from the ``-x|-y|-z`` suffix).
"""
import numpy as np

from .gmsh_mesh import write_msh

BOX_TAG = {"back": 1, "front": 2, "down": 3, "up": 4, "right": 5, "left": 6}      # dmplex.py:29, ids 1..6
CYL_TAG = {"cylinder-x": 7, "cylinder-y": 8}
# Gmsh hexahedron: vertices 0-3 bottom face counter-clockwise, 4-7 the top face above them
_HEX = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]])
# the six faces of a cell as (axis, side, the four local Gmsh vertices)
_FACES = [(ax, sd, [k for k in range(8) if _HEX[k, ax] == sd]) for ax in range(3) for sd in (0, 1)]


def blocked_cylinder_channel(h=0.5, lower=(-4.0, -4.0, 0.0), upper=(12.0, 4.0, 2.0), center=(0.0, 0.0),
                             diameter=1.0, seed=7):
    """-> (vertex coordinates [nv, 3], cells [nel, 8] in Gmsh order, boundary [(tag, 4 vertex ids)], names).
    The channel extents must be multiples of h.  Numbering of vertices and cells is shuffled (seed)."""
    lower, upper = np.asarray(lower, float), np.asarray(upper, float)
    n = np.rint((upper - lower) / h).astype(int)
    if np.abs(lower + n * h - upper).max() > 1e-12 * h or n.min() < 1:
        raise ValueError("channel extents must be multiples of h")
    ijk = np.stack(np.meshgrid(*[np.arange(m) for m in n], indexing="ij"), -1).reshape(-1, 3)
    cen = lower + (ijk + 0.5) * h
    solid = (cen[:, 0] - center[0]) ** 2 + (cen[:, 1] - center[1]) ** 2 < (0.5 * diameter) ** 2
    if not solid.any():
        raise ValueError("h is too coarse: no cell centre lies inside the cylinder")
    fluid = np.zeros(tuple(n), bool)
    fluid[tuple(ijk[~solid].T)] = True
    cells_ijk = ijk[~solid]
    vid_of = lambda v: v[..., 0] + (n[0] + 1) * (v[..., 1] + (n[1] + 1) * v[..., 2])       # noqa: E731
    cell_v = vid_of(cells_ijk[:, None, :] + _HEX[None, :, :])                             # lattice vertex ids
    used, inv = np.unique(cell_v, return_inverse=True)
    rng = np.random.default_rng(seed)
    vperm = rng.permutation(len(used))                                                     # scrambled numbering
    cells = vperm[inv.reshape(cell_v.shape)]
    vx = np.stack(np.unravel_index(used, tuple(n + 1), order="F"), 1) * h + lower
    vxyz = np.empty_like(vx)
    vxyz[vperm] = vx
    boundary = []
    side_of = {(0, 0): "left", (0, 1): "right", (1, 0): "down", (1, 1): "up", (2, 0): "back", (2, 1): "front"}
    for c, (i, j, k) in enumerate(cells_ijk):
        for ax, sd, loc in _FACES:
            nb = np.array([i, j, k])
            nb[ax] += 1 if sd else -1
            if 0 <= nb[ax] < n[ax]:
                if fluid[tuple(nb)]:
                    continue
                tag = CYL_TAG["cylinder-x" if ax == 0 else "cylinder-y"]     # a removed cell: the staircase wall
            else:
                tag = BOX_TAG[side_of[(ax, sd)]]
            boundary.append((tag, cells[c, loc][[0, 1, 3, 2]]))          # the four vertices, around the face
    order = rng.permutation(len(cells))
    names = {(2, t): nm for nm, t in {**BOX_TAG, **CYL_TAG}.items()}
    return vxyz, cells[order], boundary, names


def write_cylinder_msh(path, **kw):
    vxyz, cells, boundary, names = blocked_cylinder_channel(**kw)
    write_msh(path, vxyz, cells, boundary, physical_names=names)
    return len(cells), len(vxyz)


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("path")
    ap.add_argument("--h", type=float, default=0.25)
    a = ap.parse_args()
    nel, nv = write_cylinder_msh(a.path, h=a.h)
    print(f"{a.path}: {nel} hexahedra, {nv} vertices")
