"""Structured quad/hex spectral-element mesh tables (inputs A0 of the hot path).

Produces what the reference's L2 layer hands to the assembly loop
 without PETSc:

* corner coordinates per cell in DMPlex closure-vertex order
  (DMPlexDom.getCellCornersCoords, dmplex.py:93-100; order pinned by
  src/tests/test_dmplex_wrapper.py:23-27 and :105-115);
* element -> global node table in the reference's local order
  (getGlobalNodesFromCell, dmplex.py:193-196 -> indices.py:70-92; local order =
  Spectral.getSpectralOrder, spectral.py:356-441);
* node coordinates as the reference computes them, corner basis at the nodal
  points times the corner coordinates (computeFullCoordinates, dmplex.py:73-88);
* the node sets of the box sides, named as dmplex.py:26-29.

Global node numbering is lexicographic (x fastest) over the full GLL grid.  For
ngl = 2 that is the vertex numbering the reference's tests pin
(test_dmplex_wrapper.py:223-227); for ngl >= 3 the reference's numbering comes
from PetscSection internals that are not in its tree, so matrices agree with a
real PETSc run up to a symmetric permutation ("numbering").
"""
import numpy as np

from .. import _lib

SIDES = {2: ["down", "right", "up", "left"],                          # dmplex.py:27
         3: ["back", "front", "down", "up", "right", "left"]}         # dmplex.py:29
# side name -> (axis, 0 = lower / 1 = upper)
SIDE_PLANE = {"left": (0, 0), "right": (0, 1), "down": (1, 0), "up": (1, 1),
              "back": (2, 0), "front": (2, 1)}
# unit-cell corners in DMPlex closure order
CORNERS = {2: np.array([[0, 0], [1, 0], [1, 1], [0, 1]], dtype=np.int64),
           3: np.array([[0, 0, 0], [0, 1, 0], [1, 1, 0], [1, 0, 0],
                        [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]], dtype=np.int64)}


def corner_basis_at_nodes(dim, ngl):
    """Hc[a, c]: multilinear corner basis of local corner c at local node a
    (the reference's HCooOp, spectral.py:63-64,88-89), from the library's tables."""
    nodes1d, _ = _lib.rule_1d("lobatto", ngl)
    hc, _ = _lib.lagrange_1d([-1.0, 1.0], nodes1d)            # [node point, corner 1-D index]
    node_t = _lib.local_order(dim, ngl)
    corn_t = _lib.local_order(dim, 2)
    H = np.ones((ngl ** dim, 2 ** dim))
    first = True
    for d in range(dim):
        f = hc[node_t[:, d]][:, corn_t[:, d]]
        H = f.copy() if first else H * f
        first = False
    return H, nodes1d


class BoxMesh:
    def __init__(self, nelem, lower, upper, ngl, perturb=0.0, seed=20240229):
        nelem = [int(n) for n in nelem]
        self.dim = dim = len(nelem)
        assert dim in (2, 3), "box mesh must be 2-D or 3-D"
        self.ngl, self.nelem = int(ngl), nelem
        self.lower = np.asarray(lower, dtype=np.float64)
        self.upper = np.asarray(upper, dtype=np.float64)
        p = self.ngl - 1
        self.nn = self.ngl ** dim
        self.nel = int(np.prod(nelem))
        self.grid = np.array([n * p + 1 for n in nelem], dtype=np.int64)   # GLL points per axis
        self.nnodes = int(np.prod(self.grid))
        self.sides = SIDES[dim]

        # vertices (optionally displaced: the seeded non-affine geometry of SURVEY 8d-D2)
        axes = [np.linspace(self.lower[d], self.upper[d], nelem[d] + 1) for d in range(dim)]
        vgrid = np.stack(np.meshgrid(*axes, indexing="ij"), axis=-1)        # [vx, vy(, vz), dim]
        if perturb:
            rng = np.random.default_rng(seed)
            h = (self.upper - self.lower) / np.array(nelem)
            disp = perturb * h * rng.uniform(-1.0, 1.0, vgrid.shape)
            inner = tuple(slice(1, -1) for _ in range(dim))
            vgrid[inner] += disp[inner]
        self.vertex_grid = vgrid

        # cells, lexicographic with x fastest
        cidx = np.stack(np.meshgrid(*[np.arange(n) for n in nelem], indexing="ij"), axis=-1)
        cidx = cidx.reshape(-1, dim)
        order = np.lexsort([cidx[:, d] for d in range(dim)])               # x fastest
        self.cell_ijk = cidx[order]
        corner_ijk = self.cell_ijk[:, None, :] + CORNERS[dim][None, :, :]   # [nel, 2^dim, dim]
        self.corner_xyz = np.ascontiguousarray(
            vgrid[tuple(corner_ijk[..., d] for d in range(dim))], dtype=np.float64)

        # local node -> offset inside the element's GLL grid, from the tables themselves
        Hc, nodes1d = corner_basis_at_nodes(dim, self.ngl)
        unit = Hc @ CORNERS[dim].astype(np.float64)                         # [nn, dim] in [0,1]
        ref = (1.0 + nodes1d) / 2.0
        self.local_offset = np.abs(unit[:, :, None] - ref[None, None, :]).argmin(axis=2)
        assert np.allclose(ref[self.local_offset], unit, atol=1e-13)
        gpos = self.cell_ijk[:, None, :] * p + self.local_offset[None, :, :]  # [nel, nn, dim]
        stride = np.concatenate([[1], np.cumprod(self.grid[:-1])])
        self.elem_nodes = np.ascontiguousarray((gpos * stride).sum(axis=2), dtype=np.int32)

        # coordinates: last element writing a shared node wins, as Vec.setValues(INSERT)
        self.coords = np.zeros((self.nnodes, dim))
        xyz = np.einsum("ac,ecd->ead", Hc, self.corner_xyz)
        self.coords[self.elem_nodes.ravel()] = xyz.reshape(-1, dim)
        self.node_ijk = np.stack(np.unravel_index(np.arange(self.nnodes), self.grid, order="F"),
                                 axis=1)

    # ---- the L2 queries the builders use ---------------------------------------------------
    def getDimension(self):
        return self.dim

    def getBordersNames(self):
        return self.sides

    def getBorderNodes(self, name):
        axis, hi = SIDE_PLANE[name]
        if axis >= self.dim:
            return np.zeros(0, dtype=np.int64)
        want = self.grid[axis] - 1 if hi else 0
        return np.nonzero(self.node_ijk[:, axis] == want)[0]

    def getBordersNodes(self):
        out = [self.getBorderNodes(s) for s in self.sides]
        return np.unique(np.concatenate(out))

    def getCellCornersCoords(self, cell):
        return self.corner_xyz[cell].ravel().copy()

    def getGlobalNodesFromCell(self, cell, shared=True):
        return self.elem_nodes[cell].tolist()

    def getNodesCoordinates(self, nodes):
        return self.coords[np.asarray(nodes, dtype=np.int64)]
