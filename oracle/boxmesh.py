"""Oracle: structured box mesh tables (what the reference's DMPlex layer hands to the assembly).

TEST INFRASTRUCTURE -- see oracle/__init__.py.  A standalone generator so that the CPU arm of
bench.py (`--impl reference`) runs without importing the product package.  Conventions restated:

* cell corners in DMPlex closure order          src/tests/test_dmplex_wrapper.py:23-27, :105-115
* node positions = corner basis at the nodal points times the corner coordinates
                                                 src/domain/dmplex.py:73-88 (oracle.sem.node_coords)
* side names                                     src/domain/dmplex.py:26-29
* global node numbering: lexicographic over the GLL grid, x fastest (the numbering the product
  uses; PETSc's own for ngl >= 3 is not in the reference tree -- DESIGN.md section 4)

tests/test_oracle_rows.py checks these tables against pynama_b200.domain.box_mesh.BoxMesh.
"""
from types import SimpleNamespace

import numpy as np

from . import sem

UNIT_CORNERS = {2: [(0, 0), (1, 0), (1, 1), (0, 1)],
                3: [(0, 0, 0), (0, 1, 0), (1, 1, 0), (1, 0, 0), (0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1)]}
SIDE_AXIS_END = {"left": (0, 0), "right": (0, 1), "down": (1, 0), "up": (1, 1), "back": (2, 0), "front": (2, 1)}


def box_mesh(nelem, lower, upper, ngl, perturb=0.0, seed=20240229):
    dim = len(nelem)
    p = ngl - 1
    ne = np.asarray(nelem, dtype=np.int64)
    lo, hi = np.asarray(lower, float), np.asarray(upper, float)
    grid = ne * p + 1
    # vertex lattice, interior vertices optionally displaced
    lattice = np.stack(np.meshgrid(*[np.linspace(lo[d], hi[d], ne[d] + 1) for d in range(dim)], indexing="ij"), -1)
    if perturb:
        shift = perturb * ((hi - lo) / ne) * np.random.default_rng(seed).uniform(-1.0, 1.0, lattice.shape)
        core = (slice(1, -1),) * dim
        lattice[core] += shift[core]
    # cells: x fastest
    cells = np.array(np.unravel_index(np.arange(int(ne.prod())), ne, order="F")).T        # [nel, dim]
    uc = np.array(UNIT_CORNERS[dim], dtype=np.int64)
    cidx = cells[:, None, :] + uc[None, :, :]
    corner_xyz = lattice[tuple(cidx[..., d] for d in range(dim))]                          # [nel, 2^dim, dim]
    # where each local node sits in the element's own GLL lattice
    tab = sem.SpectralTables(ngl, dim)
    unit = sem.node_coords(tab, uc.astype(float))                                          # in [0, 1]^dim
    x1 = (1.0 + sem.gauss_lobatto(ngl)[0]) / 2.0
    off = np.abs(unit[:, :, None] - x1[None, None, :]).argmin(2)                           # [nn, dim]
    assert np.abs(x1[off] - unit).max() < 1e-13
    gpos = cells[:, None, :] * p + off[None, :, :]
    stride = np.concatenate([[1], np.cumprod(grid[:-1])])
    elem_nodes = (gpos * stride).sum(2).astype(np.int32)
    nnodes = int(grid.prod())
    coords = np.zeros((nnodes, dim))
    Hc = tab.nodal.Hc
    coords[elem_nodes.ravel()] = np.einsum("ac,ecd->ead", Hc, corner_xyz).reshape(-1, dim)  # last writer wins
    ijk = np.array(np.unravel_index(np.arange(nnodes), grid, order="F")).T
    sides = {}
    for name, (ax, end) in SIDE_AXIS_END.items():
        if ax < dim:
            sides[name] = np.flatnonzero(ijk[:, ax] == (grid[ax] - 1 if end else 0))
    border = np.unique(np.concatenate(list(sides.values())))
    return SimpleNamespace(dim=dim, ngl=ngl, nel=len(cells), nnodes=nnodes, grid=grid,
                           corner_xyz=np.ascontiguousarray(corner_xyz), elem_nodes=elem_nodes, coords=coords,
                           sides=sides, border_nodes=border)
