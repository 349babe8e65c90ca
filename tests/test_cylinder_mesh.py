"""CPU: the mesh of BASELINE.json configs[3] (domain/cylinder_mesh.py: blocked-cell cylinder in a
channel, Gmsh 4.1 file) through the unstructured reader, the two extra wall borders in the boundary
layer, and the no-slip dof classes of the drop-in Domain against the oracle's restatement of
boundary_conditions.py:214-243 / mat_ns.py:59-62."""
import os

import numpy as np
import pytest

from oracle import timestep
from pynama_b200.domain.cylinder_mesh import blocked_cylinder_channel, write_cylinder_msh
from pynama_b200.domain.gmsh_mesh import GmshMesh


@pytest.fixture(scope="module")
def small(tmp_path_factory):
    path = os.path.join(tmp_path_factory.mktemp("cyl"), "cyl.msh")
    nel, nv = write_cylinder_msh(path, h=0.25, lower=(-2.0, -1.5, 0.0), upper=(3.0, 1.5, 0.5))
    return path, nel, nv


def test_cells_borders_and_counts(small):
    path, nel, nv = small
    # 20 x 12 x 2 cubes minus the 12 per layer whose centre lies inside r = 0.5
    assert nel == 20 * 12 * 2 - 12 * 2
    m = GmshMesh(path, 3)
    assert m.nel == nel and m.sides == ["back", "front", "down", "up", "right", "left", "cylinder-x", "cylinder-y"]
    # every exterior face belongs to exactly one border
    ext = m.getBordersNodes()
    union = np.unique(np.concatenate([m.getBorderNodes(s) for s in m.sides]))
    assert np.array_equal(ext, union)
    for name, axis in (("cylinder-x", 0), ("cylinder-y", 1)):
        X = m.coords[m.getBorderNodes(name)]
        assert len(X) and np.abs(X[:, :2]).max() <= 0.5 + 1e-12              # on the staircase around the axis
        # faces of this group are planes of constant x (resp. y): multiples of h
        assert np.abs(X[:, axis] / 0.25 - np.rint(X[:, axis] / 0.25)).max() < 1e-12
    # the fluid volume: channel minus the blocked cubes (lumped mass is checked on the GPU side)
    vol = sum(np.prod(np.ptp(c, axis=0)) for c in m.corner_xyz)
    assert abs(vol - (5.0 * 3.0 * 0.5 - 12 * 2 * 0.25 ** 3)) < 1e-12


def test_numbering_is_scrambled_but_the_mesh_is_the_same():
    a = blocked_cylinder_channel(h=0.5, lower=(-2, -1.5, 0), upper=(3, 1.5, 1), seed=1)
    b = blocked_cylinder_channel(h=0.5, lower=(-2, -1.5, 0), upper=(3, 1.5, 1), seed=2)
    assert not np.array_equal(a[1], b[1])
    ca = np.sort(a[0][a[1]].mean(axis=1).round(9).view([("", float)] * 3), axis=0)
    cb = np.sort(b[0][b[1]].mean(axis=1).round(9).view([("", float)] * 3), axis=0)
    assert np.array_equal(ca, cb)


def test_dof_classes_of_the_case_equal_the_oracle_rule(small, monkeypatch):
    """Domain + BoundaryConditions on the 8-border mesh (no GPU needed up to the dof classes)"""
    from pynama_b200.domain.domain import Domain
    from pynama_b200.run_case import loadCase
    path, _, _ = small
    cfg = loadCase("cylinder")
    cfg["domain"]["gmsh-file"] = path
    dom = Domain()
    dom.configure(cfg)
    dom.setOptions(ngl=3)
    dom.create()
    dom.setUpBoundaryConditions()
    assert dom.getBoundaryType() == "NS"
    mesh = dom.getMesh()
    walls = cfg["boundary-conditions"]["no-slip"]
    sides = list(walls.keys())
    want = timestep.ns_dof_classes(sides, {s: mesh.getBorderNodes(s) for s in sides}, mesh.nnodes, 3)
    got = dom.getDofClasses()
    assert np.array_equal(got, want)
    # a convex corner line of the staircase belongs to an x-wall and a y-wall: x and y normal, z tangential
    both = np.intersect1d(mesh.getBorderNodes("cylinder-x"), mesh.getBorderNodes("cylinder-y"))
    z = mesh.coords[both, 2]
    both = both[(z > 1e-9) & (z < 0.5 - 1e-9)]                # away from the back / front planes (z normal there)
    assert len(both) and (got[both * 3] == 2).all() and (got[both * 3 + 1] == 2).all() and (got[both * 3 + 2] == 3).all()
    assert timestep.side_axis("cylinder-y") == 1 and timestep.side_axis("left") == 0
