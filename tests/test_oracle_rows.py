"""CPU: the oracle's row-sample assembly (oracle/assembly.py assemble_rows, used by the GPU parity
test at BASELINE.json's full size) equals the same rows of the whole-mesh oracle assembly, bit for
bit -- same cells, same ascending order."""
import numpy as np
import pytest

from oracle import assembly, sem
from tests.util import fs_classes, ns_classes


def _rows_of(m, nodes, br):
    rows = (np.asarray(nodes)[:, None] * br + np.arange(br)[None, :]).ravel()
    return m.tocsr()[rows]


@pytest.mark.parametrize("nelem,ngl,mode,perturb", [([4, 3, 3], 3, "FS", 0.15), ([3, 3, 3], 3, "NS", 0.0),
                                                     ([5, 4], 4, "FS", 0.0)])
def test_sampled_rows_equal_full_assembly(nelem, ngl, mode, perturb):
    from pynama_b200.domain.box_mesh import BoxMesh
    dim = len(nelem)
    mesh = BoxMesh(nelem, [0.0] * dim, [1.0] * dim, ngl, perturb=perturb)
    tab = sem.SpectralTables(ngl, dim)
    cls = fs_classes(mesh) if mode == "FS" else ns_classes(mesh, None)
    full = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, mode)
    full.update(assembly.assemble_ops(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes))
    rng = np.random.default_rng(1)
    nodes = np.unique(np.concatenate([[0, mesh.nnodes - 1], rng.integers(0, mesh.nnodes, 12)]))
    got, els = assembly.assemble_rows(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, nodes, mode)
    assert len(els) < mesh.nel
    br = {"Curl": tab.dim_w, "SrT": tab.dim_s}
    for name, m in got.items():
        if name == "W":
            assert np.array_equal(m, full["W"][nodes])
            continue
        a, b = _rows_of(m, nodes, br.get(name, dim)), _rows_of(full[name], nodes, br.get(name, dim))
        assert np.array_equal(a.indptr, b.indptr) and np.array_equal(a.indices, b.indices), name
        assert np.array_equal(a.data, b.data), name
        assert m.nnz == a.nnz, name                        # nothing outside the sampled rows


@pytest.mark.parametrize("nelem,ngl,perturb", [([4, 3], 3, 0.0), ([3, 2, 4], 4, 0.15), ([2, 2, 2], 2, 0.1)])
def test_oracle_box_mesh_equals_the_product_box_mesh(nelem, ngl, perturb):
    """the CPU arm of bench.py builds its mesh from oracle/boxmesh.py (no product import there)"""
    from oracle import boxmesh
    from pynama_b200.domain.box_mesh import BoxMesh
    dim = len(nelem)
    a = BoxMesh(nelem, [0.0] * dim, [1.0, 2.0, 1.5][:dim], ngl, perturb=perturb)
    b = boxmesh.box_mesh(nelem, [0.0] * dim, [1.0, 2.0, 1.5][:dim], ngl, perturb=perturb)
    assert np.array_equal(a.corner_xyz, b.corner_xyz)
    assert np.array_equal(a.elem_nodes, b.elem_nodes)
    assert np.abs(a.coords - b.coords).max() <= 1e-14
    assert np.array_equal(a.getBordersNodes(), b.border_nodes)
    for s in a.sides:
        assert np.array_equal(a.getBorderNodes(s), b.sides[s])
