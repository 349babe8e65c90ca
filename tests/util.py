"""Shared helpers of the test-suite."""
import os

import numpy as np
import scipy.sparse as sp

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def fs_classes(mesh):
    """all dofs of all boundary nodes are Dirichlet (BC types 'uniform' / 'custom-func')"""
    cls = np.zeros(mesh.nnodes * mesh.dim, np.uint8)
    b = mesh.getBordersNodes()
    for d in range(mesh.dim):
        cls[b * mesh.dim + d] = 1
    return cls


def ns_classes(mesh, sides=None):
    """no-slip on the given sides: normal dof class 2, the others class 3
    (normal wins where sides meet, mat_ns.py:60-62)"""
    from pynama_b200.domain.box_mesh import SIDE_PLANE
    cls = np.zeros(mesh.nnodes * mesh.dim, np.uint8)
    sides = mesh.sides if sides is None else sides
    for s in sides:
        nodes = mesh.getBorderNodes(s)
        for d in range(mesh.dim):
            idx = nodes * mesh.dim + d
            cls[idx] = np.where(cls[idx] == 2, 2, 3)
    for s in sides:
        nodes = mesh.getBorderNodes(s)
        cls[nodes * mesh.dim + SIDE_PLANE[s][0]] = 2
    return cls


def csr_of(mat):
    ip, ix, v = mat.getValuesCSR()
    return sp.csr_matrix((v, ix, ip), shape=mat.getSize())


def assert_same_csr(got, want, rtol=1e-12, name=""):
    """bit-exact pattern, values within rtol of the largest entry"""
    want = want.tocsr()
    want.sort_indices()
    assert got.shape == want.shape, f"{name}: shape {got.shape} vs {want.shape}"
    assert got.nnz == want.nnz, f"{name}: nnz {got.nnz} vs {want.nnz}"
    assert np.array_equal(got.indptr, want.indptr), f"{name}: row pointers differ"
    assert np.array_equal(got.indices, want.indices), f"{name}: column indices differ"
    scale = np.abs(want.data).max() if want.nnz else 1.0
    err = np.abs(got.data - want.data).max() if want.nnz else 0.0
    assert err <= rtol * max(scale, 1e-300), f"{name}: max |diff| {err:.3e} vs scale {scale:.3e}"
    return err / max(scale, 1e-300)
