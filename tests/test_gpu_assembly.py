"""GPU parity: assembled matrices (k4, k5) through the C ABI against the oracle's
scipy restatement of mat_fs.py / mat_ns.py.  The sparsity pattern (scalar CSR
with the reference's explicit zeros) must be bit-exact, values within 1e-12 of
the largest entry."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import assembly, sem
from tests.util import assert_same_csr, csr_of, fs_classes, ns_classes

pytestmark = pytest.mark.gpu


def _setup(nelem, ngl, perturb=0.0, lower=None, upper=None):
    from pynama_b200.device_mesh import DeviceMesh
    from pynama_b200.domain.box_mesh import BoxMesh
    dim = len(nelem)
    mesh = BoxMesh(nelem, lower or [0.0] * dim, upper or [1.0] * dim, ngl, perturb=perturb)
    dm = DeviceMesh(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
    return mesh, dm, sem.SpectralTables(ngl, dim)


FS_CASES = [([4, 4], 3, 0.0), ([3, 5], 4, 0.15), ([2, 2], 2, 0.0), ([3, 3, 3], 3, 0.0),
            ([2, 3, 2], 4, 0.15), ([1, 1, 1], 3, 0.0), ([2, 2, 2], 2, 0.1)]


@pytest.mark.parametrize("nelem,ngl,perturb", FS_CASES)
def test_free_slip_matrices(nelem, ngl, perturb):
    mesh, dm, tab = _setup(nelem, ngl, perturb)
    cls = fs_classes(mesh)
    dm.set_dof_class(cls)
    got = dm.assemble_kle("FS")
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    for name in ("K", "Krhs", "Rw"):
        assert_same_csr(csr_of(got[name]), want[name], name=name)
    # Rd is created but never filled in free-slip mode (mat_fs.py:155)
    assert got["Rd"].nnz()[0] == 0
    # node-graph statistics: pairs = prod(nel_d (p+1)^2 - (nel_d - 1))  
    p1 = ngl
    pairs = int(np.prod([n * p1 * p1 - (n - 1) for n in nelem]))
    info = dm.graph_info()
    assert info["node_pairs"] == pairs
    assert info["coo_length"] == mesh.nel * mesh.nn ** 2


NS_CASES = [([4, 3], 3, None), ([3, 3], 4, ["up", "left"]), ([3, 2, 2], 3, None),
            ([2, 2, 2], 4, ["up", "down", "right"]),
            ([8, 6, 6], 3, None)]         # a cavity mesh well above 3^3 cells (added after the last GPU call of round 2)


@pytest.mark.parametrize("nelem,ngl,sides", NS_CASES)
def test_no_slip_matrices(nelem, ngl, sides):
    mesh, dm, tab = _setup(nelem, ngl)
    cls = ns_classes(mesh, sides)
    dm.set_dof_class(cls)
    got = dm.assemble_kle("NS")
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "NS")
    for name in ("K", "Krhs", "Rw", "Rd", "Kfs", "Krhsfs", "Rwfs", "Rdfs"):
        assert_same_csr(csr_of(got[name]), want[name], name=name)
    # K + Kfs on the union pattern (MatAXPY, kle_solver.py:25)
    ksum = got["K"] + got["Kfs"]
    ref = (want["K"] + want["Kfs"]).tocsr()
    g = csr_of(ksum)
    ones = lambda m: sp.csr_matrix((np.ones(m.nnz), m.indices, m.indptr), shape=m.shape)
    pat_ref = (ones(want["K"]) + ones(want["Kfs"])).tocsr()
    assert (pat_ref.sign() != ones(g)).nnz == 0, "K+Kfs: union pattern differs"
    assert abs(g - ref).max() <= 1e-12 * abs(ref).max()


OPS_CASES = [([4, 4], 3, 0.0), ([3, 2], 5, 0.15), ([3, 3, 3], 3, 0.0), ([2, 2, 3], 4, 0.15)]


@pytest.mark.parametrize("nelem,ngl,perturb", OPS_CASES)
@pytest.mark.parametrize("compact", [True, False])
def test_operator_matrices(nelem, ngl, perturb, compact):
    mesh, dm, tab = _setup(nelem, ngl, perturb)
    curl, srt, div, w = dm.assemble_ops(compact=compact)
    want = assembly.assemble_ops(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
    # the exported CSR carries the reference's explicit zeros in both storage modes
    assert_same_csr(csr_of(curl), want["Curl"], name="Curl")
    assert_same_csr(csr_of(srt), want["SrT"], name="SrT")
    assert_same_csr(csr_of(div), want["DivSrT"], name="DivSrT")
    assert np.abs(w.getArray() - want["W"]).max() <= 1e-13 * want["W"].max()
    if compact:
        dim = len(nelem)
        assert curl.nblocks() < want["Curl"].nnz / (dim * tab.dim_w)    # zero blocks not stored
    # SpMV of both storage modes equals the reference operator applied to a vector
    from pynama_b200.petsc import Vec
    rng = np.random.default_rng(12345)
    x = rng.uniform(-1, 1, curl.getSize()[1])
    y = curl * Vec().createWithArray(x)
    ref = want["Curl"] @ x
    assert np.abs(y.getArray() - ref).max() <= 1e-12 * np.abs(ref).max()


def test_assembly_is_reproducible():
    """fixed summation order: two assemblies are bit-identical"""
    mesh, dm, _ = _setup([3, 3, 2], 3, 0.15)
    dm.set_dof_class(fs_classes(mesh))
    a = dm.assemble_kle("FS")["K"].getValuesCSR()[2]
    b = dm.assemble_kle("FS")["K"].getValuesCSR()[2]
    assert np.array_equal(a, b)


@pytest.mark.parametrize("nelem,ngl,upper", [([4, 3], 3, [1.0, 1.0]), ([3, 2, 4], 3, [3.0, 1.0, 1.0]),
                                              ([2, 2, 2], 4, [1.0, 1.0, 1.0])])
def test_congruence_classes_uniform_box(nelem, ngl, upper):
    """a uniform box is ONE congruence class: element blocks are evaluated once and every
    assembled matrix agrees with the element-by-element evaluation (and with the oracle, which
    calls getElemKLEMatrices per cell as mat_fs.py:152-155 does) to 1e-12"""
    mesh, dm, tab = _setup(nelem, ngl, 0.0, upper=upper)
    tol, nclass = dm.congruence()
    assert tol == 1e-13 and nclass == 1
    cls = ns_classes(mesh, None)
    dm.set_dof_class(cls)
    merged = dm.assemble_kle("NS")
    ops_m = dm.assemble_ops(compact=True)
    assert dm.set_congruence(0.0) == mesh.nel
    single = dm.assemble_kle("NS")
    ops_s = dm.assemble_ops(compact=True)
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "NS")
    for name in ("K", "Krhs", "Rw", "Rd", "Kfs", "Krhsfs", "Rwfs", "Rdfs"):
        a, b = csr_of(merged[name]), csr_of(single[name])
        assert np.array_equal(a.indptr, b.indptr) and np.array_equal(a.indices, b.indices)
        scale = max(np.abs(b.data).max(), 1e-300) if b.nnz else 1.0
        assert b.nnz == 0 or np.abs(a.data - b.data).max() <= 1e-12 * scale
        assert_same_csr(a, want[name], name=name)
    for a, b in zip(ops_m[:3], ops_s[:3]):
        a, b = csr_of(a), csr_of(b)
        assert np.array_equal(a.indices, b.indices)
        assert np.abs(a.data - b.data).max() <= 1e-12 * np.abs(b.data).max()
    assert np.abs(ops_m[3].getArray() - ops_s[3].getArray()).max() <= 1e-13 * ops_s[3].getArray().max()
    assert dm.set_congruence(1e-13) == 1


def test_congruence_classes_mixed_mesh():
    """cells with a displaced corner are never merged (every cell of a 3x3x2 box touches an
    interior corner); a box with two element sizes along x has two classes"""
    mesh, dm, _ = _setup([3, 3, 2], 3, 0.15)
    assert dm.congruence()[1] == mesh.nel
    from pynama_b200.device_mesh import DeviceMesh
    from pynama_b200.domain.box_mesh import BoxMesh
    m2 = BoxMesh([4, 2], [0.0, 0.0], [1.0, 1.0], 3)
    cx = m2.corner_xyz.copy().reshape(m2.nel, 4, 2)
    # stretch the x >= 0.5 half by 2: two congruence classes, still a conforming mesh
    cx[..., 0] = np.where(cx[..., 0] > 0.5, 0.5 + 2.0 * (cx[..., 0] - 0.5), cx[..., 0])
    dm2 = DeviceMesh(2, 3, cx, m2.elem_nodes, m2.nnodes)
    assert dm2.congruence()[1] == 2
    tab = sem.SpectralTables(3, 2)
    cls = np.zeros(m2.nnodes * 2, np.uint8)
    cls[: 2 * 3] = 1
    dm2.set_dof_class(cls)
    got = dm2.assemble_kle("FS")
    want = assembly.assemble_kle(tab, cx, m2.elem_nodes, m2.nnodes, cls, "FS")
    for name in ("K", "Krhs", "Rw"):
        assert_same_csr(csr_of(got[name]), want[name], name=name)


@pytest.mark.parametrize("mode,nelem,ngl,perturb", [("FS", [3, 3, 2], 3, 0.15), ("NS", [4, 3], 4, 0.0),
                                                     ("NS", [2, 2, 2], 3, 0.1)])
def test_fused_kle_pass_equals_per_matrix_pass(mode, nelem, ngl, perturb, monkeypatch):
    """K, Krhs, Rw, Rd from the single pass over the node graph are bit-identical to the
    one-gather-per-matrix kernels (same blocks, same ascending-element summation order)"""
    mesh, dm, _ = _setup(nelem, ngl, perturb)
    dm.set_dof_class(ns_classes(mesh, None) if mode == "NS" else fs_classes(mesh))
    fused = dm.assemble_kle(mode)
    monkeypatch.setenv("PYN_ASSEMBLE_PER_MATRIX", "1")
    per = dm.assemble_kle(mode)
    for name in ("K", "Krhs", "Rw", "Rd"):
        a, b = fused[name].getValuesCSR(), per[name].getValuesCSR()
        for x, y in zip(a, b):
            assert np.array_equal(x, y), name


@pytest.mark.parametrize("mode,nelem,ngl,perturb", [("FS", [3, 3, 2], 3, 0.15), ("NS", [2, 2, 2], 4, 0.0),
                                                     ("NS", [4, 3], 4, 0.1)])
def test_64bit_index_path_equals_32bit_and_oracle(mode, nelem, ngl, perturb):
    """BASELINE.json configs[2] (32^3, ngl=5: 4.09e9 scalar non-zeros in K) is the only mesh large
    enough to take the 64-bit instantiation of the numeric kernel by itself; pyn_tune("force_idx64")
    makes a small mesh take it, and the result must be the oracle's and bit-identical to the 32-bit one"""
    from pynama_b200 import _lib
    mesh, dm, tab = _setup(nelem, ngl, perturb)
    cls = ns_classes(mesh, None) if mode == "NS" else fs_classes(mesh)
    dm.set_dof_class(cls)
    a = dm.assemble_kle(mode)
    _lib.call("pyn_tune", b"force_idx64", 1)
    try:
        b = dm.assemble_kle(mode)
    finally:
        _lib.call("pyn_tune", b"force_idx64", 0)
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, mode)
    for name in want:
        if mode == "FS" and name == "Rd":
            continue
        ga, gb = a[name].getValuesCSR(), b[name].getValuesCSR()
        for x, y in zip(ga, gb):
            assert np.array_equal(x, y), name
        assert_same_csr(csr_of(b[name]), want[name], name=name)
