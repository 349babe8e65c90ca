"""GPU checks at BASELINE.json's full single-GPU size (configs[1]: 3-D, 16^3 elements, ngl = 4,
352 947 velocity dofs), where the CPU oracle no longer finishes in seconds: size-independent
properties of the assembled operators and of the solve (symmetry, null space of constants, exact
differentiation of polynomials, partition of unity, linearity, reproducibility, round trip through
the Krylov solve, stored-block product == matrix-free product)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NEL, NGL = 16, 4


@pytest.fixture(scope="module")
def full():
    from pynama_b200.device_mesh import DeviceMesh
    from pynama_b200.domain.box_mesh import BoxMesh
    mesh = BoxMesh([NEL] * 3, [0.0] * 3, [1.0] * 3, NGL, perturb=0.15)     # graded (non-affine) geometry
    dm = DeviceMesh(3, NGL, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
    cls = np.zeros(mesh.nnodes * 3, np.uint8)
    b = mesh.getBordersNodes()
    for d in range(3):
        cls[b * 3 + d] = 1
    dm.set_dof_class(cls)
    mats = dm.assemble_kle("FS")
    curl, srt, div, w = dm.assemble_ops(compact=True)
    mats.update({"Curl": curl, "SrT": srt, "DivSrT": div, "W": w})
    return mesh, dm, cls, mats


def _vec(a):
    from pynama_b200.petsc import Vec
    return Vec().createWithArray(np.ascontiguousarray(a, dtype=np.float64))


def test_sizes_match_the_survey_table(full):
    mesh, dm, cls, mats = full
    assert mesh.nnodes == 117649 and mats["K"].getSize() == (352947, 352947)
    # reference pattern = node graph (x) dense block: SURVEY 8 table, nnz(K) before elimination 1.26e8
    pairs = (NEL * (NGL * NGL) - (NEL - 1)) ** 3
    assert dm.graph_info()["node_pairs"] == pairs
    assert abs(pairs * 9 - 1.26e8) < 0.01e8


def test_K_is_symmetric_and_linear(full):
    mesh, dm, cls, mats = full
    K = mats["K"]
    rng = np.random.default_rng(0)
    x, y = rng.uniform(-1, 1, (2, mesh.nnodes * 3))
    Kx, Ky = (K * _vec(x)).getArray(), (K * _vec(y)).getArray()
    scale = np.abs(Kx).max() * np.sqrt(x.size)
    assert abs(y @ Kx - x @ Ky) <= 1e-12 * scale * np.sqrt(x.size)
    Kz = (K * _vec(2.0 * x - 3.0 * y)).getArray()
    assert np.abs(Kz - (2.0 * Kx - 3.0 * Ky)).max() <= 1e-12 * np.abs(Kx).max()


def test_constants_are_in_the_null_space_of_the_element_operator(full):
    """grad, div and curl of a constant vanish, so K v = Krhs v for a uniform field v (the identity
    rows of both matrices carry the boundary values): KAT-1 of the reference at full size"""
    mesh, dm, cls, mats = full
    v = np.tile([4.0, -1.0, 0.5], mesh.nnodes)
    vbc = np.where(cls == 1, v, 0.0)
    lhs = (mats["K"] * _vec(v)).getArray()
    rhs = (mats["Krhs"] * _vec(vbc)).getArray()
    kmax = np.abs((mats["K"] * _vec(np.where(cls == 0, v, 0.0))).getArray()).max()
    assert np.abs(lhs - rhs).max() <= 1e-11 * kmax
    assert np.abs((mats["Rw"] * _vec(np.zeros(mesh.nnodes * 3))).getArray()).max() == 0.0


def test_operators_differentiate_polynomials_exactly(full):
    mesh, dm, cls, mats = full
    X = mesh.coords
    om = np.array([0.3, -1.1, 0.7])
    v = np.cross(om, X)                                    # rigid rotation: curl = 2*omega, strain rate = 0
    curl = (mats["Curl"] * _vec(v.ravel())).getArray().reshape(-1, 3)
    assert np.abs(curl - 2 * om).max() <= 1e-9
    srt = (mats["SrT"] * _vec(v.ravel())).getArray()
    assert np.abs(srt).max() <= 1e-9
    A = np.array([[0.5, 0.2, -0.3], [0.2, -1.0, 0.4], [-0.3, 0.4, 0.25]])     # v = A x, A symmetric
    s = (mats["SrT"] * _vec((X @ A.T).ravel())).getArray().reshape(-1, 6)
    want = np.array([A[0, 0], A[0, 1], A[1, 1], A[1, 2], A[2, 2], A[2, 0]])   # xx xy yy yz zz zx
    assert np.abs(s - want).max() <= 1e-9
    assert np.abs((mats["Curl"] * _vec(np.tile([1.0, 2.0, 3.0], mesh.nnodes))).getArray()).max() <= 1e-10


def test_lumped_mass_is_a_partition_of_unity(full):
    mesh, dm, cls, mats = full
    assert abs(mats["W"].sum() - 1.0) <= 1e-12              # volume of the unit box
    assert mats["W"].getArray().min() > 0.0


def test_assembly_is_bitwise_reproducible(full):
    mesh, dm, cls, mats = full
    again = dm.assemble_kle("FS")
    x = np.random.default_rng(3).uniform(-1, 1, mesh.nnodes * 3)
    for name in ("K", "Krhs", "Rw"):
        nc = again[name].getSize()[1]
        assert np.array_equal((again[name] * _vec(x[:nc])).getArray(), (mats[name] * _vec(x[:nc])).getArray()), name


def test_matrix_free_product_equals_stored_product(full):
    mesh, dm, cls, mats = full
    K = mats["K"]
    x = _vec(np.random.default_rng(4).uniform(-1, 1, mesh.nnodes * 3))
    y0 = (K * x).getArray()
    K.setMatFree(True)
    y1 = (K * x).getArray()
    K.setMatFree(False)
    assert np.abs(y1 - y0).max() <= 1e-12 * np.abs(y0).max()


@pytest.mark.parametrize("opts", [{}, {"single": True}, {"matfree": True, "single": True}])
def test_solve_round_trip(full, opts):
    """b = K x_s, solve, compare: the Krylov solve inverts the assembled operator at full size"""
    from pynama_b200.petsc import KSP, PC, Vec
    mesh, dm, cls, mats = full
    K = mats["K"]
    rng = np.random.default_rng(5)
    xs = np.where(cls == 0, rng.uniform(-1, 1, mesh.nnodes * 3), 0.25)
    b = K * _vec(xs)
    K.setMatFree(bool(opts.get("matfree")))
    ksp = KSP().create()
    ksp.setType("cg")
    pc = PC().create()
    pc.setType("bjacobi")
    ksp.setPC(pc)
    ksp.setCGSingleReduction(bool(opts.get("single")))
    ksp.setOperators(K)
    ksp.setTolerances(rtol=1e-11, max_it=20000)
    x = Vec(mesh.nnodes * 3)
    ksp.solve(b, x)
    K.setMatFree(False)
    assert ksp.getConvergedReason() == 2
    assert np.abs(x.getArray() - xs).max() <= 1e-7
    r = (K * x).getArray() - b.getArray()
    assert np.linalg.norm(r) <= 1e-9 * np.linalg.norm(b.getArray())


# ---- oracle parity at the full size: a sample of rows ------------------------------------------
def _sample_nodes(mesh):
    """48 row nodes of the 16^3 / ngl=4 GLL grid covering every kind of row: box corners, box edges,
    box faces, element vertices (valence 8), element edges (4), element faces (2), element interiors
    (1), nodes on the planes where the 2/4/8-rank partitions cut (element index 8), first and last row"""
    g = mesh.grid
    p = NGL - 1
    nid = lambda i, j, k: int(i + g[0] * (j + g[1] * k))
    last = int(g[0] - 1)
    mid = 8 * p                       # the partition planes of parallel.py (element 8 of 16)
    picks = [(0, 0, 0), (last, last, last), (last, 0, 0), (0, last, last),          # box corners
             (5, 0, 0), (0, 7, last), (last, last, 10), (mid, 0, 0),                 # box edges
             (4, 5, 0), (last, 7, 11), (6, last, mid), (mid, mid, 0),                # box faces
             (3, 3, 3), (mid, mid, mid), (3 * p, 9 * p, 15 * p), (p, 15 * p, p),     # element vertices
             (mid, mid, mid + 1), (3, 6, 9 + 1), (12 + 2, 9, 3), (mid, mid - 1, mid),  # element edges
             (mid, 10, 11), (4, mid + 1, 2 * p), (7, 8, 3 * p), (mid + 1, mid + 2, mid),  # element faces
             (1, 1, 1), (mid + 1, mid + 2, mid + 1), (46, 47, 46), (22, 23, 26),      # element interiors
             (mid, 5, 7), (11, mid, 13), (20, 22, mid), (mid, mid, 4), (mid - 1, mid, mid)]  # partition planes
    rng = np.random.default_rng(20240229)
    nodes = [nid(*q) for q in picks] + rng.integers(0, mesh.nnodes, 15).tolist()
    return np.unique(np.asarray(nodes, dtype=np.int64))


def _rows(m, nodes, br):
    rows = (np.asarray(nodes)[:, None] * br + np.arange(br)[None, :]).ravel()
    return m.tocsr()[rows]


def _check_rows(got_mat, want, nodes, br, name, rtol=1e-12):
    from tests.util import csr_of
    g = _rows(csr_of(got_mat), nodes, br)
    w = _rows(want, nodes, br)
    w.sort_indices()
    assert np.array_equal(g.indptr, w.indptr), f"{name}: row lengths differ"
    assert np.array_equal(g.indices, w.indices), f"{name}: column indices differ"
    scale = np.abs(w.data).max() if w.nnz else 1.0
    err = np.abs(g.data - w.data).max() if w.nnz else 0.0
    assert err <= rtol * scale, f"{name}: max |diff| {err:.3e} vs scale {scale:.3e}"
    return g.nnz


@pytest.mark.parametrize("perturb", [0.0, 0.15])
def test_sampled_rows_against_the_oracle(perturb):
    """configs[1] (16^3 elements, ngl = 4), affine and perturbed: the rows of 48 sampled nodes of K,
    Krhs, Rw, Curl, SrT, DivSrT and the lumped mass against the oracle's restatement of the cell
    loops (mat_fs.py:150-201, spectral.py:92-228) restricted to the cells touching those nodes:
    pattern bit-exact (explicit zeros included), values <= 1e-12 of the largest entry"""
    from oracle import assembly, sem
    from pynama_b200.device_mesh import DeviceMesh
    from pynama_b200.domain.box_mesh import BoxMesh
    mesh = BoxMesh([NEL] * 3, [0.0] * 3, [1.0] * 3, NGL, perturb=perturb)
    dm = DeviceMesh(3, NGL, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes)
    assert dm.congruence()[1] == (1 if perturb == 0.0 else mesh.nel)
    cls = np.zeros(mesh.nnodes * 3, np.uint8)
    b = mesh.getBordersNodes()
    for d in range(3):
        cls[b * 3 + d] = 1
    dm.set_dof_class(cls)
    nodes = _sample_nodes(mesh)
    want, els = assembly.assemble_rows(sem.SpectralTables(NGL, 3), mesh.corner_xyz, mesh.elem_nodes,
                                       mesh.nnodes, cls, nodes, "FS")
    assert len(els) > 80
    mats = dm.assemble_kle("FS")
    total = 0
    for name in ("K", "Krhs", "Rw"):
        total += _check_rows(mats[name], want[name], nodes, 3, name)
        mats[name].destroy()
    curl, srt, div, w = dm.assemble_ops(compact=True)           # the storage the bench runs on
    total += _check_rows(curl, want["Curl"], nodes, 3, "Curl")
    total += _check_rows(srt, want["SrT"], nodes, 6, "SrT")
    total += _check_rows(div, want["DivSrT"], nodes, 3, "DivSrT")
    assert np.abs(w.getArray()[nodes] - want["W"]).max() <= 1e-13 * want["W"].max()
    assert total > 100000
    dm.destroy()
