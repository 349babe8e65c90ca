"""CPU: the C + OpenMP port that bench.py times as the CPU arm (oracle/ckle.c, oracle/cport.py)
against the numpy/scipy oracle it restates: patterns bit-exact (explicit zeros kept), values to
rounding, products, the CG solve, and fields after RK5BS steps."""
import numpy as np
import pytest

from oracle import assembly, boxmesh, cport, sem, timestep


def _cls(mesh):
    cls = np.zeros(mesh.nnodes * mesh.dim, np.uint8)
    for d in range(mesh.dim):
        cls[mesh.border_nodes * mesh.dim + d] = 1
    return cls


@pytest.mark.parametrize("nelem,ngl,perturb,reuse", [([4, 3], 3, 0.0, True), ([3, 3, 2], 3, 0.15, True),
                                                      ([2, 2, 3], 4, 0.0, False), ([5, 4], 4, 0.1, True)])
def test_c_assembly_equals_numpy_oracle(nelem, ngl, perturb, reuse):
    dim = len(nelem)
    mesh = boxmesh.box_mesh(nelem, [0.0] * dim, [1.0] * dim, ngl, perturb=perturb)
    tab = sem.SpectralTables(ngl, dim)
    cls = _cls(mesh)
    got = cport.assemble_fs(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, reuse=reuse)
    want = assembly.assemble_kle(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes, cls, "FS")
    want.update(assembly.assemble_ops(tab, mesh.corner_xyz, mesh.elem_nodes, mesh.nnodes))
    rng = np.random.default_rng(0)
    for name in ("K", "Krhs", "Rw", "Curl", "SrT", "DivSrT"):
        g, w = got[name].tocsr(), want[name]
        assert np.array_equal(g.indptr, w.indptr), name
        assert np.array_equal(g.indices, w.indices), name
        assert np.abs(g.data - w.data).max() <= 1e-13 * np.abs(w.data).max(), name
        x = rng.uniform(-1, 1, w.shape[1])
        assert np.abs(got[name] @ x - w @ x).max() <= 1e-12 * np.abs(w @ x).max(), name
    assert np.abs(got["W"] - want["W"]).max() <= 1e-15


@pytest.mark.parametrize("pc", ["jacobi", "bjacobi"])
def test_c_problem_steps_like_the_numpy_oracle(pc):
    mesh = boxmesh.box_mesh([3, 3, 3], [0.0] * 3, [1.0] * 3, 3)
    args = (3, 3, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.border_nodes)
    a = timestep.OracleProblem(*args, rho=0.5, mu=0.01, solver="direct")
    b = cport.CProblem(*args, rho=0.5, mu=0.01, rtol=1e-13, pc=pc, fischer=4)
    ta = tb = 0.0
    ka = kb = None
    for _ in range(2):
        ta, ka = a.step(ta, 2e-3, ka)
        tb, kb = b.step(tb, 2e-3, kb)
    assert b.iterations > 0 and b.solves == a.solves
    assert np.abs(a.vel - b.vel).max() <= 1e-9 * np.abs(a.vel).max()
    assert np.abs(a.vort - b.vort).max() <= 1e-9 * np.abs(a.vort).max()
    assert cport.threads() >= 1
