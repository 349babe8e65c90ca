"""GPU parity against the reference's OWN outputs: tests/golden/refpath_*.npz (the reference's unmodified
Domain / MatFS / MatNS / KleSolver / BaseProblem.evalRHS run over the numpy stand-in for petsc4py, see
tests/golden/make_golden_refpath.py).  Ten cases (BASELINE.json configs[0] among them): free-slip 2-D / 3-D and no-slip cavity 2-D / 3-D, affine and
non-affine cells, ngl 2-5.

* every assembled matrix (K, Krhs, Rw, Rd, Kfs, Krhsfs, Rwfs, Rdfs, Curl, SrT, DivSrT) through the C ABI:
  the reference's CSR pattern bit for bit (explicit zeros included), values to 1e-12 of the largest entry;
* vorticity and velocity after the golden's fixed RK5BS steps through the reference-shaped API
  (BaseProblem / TsSolver / KleSolver on the device): 1e-8 of the field's max (KSP rtol 1e-13).

The calls are the ones tests/test_gpu_assembly.py and tests/test_gpu_timestep.py make against the oracle; the
expected values here come from the reference itself.  (Written after the round's GPU budget was spent: first
run is the driver's.)
"""
import glob
import os

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import assembly, timestep
from tests.util import assert_same_csr, csr_of

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[len("refpath_"):-4] for p in glob.glob(os.path.join(GOLD, "refpath_*.npz")))


def box_of(g):
    """(case file, box keywords) the golden was made with"""
    box = dict(lower=[float(v) for v in g["box_lower"]], upper=[float(v) for v in g["box_upper"]],
               nelem=[int(v) for v in g["box_nelem"]])
    if float(g["perturb"]):
        box["perturb"] = float(g["perturb"])
    return str(g["case"]), box


TOL = 1e-8


def load(name):
    return np.load(os.path.join(GOLD, f"refpath_{name}.npz"))


def golden_csr(g, nm):
    return sp.csr_matrix((g[f"{nm}_data"], g[f"{nm}_indices"], g[f"{nm}_indptr"]), shape=tuple(g[f"{nm}_shape"]))


def dof_classes(g):
    """as tests/test_reference_path.py (where the result is checked against the reference's own sets)"""
    dim, n = int(g["dim"]), len(g["coords"])
    sides = [str(s) for s in g["sides"]]
    side_nodes = {s: g[f"side_nodes_{s}"] for s in sides}
    if str(g["bc_type"]) == "NS":
        return timestep.ns_dof_classes(sides, side_nodes, n, dim)
    cls = np.zeros(n * dim, np.uint8)
    for s in sides:
        for d in range(dim):
            cls[side_nodes[s] * dim + d] = 1
    return cls


def _rel(a, b):
    return np.abs(np.asarray(a) - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("name", CASES)
def test_assembled_matrices_equal_the_reference_goldens(name):
    from pynama_b200.device_mesh import DeviceMesh
    g = load(name)
    dim, ngl, n = int(g["dim"]), int(g["ngl"]), len(g["coords"])
    ns = str(g["bc_type"]) == "NS"
    dm = DeviceMesh(dim, ngl, np.ascontiguousarray(g["corner_xyz"]), np.ascontiguousarray(g["elem_nodes"]), n)
    dm.set_dof_class(dof_classes(g))
    got = dm.assemble_kle("NS" if ns else "FS")
    for nm in (["K", "Krhs", "Rw", "Rd", "Kfs", "Krhsfs", "Rwfs", "Rdfs"] if ns else ["K", "Krhs", "Rw"]):
        assert_same_csr(csr_of(got[nm]), golden_csr(g, nm), name=f"{name}:{nm}")
    if not ns:
        assert got["Rd"].nnz()[0] == 0 and golden_csr(g, "Rd").nnz == 0
    else:
        ref = golden_csr(g, "KplusKfs")                       # MatAXPY on the union pattern (kle_solver.py:25)
        ksum = csr_of(got["K"] + got["Kfs"])
        ksum.sort_indices()
        assert np.array_equal(ksum.indptr, ref.indptr) and np.array_equal(ksum.indices, ref.indices)
        assert abs(ksum - ref).max() <= 1e-12 * abs(ref).max()
    curl, srt, div, _w = dm.assemble_ops(compact=False)
    assert_same_csr(csr_of(curl), golden_csr(g, "Curl"), name=f"{name}:Curl")
    assert_same_csr(csr_of(srt), golden_csr(g, "SrT"), name=f"{name}:SrT")
    assert_same_csr(csr_of(div), golden_csr(g, "DivSrT"), name=f"{name}:DivSrT")


@pytest.mark.parametrize("name", [c for c in CASES if c != "uniform2d"])
def test_fields_after_fixed_steps_equal_the_reference_goldens(name):
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.petsc import Options
    from pynama_b200.run_case import loadCase
    g = load(name)
    ns = str(g["bc_type"]) == "NS"
    o = Options()
    for k, v in (("ksp_type", "gmres" if ns else "cg"), ("pc_type", "bjacobi"), ("ksp_rtol", 1e-13),
                 ("ksp_max_it", 50000), ("ts_adapt_type", "none"), ("ts_dt", float(g["dt"]))):
        o.setValue(k, v)
    o.delValue("ksp_fischer_guess")
    case, box = box_of(g)
    cfg = loadCase(case)
    fem = BaseProblem(cfg, case=case, ngl=int(g["ngl"]), **box)
    fem.setUp()
    fem.setUpSolver()
    fem.ts._post = None
    fem.ts.setSolution(fem.vort)
    mesh = fem.dom.getMesh()
    assert np.array_equal(mesh.elem_nodes, g["elem_nodes"]) and np.abs(mesh.corner_xyz - g["corner_xyz"]).max() == 0.0
    assert np.array_equal(fem.dom.getDofClasses(), dof_classes(g))
    assert _rel(fem.vort.getArray(), g["vort0"]) < 1e-13 or np.abs(g["vort0"]).max() == 0.0
    for _ in range(int(g["nsteps"])):
        fem.ts.step()
    assert abs(fem.ts.getTime() - float(g["t_steps"])) < 1e-15
    ev = _rel(fem.solverKLE.getSolution().getArray(), g["vel_steps"])
    ew = _rel(fem.vort.getArray(), g["vort_steps"])
    print(f"{name}: vel {ev:.3e} vort {ew:.3e} against the reference's own run")
    assert ev < TOL and ew < TOL


def test_no_slip_classes_used_above_are_the_references():
    g = load("cavity3d")
    cls = dof_classes(g)
    assert np.array_equal(np.flatnonzero(cls == assembly.NORMAL), g["normal_dofs"])
