"""GPU: the whole path (assembly, KLE solve, RK5BS steps) on an unstructured .msh input through
Domain/BaseProblem ('gmsh-file', the reference's GmshDom) gives the fields of the same problem on
the structured box, compared at equal coordinates; 2-D and 3-D, scrambled numbering, rotated and
mirrored cells."""
import copy

import numpy as np
import pytest

from tests.test_gmsh_mesh import box_as_msh

pytestmark = pytest.mark.gpu


def _options():
    from pynama_b200.petsc import Options
    o = Options()
    for k, v in (("ksp_type", "cg"), ("pc_type", "bjacobi"), ("ksp_rtol", 1e-13), ("ksp_max_it", 50000),
                 ("ts_adapt_type", "none"), ("ts_dt", 2e-3)):
        o.setValue(k, v)
    for k in ("ksp_fischer_guess", "kle_matfree", "ksp_cg_single_reduction"):
        o.delValue(k)


@pytest.mark.parametrize("case,nelem,ngl", [("taylor-green", [5, 4], 4), ("taylor-green3d", [3, 3, 2], 3)])
def test_fields_on_gmsh_mesh_match_box_mesh(tmp_path, case, nelem, ngl):
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.run_case import loadCase
    _options()
    dim = len(nelem)
    path = str(tmp_path / "mesh.msh")
    box_as_msh(path, nelem, perturb=0.1)
    base = loadCase(case)
    res = {}
    for kind in ("box", "gmsh"):
        cfg = copy.deepcopy(base)
        cfg["domain"] = {"ngl": ngl}
        if kind == "box":
            cfg["domain"]["box-mesh"] = {"nelem": nelem, "lower": [0.0] * dim, "upper": [1.0, 1.5, 0.75][:dim]}
            fem = BaseProblem(cfg, case=case, perturb=0.1)
        else:
            cfg["domain"]["gmsh-file"] = path
            fem = BaseProblem(cfg, case=case)
        fem.setUp()
        fem.setUpSolver()
        fem.ts._post = None
        fem.ts.setSolution(fem.vort)
        for _ in range(2):
            fem.ts.step()
        m = fem.dom.getMesh()
        res[kind] = (m.coords.copy(), fem.solverKLE.getSolution().getArray().reshape(-1, dim).copy(),
                     fem.vort.getArray().reshape(m.nnodes, -1).copy())
        assert fem.dom.getMeshType() == kind
    kb = {tuple(k): i for i, k in enumerate(np.round(res["box"][0] * 1e9).astype(np.int64))}
    idx = np.array([kb[tuple(k)] for k in np.round(res["gmsh"][0] * 1e9).astype(np.int64)])
    for f in (1, 2):
        ref = res["box"][f][idx]
        assert np.abs(res["gmsh"][f] - ref).max() <= 1e-8 * np.abs(ref).max()
