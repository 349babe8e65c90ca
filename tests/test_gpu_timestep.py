"""GPU parity of the whole time loop through the reference-shaped API (BaseProblem / TsSolver /
KleSolver): velocity and vorticity after N fixed RK5BS steps against the CPU oracle
(oracle/timestep.py, direct solves).  Tolerance 1e-8 relative to the field's max (KSP rtol 1e-13)."""
import numpy as np
import pytest

from oracle import timestep

pytestmark = pytest.mark.gpu
TOL = 1e-8


def _options(ksp="cg"):
    from pynama_b200.petsc import Options
    o = Options()
    for k, v in (("ksp_type", ksp), ("pc_type", "bjacobi"), ("ksp_rtol", 1e-13), ("ksp_max_it", 50000),
                 ("ts_adapt_type", "none"), ("ts_dt", 2e-3)):
        o.setValue(k, v)
    o.delValue("ksp_fischer_guess")
    return o


def _rel(a, b):
    return np.abs(np.asarray(a) - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("case,nelem,ngl", [("taylor-green", [4, 4], 3),          # BASELINE.json configs[0]
                                            ("taylor-green", [5, 4], 4), ("taylor-green3d", [3, 3, 3], 3)])
def test_free_slip_fields_after_n_steps(case, nelem, ngl):
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.run_case import loadCase
    _options()
    cfg = loadCase(case)
    fem = BaseProblem(cfg, case=case, nelem=nelem, ngl=ngl)
    fem.setUp()
    fem.setUpSolver()
    fem.ts._post = None
    fem.ts.setSolution(fem.vort)
    mesh = fem.dom.getMesh()
    dim = len(nelem)
    orc = timestep.OracleProblem(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes(),
                                 rho=fem.rho, mu=fem.mu, solver="direct")
    assert _rel(fem.vort.getArray(), orc.vort) < 1e-13
    t, k0 = 0.0, None
    for _ in range(3):
        fem.ts.step()
        t, k0 = orc.step(t, 2e-3, k0)
    assert abs(fem.ts.getTime() - t) < 1e-15
    assert _rel(fem.vort.getArray(), orc.vort) < TOL
    assert _rel(fem.solverKLE.getSolution().getArray(), orc.vel) < TOL
    # and the discrete solution stays close to the analytic Taylor-Green field
    a = timestep.tg_alpha(fem.nu, t, dim)
    assert _rel(fem.solverKLE.getSolution().getArray(), timestep.tg_velocity(mesh.coords, a)) < 0.15   # coarse meshes


@pytest.mark.parametrize("case,nelem,ngl", [("cavity-2d", [5, 5], 3), ("cavity", [3, 3, 3], 3)])
def test_no_slip_cavity_fields_after_n_steps(case, nelem, ngl):
    """MatNS + the two-solve NS branch of evalRHS (base_problem.py:119-123, kle_solver.py:22-41)"""
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.run_case import loadCase
    _options("gmres")            # K + Kfs is symmetric but need not be definite
    cfg = loadCase(case)
    fem = BaseProblem(cfg, case=case, nelem=nelem, ngl=ngl)
    fem.setUp()
    fem.setUpSolver()
    fem.ts._post = None
    fem.ts.setSolution(fem.vort)
    mesh = fem.dom.getMesh()
    dim = len(nelem)
    walls = cfg["boundary-conditions"]["no-slip"]
    sides = list(walls.keys())
    orc = timestep.OracleProblemNS(dim, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, sides,
                                   {s: mesh.getBorderNodes(s) for s in sides}, walls, rho=fem.rho, mu=fem.mu)
    assert np.array_equal(fem.dom.getDofClasses(), orc.cls)
    t, k0 = 0.0, None
    for _ in range(2):
        fem.ts.step()
        t, k0 = orc.step(t, 2e-3, k0)
    assert np.abs(orc.vel).max() > 0.5
    assert _rel(fem.solverKLE.getSolution().getArray(), orc.vel) < TOL
    assert _rel(fem.solverKLE.getFreeSlipSolution().getArray(), orc.velfs) < TOL
    assert _rel(fem.vort.getArray(), orc.vort) < TOL


_ORACLE_8 = {}


def _oracle_8cube(mesh, rho, mu):
    """oracle fields after 3 steps of dt = 1e-3 on the 8^3 / ngl=4 mesh, computed once per session
    (about 90 s of scipy CG to a 1e-13 residual)"""
    if not _ORACLE_8:
        orc = timestep.OracleProblem(3, 4, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes(),
                                     rho=rho, mu=mu, solver="cg", rtol=1e-13, pc="bjacobi")
        t, k0 = 0.0, None
        for _ in range(3):
            t, k0 = orc.step(t, 1e-3, k0)
        _ORACLE_8.update(vel=orc.vel.copy(), vort=orc.vort.copy(), t=t)
    return _ORACLE_8


@pytest.mark.parametrize("extra", [{}, {"ksp_cg_single_reduction": 1}, {"kle_matfree": 1}],
                         ids=["bench-options", "single-reduction", "matrix-free"])
def test_bench_options_fields_at_8cube(extra):
    """The options bench.py times (cg + bjacobi, rtol 1e-10, -ksp_fischer_guess 1,16, dt 1e-3; and the
    multi-GPU default -ksp_cg_single_reduction, and -kle_matfree) on the 8^3-element, ngl=4 mesh
    (46 875 dofs: the largest the oracle steps in about a minute): velocity and vorticity after 3 RK5BS
    steps against the oracle run to a 1e-13 residual (base_problem.py:111-136, kle_solver.py:33-37)."""
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.petsc import Options
    from pynama_b200.run_case import loadCase
    o = _options()
    o.setValue("ksp_rtol", 1e-10)
    o.setValue("ksp_fischer_guess", "1,16")
    o.setValue("ts_dt", 1e-3)
    for k, v in extra.items():
        o.setValue(k, v)
    try:
        cfg = loadCase("taylor-green3d")
        fem = BaseProblem(cfg, case="taylor-green3d", nelem=[8, 8, 8], ngl=4)
        fem.setUp()
        fem.setUpSolver()
        fem.ts._post = None
        fem.ts.setSolution(fem.vort)
        mesh = fem.dom.getMesh()
        for _ in range(3):
            fem.ts.step()
        orc = _oracle_8cube(mesh, fem.rho, fem.mu)
        assert abs(fem.ts.getTime() - orc["t"]) < 1e-15
        ev = _rel(fem.solverKLE.getSolution().getArray(), orc["vel"])
        ew = _rel(fem.vort.getArray(), orc["vort"])
        print(f"bench-option parity at 8^3: vel {ev:.3e} vort {ew:.3e} its {fem.kspIterations}")
        assert ev < TOL and ew < TOL
    finally:
        for k in list(extra) + ["ksp_fischer_guess"]:
            Options().delValue(k)


def test_adaptive_steps_follow_the_oracle_controller():
    """-ts_adapt_type basic (the reference never pins a step: ts_solver.py:5-23 leaves dt and the controller
    to PETSc's defaults): the device time stepper against the oracle's restatement of the same controller --
    same accepted step sizes, same rejections, same fields.  Unpinned against PETSc itself."""
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.petsc import Options
    from pynama_b200.run_case import loadCase
    o = _options()
    o.setValue("ts_adapt_type", "basic")
    o.setValue("ts_dt", 0.05)                       # the controller overshoots after the first step: rejections follow
    o.setValue("ts_rtol", 1e-6)
    o.setValue("ts_atol", 1e-6)
    try:
        fem = BaseProblem(loadCase("taylor-green"), case="taylor-green", nelem=[4, 4], ngl=4)
        fem.setUp()
        fem.setUpSolver()
        fem.ts._post = None
        fem.ts.setSolution(fem.vort)
        mesh = fem.dom.getMesh()
        orc = timestep.OracleProblem(2, 4, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes(),
                                     rho=fem.rho, mu=fem.mu, solver="direct")
        t, k0, h = 0.0, None, 0.05
        taken, rej = [], 0
        for _ in range(5):
            fem.ts.step()
            t, k0, hnext, r = timestep.step_adaptive(orc, t, h, k0, rtol=1e-6, atol=1e-6)
            taken.append((fem.ts._last_dt, t))
            rej += r
            # (the device solves with CG to 1e-13, the oracle directly: the error norms differ in the ~10th
            # digit and the controller turns that into step sizes that agree to ~1e-8)
            assert abs(fem.ts.getTime() - t) <= 1e-6 * t, (fem.ts.getTime(), t)
            assert abs(fem.ts.getTimeStep() - hnext) <= 1e-4 * hnext
            h = hnext
        assert rej >= 1 and fem.ts.rejections == rej
        assert _rel(fem.vort.getArray(), orc.vort) < 1e-7
        assert _rel(fem.solverKLE.getSolution().getArray(), orc.vel) < 1e-7
    finally:
        for k in ("ts_rtol", "ts_atol"):
            Options().delValue(k)
        Options().setValue("ts_adapt_type", "none")
