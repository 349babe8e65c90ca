"""Command-line driver, same options as the reference's ``src/run_case.py``:

    python -m pynama_b200.run_case -case taylor-green3d [-test kle] [-log INFO]
        [-ksp_type cg -pc_type bjacobi -ksp_rtol 1e-8] [-ts_adapt_type none -ts_dt 1e-3]

``-case`` names a YAML file under pynama_b200/cases (same schema as the
reference's src/cases/*.yaml) or a path to one.  The time loop is the
reference's (run_case.py:152-162): setUp, setUpSolver, ts.solve(fem.vort).
"""
import logging
import os
import sys

import yaml

from .petsc import Options

CASE_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cases")


def loadCase(case):
    path = case if os.path.exists(case) else os.path.join(CASE_DIR, f"{case}.yaml")
    with open(path) as f:
        return yaml.load(f, Loader=yaml.Loader)


def timeSolving(config, FemProblem, **kwargs):
    fem = FemProblem(config, **kwargs)
    fem.setUp()
    fem.setUpSolver()
    if not fem.comm.rank:
        fem.logger.info("Solving problem...")
    fem.timer.tic()
    if hasattr(fem, "startSolver"):
        # immersed-boundary cases own their time loop (immersed_boundary.py:56-81, :303-331); the
        # reference selects them at run_case.py:20-23
        fem.startSolver()
    else:
        fem.ts.solve(fem.vort)
    if not fem.comm.rank:
        fem.logger.info(f"Solver Finished in {fem.timer.toc()} seconds")
        fem.logger.info(f"Total time: {fem.timerTotal.toc()} seconds")
    return fem


def main(argv=None):
    opts = Options()
    case = opts.getString("case", None)
    log = opts.getString("log", "INFO")
    runTests = opts.getString("test", None)
    logging.basicConfig(level=log.upper())
    logger = logging.getLogger("Init")
    if not case:
        logger.error("usage: python -m pynama_b200.run_case -case <name|file.yaml> [...]")
        return 2
    try:
        config = loadCase(case)
    except OSError:
        logger.error(f"Case '{case}' Not Found")
        return 2
    if runTests:
        from .cases.base_problem import BaseProblemTest as FemProblem
    elif config.get("name") in ("ibm-static", "ibm-dynamic") or case in ("ibm-static", "ibm-dynamic"):
        from .cases import immersed_boundary as ib
        dynamic = "dynamic" in (case if case in ("ibm-static", "ibm-dynamic") else config.get("name"))
        FemProblem = ib.ImmersedBoundaryDynamic if dynamic else ib.ImmersedBoundaryStatic
    else:
        from .cases.base_problem import BaseProblem as FemProblem
    if runTests == "kle":
        fem = FemProblem(config, case=case)
        fem.setUp()
        fem.setUpSolver()
        errs = fem.getKLEError(viscousTimes=[0.01, 0.05, 0.1, 0.2, 0.3, 0.4, 0.5])
        logger.info(f"KLE errors: {errs}")
    elif runTests == "operators":
        fem = FemProblem(config, case=case)
        fem.setUp()
        fem.setUpSolver()
        logger.info(f"operator errors (conv, diff, curl): {fem.OperatorsTests()}")
    else:
        timeSolving(config, FemProblem, case=case)
    return 0


if __name__ == "__main__":
    sys.exit(main())
