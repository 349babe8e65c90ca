#!/usr/bin/env python
"""Generate tests/golden/refpath_*.npz: outputs of the REFERENCE'S OWN hot-path code run in this container.

    python tests/golden/make_golden_refpath.py            (needs /root/reference; build container only)

The reference's modules are imported UNMODIFIED (tests/refshim/loader.py) over a numpy stand-in for
petsc4py (tests/refshim/petsc_standin.py) and a box-mesh stand-in for PETSc's DMPlex
(tests/refshim/dm_standin.py).  Everything recorded below is computed by the reference's code:

    Domain.setUp, BoundaryConditions / Boundary / FunctionBoundary          domain/domain.py, boundaries/*.py
    Spectral.getElemKLEMatrices / getElemKLEOperators                       domain/elements/spectral.py
    MatFS.build / MatNS.build (pre-allocation, buildFS / buildNS, buildOperators, lumped-mass scaling)
                                                                            matrices/mat_fs.py, mat_ns.py
    KleSolver.setUp / solve / solveFS                                       solver/kle_solver.py
    BaseProblem.setUpSolver, setUpInitialConditions, evalRHS, computeVtensV cases/base_problem.py
    the case files src/cases/{uniform,taylor-green,taylor-green3d,cavity-2d,cavity}.yaml with smaller boxes

What the stand-ins supply, and what is therefore NOT pinned by these vectors: PETSc's AIJ zero-retention rule and
MatAXPY pattern rule (restated from the manual pages), the direct solve (scipy LU for PETSc's LU), the global node
numbering of the mesh (lexicographic), and the Runge-Kutta stage loop that calls evalRHS in the `steps` records
(Bogacki-Shampine tableau, fixed step; PETSc's TS is not available).

Per case the file holds: the mesh tables the run used (so the test rebuilds nothing from the reference), every
matrix as CSR (explicit zeros included), the boundary sides in case-file order with their node lists and wall
velocities, vort0 / vel0 (initial conditions), vel1 (velocity after the first KLE solve at t0 with boundary values
applied), rhs0 = evalRHS(t0), and vort / vel after `nsteps` fixed steps of size dt.
"""
import logging
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from tests.refshim.loader import load_case, reference_modules  # noqa: E402

CASES = {
    # name: (case file, box overrides, ngl, dt, steps)
    # BASELINE.json configs[0]: 2-D Taylor-Green vortex, 4x4 spectral elements, ngl = 3 (the reference's correctness run)
    "config0": ("taylor-green", dict(lower=[0, 0], upper=[1, 1], nelem=[4, 4]), 3, 2e-3, 3),
    "uniform2d": ("uniform", dict(lower=[0, 0], upper=[1, 1], nelem=[3, 3]), 3, 1e-2, 2),
    "tg2d": ("taylor-green", dict(lower=[0, 0], upper=[1, 1], nelem=[4, 3]), 4, 2e-3, 2),
    "tg3d": ("taylor-green3d", dict(lower=[0, 0, 0], upper=[1, 1, 1], nelem=[2, 3, 2]), 3, 2e-3, 2),
    "cavity2d": ("cavity-2d", dict(lower=[0, 0], upper=[1, 1], nelem=[4, 3]), 3, 5e-3, 2),
    "cavity3d": ("cavity", dict(lower=[0, 0, 0], upper=[3, 1, 1], nelem=[3, 2, 2]), 3, 5e-3, 2),
    # non-affine cells (interior vertices displaced): no two element matrices are equal, so a wrong cell -> corner or
    # cell -> node association cannot hide behind identical blocks
    "tg2d_p": ("taylor-green", dict(lower=[0, 0], upper=[1, 1], nelem=[3, 2], perturb=0.15), 5, 2e-3, 2),
    "tg3d_p": ("taylor-green3d", dict(lower=[0, 0, 0], upper=[1, 1, 1], nelem=[2, 2, 2], perturb=0.2), 3, 2e-3, 2),
    "cavity3d_p": ("cavity", dict(lower=[0, 0, 0], upper=[3, 1, 1], nelem=[3, 3, 2], perturb=0.2), 2, 5e-3, 2),
    "tg3d_ngl4": ("taylor-green3d", dict(lower=[0, 0, 0], upper=[2, 1, 1], nelem=[2, 1, 1]), 4, 2e-3, 2),
}


def csr_fields(prefix, mat):
    m = mat.tocsr()
    return {f"{prefix}_indptr": m.indptr.astype(np.int64), f"{prefix}_indices": m.indices.astype(np.int32),
            f"{prefix}_data": m.data.astype(np.float64), f"{prefix}_shape": np.array(m.shape, dtype=np.int64)}


def run_case(ref, name):
    case, box, ngl, dt, nsteps = CASES[name]
    cfg = load_case(case)
    cfg["domain"]["ngl"] = ngl
    cfg["domain"]["box-mesh"].update(box)
    fem = ref.base_problem.BaseProblem(cfg, case=case)
    fem.setUp()
    fem.setUpSolver()
    out = {"box_lower": np.array(box["lower"], dtype=np.float64), "box_upper": np.array(box["upper"], dtype=np.float64),
           "box_nelem": np.array(box["nelem"], dtype=np.int64),
           "perturb": np.float64(box.get("perturb", 0.0)), "ngl": np.int64(ngl), "dim": np.int64(fem.dim), "dt": np.float64(dt), "nsteps": np.int64(nsteps),
           "rho": np.float64(fem.rho), "mu": np.float64(fem.mu), "bc_type": np.array(fem.dom.getBoundaryType()),
           "case": np.array(case)}
    dm = fem.dom.getDM()
    mesh = dm._m
    out.update(corner_xyz=mesh.corner_xyz, elem_nodes=mesh.elem_nodes.astype(np.int32), coords=mesh.coords,
               coords_reference=fem.dom.getFullCoordArray().copy())
    # boundary sides in the order the reference's BoundaryConditions holds them (= case-file order)
    bcs = fem.dom._Domain__bc
    sides = bcs.getNames()
    out["sides"] = np.array(sides)
    for s in sides:
        b = bcs._BoundaryConditions__ByName[s]
        out[f"side_nodes_{s}"] = np.asarray(b.getNodes(), dtype=np.int64)
        if hasattr(b, "velocity"):
            out[f"side_vel_{s}"] = np.asarray(b.velocity, dtype=np.float64)
    # matrices
    for m in fem.mat.kle:
        assert m.prealloc_exceeded == [], (name, m.getName(), m.prealloc_exceeded[:5])
        out.update(csr_fields(m.getName(), m))
    for nm in ("Curl", "SrT", "DivSrT"):
        m = getattr(fem.operator, nm)
        assert m.prealloc_exceeded == [], (name, nm)
        out.update(csr_fields(nm, m))
    if fem.solverKLE.isNS():
        out.update(csr_fields("KplusKfs", fem.solverKLE.solverFS.getOperators()[0]))
        out["tang_dofs"] = np.array(sorted(fem.dom.getTangDofs(collect=True)), dtype=np.int64)
        out["normal_dofs"] = np.array(sorted(fem.dom.getNormalDofs(collect=True)), dtype=np.int64)
    # initial conditions, first solve, first right-hand side
    t0 = fem.ts.getTime()
    out["t0"] = np.float64(t0)
    out["vort0"] = fem.vort.getArray().copy()
    out["vel0"] = fem.solverKLE.getSolution().getArray().copy()
    f = fem.vort.duplicate()
    fem.evalRHS(fem.ts, t0, fem.vort, f)
    out["rhs0"] = f.getArray().copy()
    out["vel1"] = fem.solverKLE.getSolution().getArray().copy()
    out["vort_after_rhs0"] = fem.vort.getArray().copy()          # evalRHS writes boundary values (NS: Curl velFS) into it
    if fem.solverKLE.isNS():
        out["velfs1"] = fem.solverKLE.getFreeSlipSolution().getArray().copy()
    # a fresh problem for the time loop (evalRHS above has side effects on vort / vel)
    fem2 = ref.base_problem.BaseProblem(cfg, case=case)
    fem2.setUp()
    fem2.setUpSolver()
    fem2.ts.setTimeStep(dt)
    fem2.ts.setMaxSteps(nsteps)
    fem2.ts.solve(fem2.vort)                                     # run_case.py:152-162
    assert fem2.ts.getStepNumber() == nsteps
    out["vort_steps"] = fem2.vort.getArray().copy()
    out["vel_steps"] = fem2.solverKLE.getSolution().getArray().copy()
    out["t_steps"] = np.float64(fem2.ts.getTime())
    return out


def main():
    logging.disable(logging.CRITICAL)
    import contextlib
    import io
    for name in CASES:
        with reference_modules() as ref, contextlib.redirect_stdout(io.StringIO()):
            out = run_case(ref, name)
        path = os.path.join(HERE, f"refpath_{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{path}: {os.path.getsize(path) / 1024:.0f} KiB, {len(out)} arrays, "
              f"dofs {out['coords'].shape[0] * int(out['dim'])}, |rhs0|max {np.abs(out['rhs0']).max():.3e}, "
              f"|vort_steps|max {np.abs(out['vort_steps']).max():.3e}")


if __name__ == "__main__":
    main()
