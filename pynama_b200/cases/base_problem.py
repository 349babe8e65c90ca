"""Problem orchestration on the GPU path -- same surface as the reference's
``cases.base_problem`` (``BaseProblem``, ``BaseProblemTest``;
src/cases/base_problem.py).  The two hot-path members are

* ``evalRHS``        base_problem.py:111-136  (A10): BC scatter -> [NS pre-solve] ->
  KLE solve -> T = 2 mu SrT v - rho v(x)v -> r = DivSrT T / rho -> f = Curl r
* ``computeVtensV``  base_problem.py:139-154  (A11)

Everything runs on device Vecs; the pointwise pieces are fused into single
kernels where the result is bit-identical (``pyn_stress_combine`` = VtensV +
``*= 2 mu`` + ``axpy(-rho)``; ``scale(1/rho)`` folded into the DivSrT SpMV).

Faithful quirk: ``evalRHS`` ignores the stage vector TS hands it
and always differentiates ``self.vort`` (base_problem.py:115,125).  Opt in to the
fixed behaviour with ``-pynama_rhs_uses_stage 1``.
"""
import importlib
import logging
from math import sqrt

import numpy as np

from .. import _lib
from ..domain.domain import Domain
from ..matrices.mat_fs import MatFS, Operators  # noqa: F401
from ..matrices.mat_ns import MatNS
from ..petsc import COMM_WORLD, Options
from ..solver.kle_solver import KleSolver
from ..solver.ts_solver import TsSolver
from ..viewer.paraviewer import Paraviewer


class Timer:
    def __init__(self):
        import time
        self._time = time
        self.t0 = time.perf_counter()

    def tic(self):
        self.t0 = self._time.perf_counter()

    def toc(self):
        return self._time.perf_counter() - self.t0


class BaseProblem(object):
    def __init__(self, config, **kwargs):
        self.comm = COMM_WORLD
        self.timerTotal = Timer()
        self.timer = Timer()
        self.case = kwargs.get("case", Options().getString("case", "uniform"))
        self.config = config
        self.logger = logging.getLogger(f"[{self.comm.rank}] {self.config.get('name')}")
        self.caseName = self.config.get("name")
        self.readMaterialData()
        self.opts = kwargs
        self.rhsUsesStage = bool(Options().getBool("pynama_rhs_uses_stage", False))
        self.kspIterations = 0
        self.kspSolves = 0
        if "chart" in kwargs:
            self.setUpTimeSolverTest()
        elif "time-solver" in self.config:
            self.setUpTimeSolver()

    def setUp(self):
        self.setUpDomain()
        self.setUpViewer()
        self.createMesh()

    def setUpViewer(self):
        self.viewer = Paraviewer()

    def setUpDomain(self):
        self.dom = Domain()
        self.dom.configure(self.config)
        self.dom.setOptions(**self.opts)
        self.dom.setUp()
        self.dim = self.dom.getDimension()
        self.dim_w = 1 if self.dim == 2 else 3
        self.dim_s = 3 if self.dim == 2 else 6

    def readMaterialData(self):
        materialData = self.config.get("material-properties")
        self.rho = materialData["rho"]
        self.mu = materialData["mu"]
        self.nu = self.mu / self.rho

    def createMesh(self, saveMesh=True):
        self.viewer.configure(self.dim, self.config.get("save-dir"))
        if saveMesh:
            self.viewer.saveMesh(self.dom.getFullCoordArray())
        if not self.comm.rank:
            self.logger.info("Mesh created")

    def setUpTimeSolver(self):
        options = self.config.get("time-solver")
        self.ts = TsSolver(self.comm)
        self.ts.setUpTimes(options["start-time"], options["end-time"], options["max-steps"])
        self.ts.initSolver(self.evalRHS, self.convergedStepFunction)

    def convergedStepFunction(self, ts):
        time, step, incr = ts.time, ts.step_number, ts.getTimeStep()
        vort = ts.getSolution()
        vel = self.solverKLE.getSolution()
        self.viewer.saveData(step, time, vel, vort)
        self.viewer.writeXmf(self.caseName)
        if not self.comm.rank:
            self.logger.info(f"Converged: Step {step:4} | Time {time:.4e} | Increment Time: {incr:.2e} ")

    # ---- hot path (4): the vorticity-transport right-hand side ------------------------------------
    def evalRHS(self, ts, t, vort, f):
        """Evaluate the KLE right hand side."""
        w = vort if self.rhsUsesStage else self.vort
        # boundary values of the NEXT evaluation are computed on a host thread during this solve
        self.dom.prefetchBoundaryConditions(("vorticity", "velocity"), getattr(ts, "next_rhs_time", None), self.nu)
        self.dom.applyBoundaryConditions(w, "vorticity", t, self.nu)
        vel = self.solverKLE.getSolution()
        self.dom.applyBoundaryConditions(vel, "velocity", t, self.nu)

        if self.solverKLE.isNS():
            self.solverKLE.solveFS(w)
            velFS = self.solverKLE.getFreeSlipSolution()
            self.dom.applyBoundaryConditionsNS(velFS, "velocity", t, self.nu)
            self.operator.Curl.mult(velFS, w)

        # the same stage of the last step is the best predictor of this solve's iteration count
        stage = getattr(ts, "stage_index", None)
        hints = self.__dict__.setdefault("_itsOfStage", {})
        if stage in hints and hasattr(self.solverKLE.solver, "setIterationHint"):
            self.solverKLE.solver.setIterationHint(hints[stage])
        self.solverKLE.solve(w)
        hints[stage] = self.solverKLE.solver.getIterationNumber()
        self.kspIterations += self.solverKLE.solver.getIterationNumber()
        self.kspSolves += 1

        # _Aux1 = 2 mu SrT v - rho v(x)v     (base_problem.py:127-130)
        self.operator.SrT.mult(vel, self._Aux1)
        _lib.call("pyn_stress_combine", self._Aux1._h, vel._h, float(self.mu), float(self.rho), self.dim)
        # rhs = DivSrT _Aux1 / rho          (base_problem.py:132-134)
        _lib.call("pyn_mat_mult_axpby", self.operator.DivSrT._h, 1.0 / self.rho, self._Aux1._h, 0.0,
                  None, self._rhsWork._h)
        self.operator.Curl.mult(self._rhsWork, f)

    def computeVtensV(self, vec):
        """packed v (x) v per node: [xx, xy, yy] / [xx, xy, yy, yz, zz, zx]"""
        _lib.call("pyn_vtensv", vec._h, self._VtensV._h, self.dim)

    def setUpSolver(self):
        bcType = self.dom.getBoundaryType()
        if bcType == "FS":
            mat = MatFS()
        elif bcType == "NS":
            mat = MatNS()
        else:
            raise Exception("FSNS Mat not implemented")
        mat.setDomain(self.dom)
        mat.build()
        self.mat = mat
        self.solverKLE = KleSolver()
        self.solverKLE.setMat(mat)
        self.solverKLE.setUp()
        self.operator = mat.getOperators()
        self._VtensV = self.operator.SrT.createVecLeft()
        self._Aux1 = self._VtensV.duplicate()
        self._rhsWork = self.mat.K.createVecRight()
        assert "initial-conditions" in self.config, "Initial conditions not defined"
        self.setUpInitialConditions()

    def setUpInitialConditions(self):
        self.logger.info("Computing initial conditions")
        initTime = self.ts.getTime()
        vort = self.operator.Curl.createVecLeft()
        vort.setName("vorticity")
        vel = self.solverKLE.getSolution()
        ic = self.config["initial-conditions"]
        nodes = np.asarray(self.dom.getAllNodes(), dtype=np.int64)
        inds = (nodes[:, None] * self.dim + np.arange(self.dim)[None, :]).ravel()
        if "custom-func" in ic:
            lib = importlib.import_module(f".{ic['custom-func']['name']}", package="pynama_b200.functions")
            alpha = lib.alpha(self.nu, initTime)
            coords = self.dom.getFullCoordArray()
            arrVel = lib.velocity(coords, alpha)
            arrVort = lib.vorticity(coords, alpha)
            vort.setValues(nodes if self.dim == 2 else inds, arrVort, addv=False)
            vel.setValues(inds, arrVel, addv=False)
        else:
            if "velocity" in ic and "vorticity" not in ic:
                vel.setValues(inds, np.tile(ic["velocity"], len(nodes)))
        vort.assemble()
        vel.assemble()
        self.vort = vort

    def view(self):
        print(f"Case: {self.case}")
        self.dom.view()
        print(f"NGL: {self.dom.getNGL()}")


class BaseProblemTest(BaseProblem):
    """Exact-solution harnesses of the reference (base_problem.py:233-391)."""

    def _funcLib(self):
        name = self.config["tests"]["custom-func"]["name"]
        return importlib.import_module(f".{name}", package="pynama_b200.functions")

    def generateExactVecs(self, vel=None, vort=None, time=None):
        exactVel = self.mat.K.createVecRight()
        exactVort = self.mat.Rw.createVecRight()
        exactVel.setName(f"{self.caseName}-exact-vel")
        exactVort.setName(f"{self.caseName}-exact-vort")
        n = self.dom.getNumOfNodes()
        if vel and vort:
            arrVel, arrVort = np.tile(vel, n), np.tile(vort, n)
        else:
            lib = self._funcLib()
            alpha = lib.alpha(self.nu, time)
            coords = self.dom.getFullCoordArray()
            arrVel, arrVort = lib.velocity(coords, alpha), lib.vorticity(coords, alpha)
        exactVort.setArray(arrVort)
        exactVel.setArray(arrVel)
        return exactVel, exactVort

    def generateExactOperVecs(self, time):
        lib = self._funcLib()
        alpha = lib.alpha(self.nu, time)
        coords = self.dom.getFullCoordArray()
        out = []
        for name, proto in (("velocity", self.mat.K), ("vorticity", self.mat.Rw), ("convective", self.mat.Rw),
                            ("diffusive", self.mat.Rw)):
            v = proto.createVecRight()
            v.setName(name)
            v.setArray(getattr(lib, name)(coords, alpha))
            out.append(v)
        return out

    def getKLEError(self, viscousTimes=None, startTime=0.0, endTime=1.0, steps=10):
        if viscousTimes is None:
            viscousTimes = np.arange(startTime, endTime, (endTime - startTime) / steps)
        errors = []
        for tau in viscousTimes:
            time = (tau ** 2) / (4 * self.nu)
            exactVel, exactVort = self.generateExactVecs(time=time)
            self.dom.applyBoundaryConditions(self.vort, "vorticity", time, self.nu)
            vel = self.solverKLE.getSolution()
            self.dom.applyBoundaryConditions(vel, "velocity", time, self.nu)
            self.solverKLE.solve(exactVort)
            errors.append((exactVel - vel).norm(norm_type=2))
        return errors

    def getConvective(self, exactVel, exactConv):
        convective = exactConv.copy()
        vel = self.solverKLE.getSolution()
        self.computeVtensV(vel)
        aux = vel.copy()
        self.operator.DivSrT.mult(self._VtensV, aux)
        self.operator.Curl.mult(aux, convective)
        return convective

    def getDiffusive(self, exactVel, exactDiff):
        diffusive = exactDiff.copy()
        vel = self.solverKLE.getSolution()
        self.operator.SrT.mult(exactVel, self._Aux1)
        aux = vel.copy()
        self._Aux1 *= (2.0 * self.mu)
        self.operator.DivSrT.mult(self._Aux1, aux)
        aux.scale(1 / self.rho)
        self.operator.Curl.mult(aux, diffusive)
        return diffusive

    def OperatorsTests(self, viscousTime=1):
        time = (viscousTime ** 2) / (4 * self.nu)
        vel = self.solverKLE.getSolution()
        exactVel, exactVort, exactConv, exactDiff = self.generateExactOperVecs(time)
        self.dom.applyBoundaryConditions(vel, "velocity", time, self.nu)
        self.solverKLE.solve(exactVort)
        convective = self.getConvective(exactVel, exactConv)
        diffusive = self.getDiffusive(exactVel, exactDiff)
        self.operator.Curl.mult(exactVel, self.vort)
        wgt = self.operator.weigCurl.copy()
        wgt.reciprocal()                       # lumped mass (base_problem.py:332)
        out = []
        for got, exact in ((convective, exactConv), (diffusive, exactDiff), (self.vort, exactVort)):
            err = got - exact
            out.append(sqrt((err * err).dot(wgt)))
        return tuple(out)

    def setUpTimeSolverTest(self):
        self.setUpTimeSolver()
