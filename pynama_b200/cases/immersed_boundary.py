"""2-D immersed-boundary cases: host mirror of src/cases/immersed_boundary.py
(ImmersedBoundaryStatic :19-298, ImmersedBoundaryDynamic :300-384).

The marker -> Euler-node search, the delta weights and the matrices H, S, A = H S are built by ONE
library call (``pyn_ibm_build``, csrc/ibm.cu) instead of the reference's per-marker Python loops
(collectNodes :210-229, computeClose :263-274 / :346-358, computeDirac :255-262, buildIBMMatrix
:170-193); ``collectedNodes`` is recovered from the rows of H for the callers that read it.

Differences kept on purpose:
* the reference's time loops call ``self.solveKLE``, ``self.bcNodes`` and ``self.cteValue``, which
  its BaseProblem no longer defines (startSolver :56-78, applyBoundaryConditions :155-160 cannot
  run there); here ``solveKLE`` is the boundary-condition + KLE solve pair of evalRHS
  (base_problem.py:115-125);
* bodies' velocities are concatenated instead of nested (immersed_body.py:116-124).
"""
import ctypes

import numpy as np

from .. import _lib
from .._lib import call, ptr, vp, c_i64
from ..petsc import Context, Mat
from ..domain.immersed_body import BodiesContainer
from ..solver.kle_solver import KspSolver
from .base_problem import BaseProblem, Timer


class ImmersedBoundaryStatic(BaseProblem):
    DYNAMIC_SEARCH = 0          # computeClose rule: 0 static (:263-274), 1 dynamic (:346-358)

    def setUp(self):
        super().setUp()
        self.U_ref = self.dom.getFreeStreamVelocity()
        self.setUpBodies()
        cells = self.getAffectedCells(10)
        self.collectedNodes, self.maxNodesPerLag = self.collectNodes(cells)
        self.totalEulerNodes = len(self.dom.getAllNodes())
        self.createEmptyIBMMatrix()
        self.mNodes = self.buildIBMMatrix()

    def setUpDomain(self):
        super().setUpDomain()
        meshType = self.dom.getMeshType()
        ngl = self.dom.getNGL()
        assert ngl < 4, "IBM Only implemented for NGL lower than 4"
        if meshType == "box":
            self.nodeSeparation = self.dom.getNodeSeparationIBM()
        elif meshType == "gmsh":
            self.nodeSeparation = self.config["domain"]["h-min"] / (ngl - 1)
        else:
            raise Exception("Mesh not defined")
        self.h = self.nodeSeparation

    def setUpBodies(self):
        if "bodies" not in self.config:
            raise Exception("Bodies not defined")
        self.body = BodiesContainer(self.config["bodies"])
        self.logger.info(f"Node separation: {self.nodeSeparation}")
        self.body.createBodies(self.nodeSeparation)
        self.body.setVelRef(self.U_ref)

    # ---- search + matrices (device) -------------------------------------------------------------
    def getAffectedCells(self, xSide, ySide=None, center=(0, 0)):
        """cells whose centroid lies inside the box of half sides (xSide, ySide) around center"""
        ySide = ySide if ySide else xSide
        first, last = self.dom.getLocalCellRange()
        d = self.dom.getCellCentroids()[first:last, :2] - np.asarray(center, dtype=np.float64)
        hit = (np.abs(d[:, 0]) < xSide) & (np.abs(d[:, 1]) < ySide)
        return (np.nonzero(hit)[0] + first).tolist()

    def collectNodes(self, cells):
        """Euler nodes of the affected cells; the per-marker search itself runs inside
        buildIBMMatrix.  Returns the lazily filled marker -> nodes mapping and its longest row."""
        self.ibmNodes, self.eulerCoords = self.dom.getNodesCoordsFromEntities(cells)
        self.ibmNodes = np.ascontiguousarray(self.ibmNodes, dtype=np.int32)
        self.totalEulerNodes = len(self.dom.getAllNodes())
        self._assembleOnDevice()
        return self.collectedNodes, self.maxNodesPerLag

    def _assembleOnDevice(self):
        lag = np.ascontiguousarray(self.body.getAllCoordinates(), dtype=np.float64)
        cand = np.ascontiguousarray(self.eulerCoords[:, :2], dtype=np.float64)
        h, dl = float(self.nodeSeparation), float(self.body.getElementLong())
        H, S, A, mx = vp(), vp(), vp(), c_i64()
        mark = np.zeros(self.totalEulerNodes, dtype=np.uint8)
        # H.scale(h^2), S.scale(dl h)  (:185-187)
        call("pyn_ibm_build", Context.get().handle, self.totalEulerNodes, len(lag), ptr(lag), len(cand),
             ptr(self.ibmNodes), ptr(cand), h, dl, h ** 2, dl * h, self.body.getDiracKind(),
             self.DYNAMIC_SEARCH, ctypes.byref(H), ctypes.byref(S), ctypes.byref(A), ctypes.byref(mx),
             ptr(mark))
        for old in ("H", "S", "A"):
            if getattr(self, old, None) is not None:
                getattr(self, old).destroy()
        self.H, self.S, self.A = Mat(H, "ibm-H"), Mat(S, "ibm-S"), Mat(A, "ibm-A")
        self.maxNodesPerLag = mx.value
        self._affected = np.nonzero(mark)[0]
        self.collectedNodes = _CollectedNodes(self)
        for lagNode, n in enumerate(self.collectedNodes.counts()):
            self.body.setEulerNodes(lagNode, int(n))

    def createEmptyIBMMatrix(self):
        """preallocation in the reference (:170-178); the device build sizes H itself"""
        return None

    def buildIBMMatrix(self):
        if getattr(self, "H", None) is None:
            self._assembleOnDevice()
        if getattr(self, "ksp", None) is not None:
            self.ksp.destroy()
        self.ksp = KspSolver()
        self.ksp.createSolver(self.A)
        self.logger.info("IBM Matrices builded")
        return self._affected.tolist()

    def getMarkedNodes(self):
        return np.unique(self.collectedNodes.all_nodes()).tolist()

    def computeDirac(self, lagPoint, eulerCoords):
        xb = self.body.getNodeCoordinates(lagPoint)
        return [self.body.getDiracs(c - xb, self.nodeSeparation) for c in eulerCoords]

    def computeClose(self, lagPoint, eulerCoords):
        dist = (np.asarray(eulerCoords) - self.body.getNodeCoordinates(lagPoint)) / self.body.getElementLong()
        return ((np.abs(dist[:, 0]) < 3) & (np.abs(dist[:, 1]) < 3)).astype(int)

    # ---- solver ---------------------------------------------------------------------------------
    def setUpSolver(self):
        super().setUpSolver()
        self.vel = self.solverKLE.getSolution()
        self.vel_correction = self.vel.copy()
        self.vel_correction.setName("velocity_correction")
        self.virtualFlux = self.S.createVecRight()
        self.virtualFlux.setName("virtual_flux")
        self.ibm_rhs = self.S.createVecRight()

    def solveKLE(self, time, vort):
        """boundary values + KLE solve, as evalRHS does (base_problem.py:115-125)"""
        self.dom.applyBoundaryConditions(vort, "vorticity", time, self.nu)
        self.dom.applyBoundaryConditions(self.vel, "velocity", time, self.nu)
        self.solverKLE.solve(vort)

    def computeVelocityCorrection(self, **kwargs):
        """vel += S q with (H S) q = -(H vel - v_body)  (:162-168)"""
        bodyVel = self.body.getVelocity()
        self.H.mult(self.vel, self.ibm_rhs)
        self.ibm_rhs.axpy(-1, bodyVel)
        self.ksp.solve(-self.ibm_rhs, self.virtualFlux)
        self.S.mult(self.virtualFlux, self.vel_correction)
        self.vel += self.vel_correction

    def computeDragForce(self, dt):
        denom = 0.5 * (self.U_ref ** 2)
        return self.body.computeForce(self.virtualFlux, denom * dt)

    def computeInitialCondition(self):
        startTime = self.ts.getTime()
        self.vort.set(0.0)
        self.ibmZone = self.vort.copy()
        self.ibmZone.setName("ibm-zone")
        self.affectedNodes = self.ibmZone.copy()
        self.affectedNodes.setName("affected-nodes")
        self.body.setVelRef(self.U_ref)
        self.solveKLE(startTime, self.vort)
        self.computeVelocityCorrection(t=startTime)
        self.operator.Curl.mult(self.vel, self.vort)

    def markAffectedNodes(self, nodes):
        self.affectedNodes.set(0.0)
        self.affectedNodes.setValues(nodes, [1] * len(nodes), addv=False)

    def markZone(self, nodes):
        self.ibmZone.setValues(nodes, [1] * len(nodes), addv=False)

    def _afterStep(self, i):
        """correction + new vorticity after one TS step (body of the loops at :70-81, :322-331)"""
        time = self.ts.time
        self.solveKLE(time, self.ts.getSolution())
        self.computeVelocityCorrection(t=time)
        self.operator.Curl.mult(self.vel, self.vort)
        self.ts.setSolution(self.vort)
        return time

    def startSolver(self):
        self.computeInitialCondition()
        self.ts.setSolution(self.vort)
        saveSteps = self.config["save-n-steps"]
        self.body.updateVelocity()
        self.markAffectedNodes(self.mNodes)
        self.markZone(self.getMarkedNodes())
        for i in range(self.ts.getMaxSteps() + 1):
            self.ts.step()
            time = self._afterStep(i)
            step, dt = self.ts.getStepNumber(), self.ts.getTimeStep()
            self.logger.info(f"Converged: Step {step:4} | Time {time:.4e} | DT: {dt:.4e}  ")
            if time > self.ts.getMaxTime():
                break
            if i % saveSteps == 0:
                self.viewer.saveData(step, time, self.vort, self.vel, self.ibmZone, self.affectedNodes)
                self.viewer.writeXmf(self.caseName)
                self.H.mult(self.vel, self.ibm_rhs)


class ImmersedBoundaryDynamic(ImmersedBoundaryStatic):
    DYNAMIC_SEARCH = 1

    def computeClose(self, lagPoint, eulerCoords):
        dl = self.body.getElementLong()
        dist = (np.asarray(eulerCoords) - self.body.getNodeCoordinates(lagPoint)) / dl
        return ((np.abs(dist[:, 0]) < 2) & (np.abs(dist[:, 1] * dl) < (1 + 2 * dl))).astype(int)

    def computeVelocityCorrection(self, t=None, **kwargs):
        """the body moves: new centre and velocity, H/S/A rebuilt, then the correction (:361-372)"""
        self.body.updateBodyParameters(t)
        affNodes = self.rebuildMatrix()
        super().computeVelocityCorrection()
        return affNodes

    def rebuildMatrix(self):
        self._assembleOnDevice()
        return self.buildIBMMatrix()

    def _afterStep(self, i):
        time = self.ts.time
        nods = self.computeVelocityCorrection(t=time)
        self.markAffectedNodes(nods)
        self.operator.Curl.mult(self.vel, self.vort)
        self.ts.setSolution(self.vort)
        return time

    def startSolver(self):
        self.computeInitialCondition()
        self.ts.setSolution(self.vort)
        saveSteps = self.config["save-n-steps"]
        self.markZone(self.getMarkedNodes())
        timer, record = Timer(), {k: [] for k in ("times", "cd", "cl", "dt", "steps", "elapsedTimes")}
        for step in range(1, self.ts.getMaxSteps() + 1):
            timer.tic()
            self.ts.step()
            time = self._afterStep(step)
            dt = self.ts.getTimeStep()
            self.logger.info(f"Converged: Step {step:4} | Time {time:.4e} | Increment {dt:.3e}")
            if time > self.ts.getMaxTime():
                break
            if step % saveSteps == 0:
                self.viewer.saveData(step, time, self.vort, self.vel, self.ibmZone, self.affectedNodes)
                self.viewer.writeXmf(self.caseName)
                self.H.mult(self.vel, self.ibm_rhs)
            if step % max(int(saveSteps / 10), 1) == 0:
                cd, cl = self.computeDragForce(dt)
                for key, val in (("times", time), ("cd", cd), ("cl", cl), ("dt", dt), ("steps", step),
                                 ("elapsedTimes", timer.toc())):
                    record[key].append(val)
        self.forceHistory = dict(record, dh=self.h, dl=self.body.getElementLong(),
                                 lagPoints=self.body.getTotalNodes(), eulerNodes=self.vort.getSizes()[0],
                                 ngl=self.dom.getNGL())


class _CollectedNodes:
    """marker -> {'nodes', 'coords', 'indices'} view on the rows of H (the reference's dict, :214-221)"""

    def __init__(self, case):
        self._case = case
        dim = case.dim
        ip, ix, _ = case.H.getValuesCSR()
        self._ptr = ip[::dim]                       # start of row l*dim of marker l
        self._len = (ip[1::dim] - ip[::dim][: len(ip[1::dim])])
        self._nodes = ix // dim

    def counts(self):
        return self._len[: len(self)]

    def all_nodes(self):
        return np.concatenate([self[l]["nodes"] for l in range(len(self))]) if len(self) else np.zeros(0, int)

    def __len__(self):
        return self._case.body.getTotalNodes()

    def keys(self):
        return range(len(self))

    def __getitem__(self, lagNode):
        dim = self._case.dim
        a = int(self._ptr[lagNode])
        nodes = self._nodes[a: a + int(self._len[lagNode])]
        return {"nodes": nodes, "coords": self._case.dom.getNodesCoordinates(nodes),
                "indices": [int(n) * dim + d for n in nodes for d in range(dim)]}
