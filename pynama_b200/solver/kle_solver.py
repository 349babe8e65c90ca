"""KSP wrapper of the KLE solve -- same surface as the reference's
``solver.kle_solver`` (src/solver/kle_solver.py:5-64).

``KleSolver.solve(vort)`` computes ``b = Rw*vort + Krhs*vel`` with two device
SpMVs (the second fused as y = v + A x, no temporaries) and runs the Krylov
solve in place on ``vel`` (kle_solver.py:33-37).  ``KspSolver`` keeps the
reference's defaults (type gmres, pc lu, then the options database wins,
kle_solver.py:54-64); how those are honoured on the GPU is described in
``pynama_b200.petsc.KSP``.
"""
import logging

from ..petsc import COMM_WORLD, KSP, PC, Options


class KleSolver:
    def __init__(self):
        self.logger = logging.getLogger("KLE Solver")

    def setMat(self, mat):
        self.mat = mat

    def setUp(self):
        solveType = self.mat.bcType
        K = self.mat.K
        # -kle_matfree 1: the products with K inside the Krylov loop are evaluated element by
        # element (sum-factorised quadrature) instead of from the stored blocks
        if Options().getBool("kle_matfree", False):
            K.setMatFree(True)
        self.solver = KspSolver()
        self.solver.createSolver(K)
        self.__vel = K.createVecRight()
        self.__vel.setName("velocity")
        self.__rhs = K.createVecLeft()
        self.__isNS = False
        if solveType == "NS":
            self.__KplusKfs = K + self.mat.Kfs           # kle_solver.py:25
            self.solverFS = KspSolver()
            self.solverFS.createSolver(self.__KplusKfs)
            self.__velFS = K.createVecRight()
            self.__velFS.setName("free-slip")
            self.__isNS = True

    def isNS(self):
        return self.__isNS

    def solve(self, vort, vec=None):
        vel = self.__vel if vec is None else vec
        b = self.__rhs
        self.mat.Rw.mult(vort, b)                  # b  = Rw * vort
        self.mat.Krhs.multAdd(vel, b, b)           # b += Krhs * vel
        self.solver(b, vel)

    def solveFS(self, vort):
        b = self.__rhs
        self.mat.Rw.mult(vort, b)
        self.mat.Rwfs.multAdd(vort, b, b)
        self.mat.Krhsfs.multAdd(self.__vel, b, b)
        self.solverFS(b, self.__velFS)

    def getFreeSlipSolution(self):
        return self.__velFS

    def getSolution(self):
        return self.__vel


class KspSolver(KSP):
    comm = COMM_WORLD

    def __init__(self):
        super().__init__()
        self.logger = None

    def createSolver(self, mat):
        self.logger = logging.getLogger("KSP Solver")
        self.logger.debug("setupKSP")
        self.create(self.comm)
        self.setType("gmres")
        pc = PC().create()
        pc.setType("lu")
        self.setPC(pc)
        self.setFromOptions()
        self.setOperators(mat)
        self.setUp()
