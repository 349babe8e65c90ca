"""KLE velocity solve: right-hand side products + the Krylov solve, on the device.

Drop-in for the reference's ``solver.kle_solver`` (src/solver/kle_solver.py:5-64).
``KleSolver.solve(vort)`` forms ``b = Rw*vort + Krhs*vel`` with two SpMVs into a persistent work
vector (the second one fused as ``y = b + A x``; the reference allocates three temporaries per call,
kle_solver.py:33-37) and runs the Krylov solve in place on ``vel``.  For no-slip cases a second
solver on ``K + Kfs`` produces the free-slip velocity first (kle_solver.py:22-28,39-41).
``KspSolver`` keeps the reference's defaults -- type gmres, pc lu, then the options database wins
(kle_solver.py:54-64); how those are honoured on the GPU is described in ``pynama_b200.petsc.KSP``.

Options read here: ``-kle_matfree 1`` evaluates the products with K inside the Krylov loop element
by element (sum-factorised quadrature) instead of from the stored blocks.
"""
import logging

from ..petsc import COMM_WORLD, KSP, PC, Options


class KspSolver(KSP):
    """KSP preconfigured like the reference's wrapper; ``ksp(b, x)`` solves"""
    comm = COMM_WORLD
    DEFAULT_TYPE, DEFAULT_PC = "gmres", "lu"

    def __init__(self):
        super().__init__()
        self.logger = None

    def createSolver(self, mat):
        self.logger = logging.getLogger("KSP Solver")
        self.create(self.comm)
        self.setType(self.DEFAULT_TYPE)
        precond = PC().create()
        precond.setType(self.DEFAULT_PC)
        self.setPC(precond)
        self.setFromOptions()                 # -ksp_type / -pc_type / tolerances / Fischer guess ...
        self.setOperators(mat)
        self.setUp()
        self.logger.debug("KSP ready: %s", self.getType())


class KleSolver:
    def __init__(self):
        self.logger = logging.getLogger("KLE Solver")
        self.mat = None
        self.solver = self.solverFS = None
        self._vel = self._velFS = self._rhs = self._opFS = None

    def setMat(self, mat):
        self.mat = mat

    def _newSolver(self, operator):
        ksp = KspSolver()
        ksp.createSolver(operator)
        return ksp

    def setUp(self):
        K = self.mat.K
        if Options().getBool("kle_matfree", False):
            K.setMatFree(True)
        self.solver = self._newSolver(K)
        self._rhs = K.createVecLeft()
        self._vel = K.createVecRight()
        self._vel.setName("velocity")
        if self.isNS():
            self._opFS = K + self.mat.Kfs              # kle_solver.py:25
            self.solverFS = self._newSolver(self._opFS)
            self._velFS = K.createVecRight()
            self._velFS.setName("free-slip")

    def isNS(self):
        return self.mat.bcType == "NS"

    def _assembleRhs(self, terms):
        """self._rhs = sum of A*x over terms [(A, x), ...] without temporaries"""
        (A0, x0), rest = terms[0], terms[1:]
        A0.mult(x0, self._rhs)
        for A, x in rest:
            A.multAdd(x, self._rhs, self._rhs)
        return self._rhs

    def solve(self, vort, vec=None):
        """velocity from vorticity; ``vec`` (default: the solver's own velocity Vec) carries the
        boundary values in and the solution out"""
        vel = self._vel if vec is None else vec
        self.solver(self._assembleRhs([(self.mat.Rw, vort), (self.mat.Krhs, vel)]), vel)

    def solveFS(self, vort):
        """free-slip pre-solve of the no-slip cases: tangential wall velocities are unknowns"""
        m = self.mat
        b = self._assembleRhs([(m.Rw, vort), (m.Rwfs, vort), (m.Krhsfs, self._vel)])
        self.solverFS(b, self._velFS)

    def getSolution(self):
        return self._vel

    def getFreeSlipSolution(self):
        return self._velFS
