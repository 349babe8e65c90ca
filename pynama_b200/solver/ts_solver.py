"""Time stepper of the vorticity-transport equation -- same surface as the
reference's ``solver.ts_solver.TsSolver`` (src/solver/ts_solver.py:3-23): explicit
RK '5bs', MATCHSTEP, an RHS and a post-step callback, options from the options
database."""
from ..petsc import COMM_WORLD
from ..ts import TS


class TsSolver(TS):
    rk_types = ["3", "5f", "5bs"]

    def __init__(self, comm=COMM_WORLD):
        super().__init__(comm)
        self.create(comm)
        self.setProblemType(self.ProblemType.NONLINEAR)
        self.setEquationType(self.EquationType.ODE_EXPLICIT)
        self.setType("rk")
        self.setRKType("5bs")

    def setUpTimes(self, sTime, eTime, steps):
        self.setTime(sTime)
        self.setMaxTime(eTime)
        self.setMaxSteps(steps)
        self.setExactFinalTime(self.ExactFinalTime.MATCHSTEP)

    def initSolver(self, rhsFunction, convergedStepFunction):
        self.setRHSFunction(rhsFunction)
        self.setPostStep(convergedStepFunction)
        self.setFromOptions()
