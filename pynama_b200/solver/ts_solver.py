"""Time stepper of the vorticity-transport equation.

Drop-in for the reference's ``solver.ts_solver.TsSolver`` (src/solver/ts_solver.py:3-23): the
cases build it with a communicator, hand it the time window (``setUpTimes``) and the two callbacks
(``initSolver``) and then drive ``solve`` / ``step``.  The integrator itself is the host controller
of ``pynama_b200/ts.py`` (explicit Runge-Kutta, Bogacki-Shampine 5(4) by default, device-resident
stage vectors); this class only fixes the configuration the reference fixes.
"""
from ..petsc import COMM_WORLD
from ..ts import TS

# what the reference selects before the options database gets its say
_DEFAULTS = {"problem": "NONLINEAR", "equation": "ODE_EXPLICIT", "type": "rk", "rk": "5bs",
             "final_time": "MATCHSTEP"}


class TsSolver(TS):
    rk_types = ["3", "5f", "5bs"]          # RK schemes the cases may ask for (-ts_rk_type)

    def __init__(self, comm=COMM_WORLD):
        super().__init__(comm)
        self.create(comm)
        self._configure(_DEFAULTS)

    def _configure(self, cfg):
        self.setProblemType(getattr(self.ProblemType, cfg["problem"]))
        self.setEquationType(getattr(self.EquationType, cfg["equation"]))
        self.setType(cfg["type"])
        if cfg["rk"] not in self.rk_types:
            raise ValueError(f"RK type {cfg['rk']!r} is not one of {self.rk_types}")
        self.setRKType(cfg["rk"])
        self._final_time_rule = getattr(self.ExactFinalTime, cfg["final_time"])

    def setUpTimes(self, sTime, eTime, steps):
        """integration window [sTime, eTime] and the step budget; the last step is shortened to
        land on eTime (the reference's MATCHSTEP)"""
        for setter, value in ((self.setTime, sTime), (self.setMaxTime, eTime), (self.setMaxSteps, steps)):
            setter(value)
        self.setExactFinalTime(self._final_time_rule)

    def initSolver(self, rhsFunction, convergedStepFunction):
        """rhsFunction(ts, t, u, f) fills f = du/dt; convergedStepFunction(ts) runs after every
        accepted step; -ts_* options are read last, so they win"""
        self.setRHSFunction(rhsFunction)
        self.setPostStep(convergedStepFunction)
        self.setFromOptions()
