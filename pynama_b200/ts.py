"""Explicit Runge-Kutta time stepper on device Vecs -- the slice of PETSc ``TS``
the reference drives (src/solver/ts_solver.py:3-23, cases/base_problem.py:73-103).

Only what the reference configures exists: type ``rk`` with the Bogacki-Shampine
5(4) pair ``5bs`` (8 stages, FSAL), explicit ODE, MATCHSTEP final time, an RHS
callback ``rhs(ts, t, u, F)`` and a post-step callback ``post(ts)``.  Stage
combinations are single-pass multi-axpy kernels (k11); the step controller is
host-side scalar logic.

PETSc's own sources are not in the reference tree, so two behaviours are
restated from its documentation and marked UNPINNED in DESIGN.md: the default
initial step (0.1) and the ``basic`` adaptive controller (safety 0.9, clip
[0.1, 10], rtol = atol = 1e-4, 2-norm).  Parity runs pin the step with
``-ts_adapt_type none -ts_dt h`` on both sides.
"""
import math

import numpy as np

from .petsc import COMM_WORLD, Options, Vec

# Bogacki-Shampine 5(4), 8 stages (checked against the order
# conditions there)
_A = [
    [],
    [1 / 6],
    [2 / 27, 4 / 27],
    [183 / 1372, -162 / 343, 1053 / 1372],
    [68 / 297, -4 / 11, 42 / 143, 1960 / 3861],
    [597 / 22528, 81 / 352, 63099 / 585728, 58653 / 366080, 4617 / 20480],
    [174197 / 959244, -30942 / 79937, 8152137 / 19744439, 666106 / 1039181, -29421 / 29068,
     482048 / 414219],
    [587 / 8064, 0, 4440339 / 15491840, 24353 / 124800, 387 / 44800, 2152 / 5985, 7267 / 94080],
]
_B = [587 / 8064, 0, 4440339 / 15491840, 24353 / 124800, 387 / 44800, 2152 / 5985, 7267 / 94080, 0]
_BE = [2479 / 34992, 0, 123 / 416, 612941 / 3411720, 43 / 1440, 2272 / 6561, 79937 / 1113912,
       3293 / 556956]
_C = [0, 1 / 6, 2 / 9, 3 / 7, 2 / 3, 3 / 4, 1, 1]

TABLEAUX = {"5bs": {"A": _A, "b": _B, "be": _BE, "c": _C, "order": 5, "fsal": True}}
# classical pairs PETSc also offers under the names TsSolver lists (ts_solver.py:4)
TABLEAUX["3"] = {"A": [[], [1 / 2], [-1, 2]], "b": [1 / 6, 2 / 3, 1 / 6], "be": None,
                 "c": [0, 1 / 2, 1], "order": 3, "fsal": False}
# Fehlberg 5(4), six stages (PETSc TSRK5F propagates the fifth-order solution)
TABLEAUX["5f"] = {"A": [[], [1 / 4], [3 / 32, 9 / 32], [1932 / 2197, -7200 / 2197, 7296 / 2197],
                        [439 / 216, -8, 3680 / 513, -845 / 4104],
                        [-8 / 27, 2, -3544 / 2565, 1859 / 4104, -11 / 40]],
                  "b": [16 / 135, 0, 6656 / 12825, 28561 / 56430, -9 / 50, 2 / 55],
                  "be": [25 / 216, 0, 1408 / 2565, 2197 / 4104, -1 / 5, 0],
                  "c": [0, 1 / 4, 3 / 8, 12 / 13, 1, 1 / 2], "order": 5, "fsal": False}


class TS:
    class ProblemType:
        LINEAR, NONLINEAR = 0, 1

    class EquationType:
        ODE_EXPLICIT = 1

    class ExactFinalTime:
        UNSPECIFIED, STEPOVER, INTERPOLATE, MATCHSTEP = 0, 1, 2, 3

    def __init__(self, comm=None):
        self._init_fields()

    def _init_fields(self):
        self._type, self._rk = "rk", "3"
        self.time, self.step_number = 0.0, 0
        self._max_time, self._max_steps = 5.0, 5000
        self._dt = 0.1                      # TS default initial step [unpinned]
        self._adapt = "basic"
        self._rtol = self._atol = 1e-4
        self._rhs = self._post = None
        self._u = None
        self._eft = TS.ExactFinalTime.UNSPECIFIED
        self._work = None
        self._fsal_valid = False
        self.rejections = 0
        self.rhs_evals = 0
        self.next_rhs_time = None
        self.stage_index = 0

    def create(self, comm=None):
        self._init_fields()
        return self

    # -- configuration
    def setProblemType(self, t):
        self._ptype = t

    def setEquationType(self, t):
        self._etype = t

    def setType(self, t):
        if t != "rk":
            raise ValueError(f"TS type '{t}' is not part of this path (rk only)")
        self._type = t

    def getType(self):
        return self._type

    def setRKType(self, t):
        if t not in TABLEAUX:
            raise ValueError(f"RK type '{t}' is not available (have {sorted(TABLEAUX)})")
        self._rk = t

    def getRKType(self):
        return self._rk

    def setTime(self, t):
        self.time = float(t)

    def getTime(self):
        return self.time

    def setMaxTime(self, t):
        self._max_time = float(t)

    def getMaxTime(self):
        return self._max_time

    def setMaxSteps(self, n):
        self._max_steps = int(n)

    def getMaxSteps(self):
        return self._max_steps

    def setTimeStep(self, dt):
        self._dt = float(dt)

    def getTimeStep(self):
        return self._dt

    def getStepNumber(self):
        return self.step_number

    def setExactFinalTime(self, eft):
        self._eft = eft

    def setTolerances(self, rtol=None, atol=None):
        if rtol is not None:
            self._rtol = rtol
        if atol is not None:
            self._atol = atol

    def setRHSFunction(self, fn, f=None):
        self._rhs = fn

    def setPostStep(self, fn):
        self._post = fn

    def setSolution(self, u):
        """TSSetSolution.  The cases call it after changing the state between steps (the immersed-
        boundary loops correct the velocity and recompute the vorticity, immersed_boundary.py:58-82),
        so the stage carried over from the last step (FSAL) no longer belongs to the solution and is
        evaluated again.  [PETSc's own rule for this is not in the reference tree: unpinned]"""
        self._u = u
        self._fsal_valid = False

    def getSolution(self):
        return self._u

    def setFromOptions(self):
        o = Options()
        self._dt = o.getReal("ts_dt", self._dt)
        self._adapt = o.getString("ts_adapt_type", self._adapt)
        self._max_steps = o.getInt("ts_max_steps", self._max_steps)
        self._max_time = o.getReal("ts_max_time", self._max_time)
        self._rtol = o.getReal("ts_rtol", self._rtol)
        self._atol = o.getReal("ts_atol", self._atol)
        rk = o.getString("ts_rk_type", None)
        if rk:
            self.setRKType(rk)

    # -- stepping
    def _ensure_work(self, u):
        tab = TABLEAUX[self._rk]
        s = len(tab["b"])
        if self._work is None or self._work["n"] != u.getSize() or len(self._work["k"]) != s:
            self._work = {"n": u.getSize(), "k": [u.duplicate() for _ in range(s)],
                          "y": u.duplicate(), "err": u.duplicate()}
            self._fsal_valid = False
        return self._work

    def _wlte(self, u_old, u_new, err):
        """weighted 2-norm of the local error estimate (TSErrorWeightedNorm, NORM_2)"""
        a, b, e = u_old.getArray(), u_new.getArray(), err.getArray()         # owned entries
        tol = self._atol + self._rtol * np.maximum(np.abs(a), np.abs(b))
        part = [float(np.sum((e / tol) ** 2)), float(len(e))]
        if COMM_WORLD.size > 1:
            # every rank must see the same norm, or the ranks accept / reject differently and the
            # collectives of the next RHS evaluation no longer match; summed in rank order on the
            # host, so the bits agree on all ranks
            parts = COMM_WORLD.allgather(part)
            part = [sum(q[0] for q in parts), sum(q[1] for q in parts)]
        return float(np.sqrt(part[0] / part[1])) if part[1] > 0 else 0.0

    def step(self):
        """one accepted step of size <= dt (retries with a smaller step on rejection)"""
        u = self._u
        tab = TABLEAUX[self._rk]
        A, b, be, c = tab["A"], tab["b"], tab["be"], tab["c"]
        s = len(b)
        w = self._ensure_work(u)
        k, y, err = w["k"], w["y"], w["err"]
        adaptive = self._adapt != "none" and be is not None
        while True:
            h = self._dt
            if self._eft == TS.ExactFinalTime.MATCHSTEP and self.time + h > self._max_time * (1 + 1e-14):
                h = self._max_time - self.time
            t = self.time
            for i in range(s):
                if i == 0 and tab["fsal"] and self._fsal_valid:
                    continue                                   # k[0] carried over from the last step
                if i == 0:
                    stage = u
                else:
                    u.copy(y)
                    coef = [h * a for a in A[i]]
                    nz = [(cf, k[j]) for j, cf in enumerate(coef) if cf != 0.0]
                    y.maxpy([cf for cf, _ in nz], [v for _, v in nz])
                    stage = y
                # time of the next RHS evaluation (next stage, or the first new stage of the next
                # step), so that the callback can prepare time-dependent host data ahead
                if i + 1 < s:
                    self.next_rhs_time = t + c[i + 1] * h
                else:
                    self.next_rhs_time = t + h + c[1 if tab["fsal"] else 0] * self._dt
                self.stage_index = i
                self._rhs(self, t + c[i] * h, stage, k[i])
                self.rhs_evals += 1
            # u_new = u + h sum b_j k_j  (into y, so a rejected step leaves u untouched)
            u.copy(y)
            nz = [(h * bj, k[j]) for j, bj in enumerate(b) if bj != 0.0]
            y.maxpy([cf for cf, _ in nz], [v for _, v in nz])
            accept, hnew = True, h
            if adaptive:
                err.set(0.0)
                nz = [(h * (b[j] - be[j]), k[j]) for j in range(s) if b[j] != be[j]]
                err.maxpy([cf for cf, _ in nz], [v for _, v in nz])
                enorm = self._wlte(u, y, err)
                order = tab["order"] - 1
                safety, clip_lo, clip_hi = 0.9, 0.1, 10.0
                if enorm > 1.0:
                    accept = False
                    safety *= 0.5                               # reject_safety
                fac = clip_hi if enorm == 0.0 else safety * enorm ** (-1.0 / order)
                hnew = h * min(max(fac, clip_lo), clip_hi)
            if accept:
                y.copy(u)
                if tab["fsal"]:
                    k[s - 1].copy(k[0])                       # first stage of the next step
                    self._fsal_valid = True
                self.time = t + h
                self.step_number += 1
                if adaptive:
                    self._dt = hnew
                self._last_dt = h
                return
            self.rejections += 1
            self._dt = hnew
            self._fsal_valid = tab["fsal"]                   # k[0] is still f(t, u)

    def solve(self, u):
        self._u = u
        self._fsal_valid = False
        while self.step_number < self._max_steps and self.time < self._max_time * (1 - 1e-14):
            self.step()
            if self._post is not None:
                self._post(self)
        return u

    def destroy(self):
        self._work = None
