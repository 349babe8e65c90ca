"""The RK time stepper (pynama_b200/ts.py behind solver/ts_solver.py: SURVEY 8a-A13, 8f-F3) on CPU.
ts.py only calls a handful of Vec methods (getSize, duplicate, copy, set, maxpy, getArray), so a
numpy stand-in runs it without a GPU: the reference's own src/tests/test_ts.py, the order of
accuracy of the 5bs pair, FSAL bookkeeping, MATCHSTEP and the `basic` step controller."""
import math

import numpy as np
import pytest

from pynama_b200.petsc import COMM_WORLD, Options
from pynama_b200.solver.ts_solver import TsSolver
from pynama_b200.ts import TS


class NpVec:
    """the Vec surface ts.py uses, on a numpy array"""

    def __init__(self, a):
        self.a = np.array(a, dtype=np.float64)

    def getSize(self):
        return self.a.size

    def duplicate(self):
        return NpVec(np.zeros_like(self.a))

    def copy(self, dst=None):
        if dst is None:
            return NpVec(self.a.copy())
        dst.a[:] = self.a
        return dst

    def set(self, v):
        self.a[:] = v

    def maxpy(self, coefs, vecs):
        for c, v in zip(coefs, vecs):
            self.a += c * v.a

    def getArray(self):
        return self.a


@pytest.fixture
def opts():
    o = Options()
    touched = []

    def setv(name, val):
        o.setValue(name, val)
        touched.append(name)
    yield setv
    for name in touched:
        o.delValue(name)


def make_ts(rhs, t0, t1, steps, post=None):
    ts = TsSolver(COMM_WORLD)
    ts.setUpTimes(t0, t1, steps)
    ts.initSolver(rhs, post)
    return ts


# ---- src/tests/test_ts.py:11-20 -------------------------------------------------------------------
def test_reference_ts_tests():
    ts = TsSolver()
    assert ts.getType() == "rk" and ts.getRKType() == "5bs"
    ts.setUpTimes(1.2, 3.4, 50)
    assert ts.getTime() == 1.2 and ts.getMaxTime() == 3.4 and ts.getMaxSteps() == 50


LAM = np.array([-1.0, -0.3, 0.7])


def linear_rhs(ts, t, u, f):
    # non-autonomous so that the stage times c_i matter: y' = lam y + cos t
    f.a[:] = LAM * u.a + math.cos(t)


def linear_exact(t, y0):
    # y = e^{lam t} (y0 - p(0)) + p(t), p = particular solution a cos t + b sin t
    a = -LAM / (1 + LAM ** 2)
    b = 1 / (1 + LAM ** 2)
    return np.exp(LAM * t) * (y0 - a) + a * np.cos(t) + b * np.sin(t)


def test_fixed_step_is_fifth_order_with_fsal(opts):
    opts("ts_adapt_type", "none")
    y0 = np.array([1.0, -2.0, 0.5])
    errs, evals = [], []
    for n in (4, 8, 16):
        opts("ts_dt", 1.0 / n)
        ts = make_ts(linear_rhs, 0.0, 1.0, 10 ** 6)
        u = NpVec(y0)
        ts.solve(u)
        assert ts.getStepNumber() == n and abs(ts.getTime() - 1.0) < 1e-13
        errs.append(np.abs(u.a - linear_exact(1.0, y0))[1:])     # (the lam = -1 error changes sign near n = 6)
        evals.append(ts.rhs_evals)
    assert evals == [8 + 7 * (n - 1) for n in (4, 8, 16)]          # first same as last: 7 new stages per step
    rates = np.array([np.log2(errs[i] / errs[i + 1]) for i in range(2)])
    assert np.all((rates > 4.8) & (rates < 5.8)), (errs, rates)


def test_single_steps_and_post_hook(opts):
    opts("ts_adapt_type", "none")
    opts("ts_dt", 0.125)
    seen = []
    ts = make_ts(linear_rhs, 0.0, 1.0, 3, post=lambda t: seen.append((t.step_number, t.time)))
    u = NpVec([1.0, 1.0, 1.0])
    ts.solve(u)                                                   # stops at max-steps
    assert [s for s, _ in seen] == [1, 2, 3] and abs(seen[-1][1] - 0.375) < 1e-15
    ts.step()                                                     # BaseProblem / IBM loops call step() directly
    assert ts.getStepNumber() == 4 and len(seen) == 3             # the hook belongs to solve()
    assert ts.getSolution() is u


def test_matchstep_lands_on_the_final_time(opts):
    opts("ts_adapt_type", "none")
    opts("ts_dt", 0.3)
    ts = make_ts(linear_rhs, 0.0, 1.0, 100)
    assert ts._eft == TS.ExactFinalTime.MATCHSTEP
    u = NpVec([1.0, 0.0, 0.0])
    ts.solve(u)
    assert ts.getStepNumber() == 4 and ts.getTime() == 1.0
    assert np.abs(u.a - linear_exact(1.0, np.array([1.0, 0.0, 0.0]))).max() < 1e-5


def test_basic_controller_rejects_and_meets_the_tolerance(opts):
    """PETSc's `basic` adaptor as restated in ts.py [unpinned, DESIGN section 4]: a too large first
    step is rejected, the step then settles, the global error stays near the tolerance"""
    opts("ts_dt", 2.0)
    y0 = np.array([1.0, -2.0, 0.5])
    ts = make_ts(linear_rhs, 0.0, 3.0, 10 ** 5)
    u = NpVec(y0)
    ts.solve(u)
    assert ts.getTime() == 3.0
    assert ts.rejections >= 1
    err = np.abs(u.a - linear_exact(3.0, y0)).max()
    assert err < 5e-4
    n_default = ts.getStepNumber()
    # tighter tolerances: more steps, smaller error
    opts("ts_rtol", 1e-8)
    opts("ts_atol", 1e-8)
    ts = make_ts(linear_rhs, 0.0, 3.0, 10 ** 5)
    u = NpVec(y0)
    ts.solve(u)
    assert ts.getStepNumber() > n_default
    assert np.abs(u.a - linear_exact(3.0, y0)).max() < 1e-7


def test_rejected_step_leaves_the_state_untouched(opts):
    opts("ts_dt", 50.0)
    calls = []

    def stiff(ts, t, u, f):
        calls.append(t)
        f.a[:] = -5.0 * u.a

    ts = make_ts(stiff, 0.0, 100.0, 1)
    u = NpVec([1.0])
    ts.setSolution(u)
    ts.step()
    assert ts.rejections >= 1 and ts.getStepNumber() == 1
    assert abs(u.a[0] - math.exp(-5.0 * ts.getTime())) < 1e-3
    assert calls[0] == 0.0


def test_fehlberg_pair_is_listed_and_fifth_order(opts):
    """TsSolver.rk_types names '5f' (ts_solver.py:4): the tableau exists and converges at order 5"""
    opts("ts_adapt_type", "none")
    opts("ts_rk_type", "5f")
    errs = []
    for n in (8, 16):
        opts("ts_dt", 1.0 / n)
        ts = make_ts(lambda ts_, t, u, f: np.copyto(f.a, -u.a * (1.0 + t)), 0.0, 1.0, 10000)
        assert ts.getRKType() == "5f"
        u = NpVec([1.0])
        ts.solve(u)
        errs.append(abs(u.a[0] - math.exp(-1.5)))
    assert 4.5 < math.log2(errs[0] / errs[1]) < 6.5


def test_set_solution_drops_the_carried_stage(opts):
    """a state changed between steps (the immersed-boundary loops) must not reuse the
    FSAL stage of the old state"""
    opts("ts_adapt_type", "none")
    opts("ts_dt", 0.1)
    ts = make_ts(lambda ts_, t, u, f: np.copyto(f.a, -u.a), 0.0, 1.0, 100)
    u = NpVec([1.0])
    ts.setSolution(u)
    ts.step()
    n1 = ts.rhs_evals
    ts.step()
    assert ts.rhs_evals - n1 == 7            # FSAL: first stage carried over
    u.a[:] = 5.0
    ts.setSolution(u)
    n2 = ts.rhs_evals
    ts.step()
    assert ts.rhs_evals - n2 == 8            # re-evaluated on the changed state
    assert abs(u.a[0] - 5.0 * math.exp(-0.1)) < 1e-8
