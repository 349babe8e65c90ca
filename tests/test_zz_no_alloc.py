"""GPU: pyn_vec_dot_noalloc (include/pynama_b200.h, csrc/noalloc.cu) -- the inner product with its
reduction scratch in a fixed file-scope device array, for callers that may not allocate -- against
the existing pyn_vec_dot through the C ABI, on two calls in a row (the scratch keeps the previous
call's contents), then the size limit.  Named to be collected last."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_N = 268435456          # PYN_NOALLOC_MAX_N


def _dot_noalloc(ctx, x, y, n, out):
    from pynama_b200._lib import call, vp
    px, py, po = vp(), vp(), vp()
    call("pyn_vec_device_ptr", x._h, ctypes.byref(px))
    call("pyn_vec_device_ptr", y._h, ctypes.byref(py))
    call("pyn_vec_device_ptr", out._h, ctypes.byref(po))
    call("pyn_vec_dot_noalloc", ctx.handle, px, py, int(n), po, None)      # NULL stream = the context's
    return out.getArray()[0]                                               # read back on the same stream


@pytest.mark.parametrize("n", [1, 255, 16384, 16385, 1_000_003, 5_000_000])
def test_noalloc_dot_equals_vec_dot_twice_in_a_row(n):
    from pynama_b200.petsc import Context, Vec
    ctx = Context.get()
    rng = np.random.default_rng(n)
    out = Vec(1)
    out.set(-7.0)                                   # stale output must be overwritten
    for rep in range(2):                            # second call: the scratch holds the first call's sums
        a, b = rng.uniform(-1, 1, (2, n))
        x, y = Vec().createWithArray(a), Vec().createWithArray(b)
        want = x.dot(y)                             # existing entry point (pyn_vec_dot)
        got = _dot_noalloc(ctx, x, y, n, out)
        scale = float(np.abs(a * b).sum())
        assert abs(got - want) <= 1e-12 * scale, (n, rep, got, want)
        assert abs(got - float(a @ b)) <= 1e-12 * scale
        again = _dot_noalloc(ctx, x, y, n, out)     # fixed summation order: bit-identical
        assert again == got


def test_noalloc_dot_of_nothing_is_zero():
    from pynama_b200.petsc import Context, Vec
    ctx = Context.get()
    x = Vec().createWithArray(np.ones(8))
    out = Vec(1)
    out.set(3.0)
    assert _dot_noalloc(ctx, x, x, 0, out) == 0.0
    assert _dot_noalloc(ctx, x, x, 8, out) == 8.0


def test_noalloc_dot_refuses_sizes_beyond_the_scratch():
    from pynama_b200._lib import PynamaError
    from pynama_b200.petsc import Context, Vec
    ctx = Context.get()
    x = Vec().createWithArray(np.ones(8))
    out = Vec(1)
    with pytest.raises(PynamaError, match="PYN_NOALLOC_MAX_N"):
        _dot_noalloc(ctx, x, x, MAX_N + 1, out)
    with pytest.raises(PynamaError, match="negative"):
        _dot_noalloc(ctx, x, x, -1, out)
    assert _dot_noalloc(ctx, x, x, 8, out) == 8.0      # still usable after a refusal
