"""The device's expf / acosf must agree bit for bit with glibc's (which the reference's
host build and the oracle call)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _libm(name, x):
    libm = ctypes.CDLL("libm.so.6")
    f = getattr(libm, name); f.restype = ctypes.c_float; f.argtypes = [ctypes.c_float]
    return np.array([f(float(v)) for v in x], dtype=np.float32)


def test_expf_matches_libm():
    from line3d_b200 import capi
    rng = np.random.default_rng(3)
    x = np.concatenate([-rng.random(100000) * 0.8, -rng.random(100000) * 30.0, -rng.random(2000) * 120.0,
                        [0.0, -0.0, -0.7, -87.0, -88.5, -100.0, -103.9, -104.5, -200.0]]).astype(np.float32)
    ctx = capi.Context(0)
    got = ctx.debug_math(0, x)
    assert np.array_equal(got.view(np.uint32), _libm("expf", x).view(np.uint32))
    ctx.close()


def test_acosf_matches_libm():
    from line3d_b200 import capi
    rng = np.random.default_rng(4)
    x = np.concatenate([rng.random(100000) * 2 - 1, 1 - rng.random(100000) * 3e-2, [1.0, -1.0, 0.0, 0.5, -0.5, 1e-9, 0.49999997, 0.99999994]]).astype(np.float32)
    ctx = capi.Context(0)
    got = ctx.debug_math(1, x)
    assert np.array_equal(got.view(np.uint32), _libm("acosf", x).view(np.uint32))
    ctx.close()
