"""Accuracy of the product's branch-free pow / division (e3sm_b200/csrc/thetal_pow.h), host build, against
40-digit mpmath.  The device build differs only in the seed of the reciprocal (MUFU.RCP instead of a float
division), which the two FMA corrections absorb; GPU parity tests cover that path end to end."""
import ctypes
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "oracle", "libpow_shim.so")
KAPPA = 287.04 / 1005.0


@pytest.fixture(scope="module")
def shim():
    if not os.path.exists(SHIM):
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "libpow_shim.so"])
    return ctypes.CDLL(SHIM)


def _ulp_err(got, exact_mpf):
    import mpmath
    ref = float(exact_mpf)
    ulp = np.spacing(abs(ref))
    return float(abs((mpmath.mpf(got) - exact_mpf) / mpmath.mpf(ulp)))


@pytest.mark.parametrize("which,g", [("shim_pow_gamma", 1.0 / (1.0 - KAPPA)), ("shim_pow_km1", KAPPA - 1.0)])
def test_pow_pos_within_2_ulp(shim, which, g):
    import mpmath
    mpmath.mp.dps = 40
    rng = np.random.default_rng(11)
    x = np.exp(rng.uniform(np.log(1e-7), np.log(1e4), 4000))
    x = np.concatenate([x, [1.0, 0.5, 2.0, np.sqrt(2.0), np.nextafter(np.sqrt(2.0), 2.0), 1e-5, 1.0 - 2e-16, 1.0 + 2e-16]])
    y = np.empty_like(x)
    getattr(shim, which)(x.ctypes.data_as(ctypes.c_void_p), y.ctypes.data_as(ctypes.c_void_p), ctypes.c_long(x.size))
    worst = max(_ulp_err(float(yi), mpmath.power(mpmath.mpf(float(xi)), mpmath.mpf(g))) for xi, yi in zip(x, y))
    assert worst <= 2.0, worst            # 1.56 ulp on 2e7 samples (oracle/check_pow.cpp); CUDA pow documents 2 ulp
    assert y[x == 1.0][0] == 1.0


def test_div_fast_within_1_ulp(shim):
    import mpmath
    mpmath.mp.dps = 40
    rng = np.random.default_rng(12)
    a = rng.uniform(1e-3, 1e5, 4000) * rng.choice([-1.0, 1.0], 4000)
    d = rng.uniform(1e-2, 1e4, 4000) * rng.choice([-1.0, 1.0], 4000)
    y = np.empty_like(a)
    shim.shim_div(a.ctypes.data_as(ctypes.c_void_p), d.ctypes.data_as(ctypes.c_void_p), y.ctypes.data_as(ctypes.c_void_p),
                  ctypes.c_long(a.size))
    worst = max(_ulp_err(float(yi), mpmath.mpf(float(ai)) / mpmath.mpf(float(di))) for ai, di, yi in zip(a, d, y))
    assert worst <= 1.0, worst
