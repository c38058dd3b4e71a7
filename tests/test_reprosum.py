"""Order-independent fixed-point sum of e3sm_b200/csrc/thetal_reprosum.h (host build through
oracle/libpow_shim.so): any permutation gives the same bits, and the value is the exact sum rounded
(math.fsum) to within one ulp.  The device path (diag_accumulate_kernel) runs the same split()."""
import ctypes
import math
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "oracle", "libpow_shim.so")


@pytest.fixture(scope="module")
def shim():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "libpow_shim.so"])
    L = ctypes.CDLL(SHIM)
    L.shim_reprosum.restype = ctypes.c_double
    L.shim_reprosum.argtypes = [ctypes.c_void_p, ctypes.c_long]
    return L


def _sum(L, x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    return L.shim_reprosum(x.ctypes.data, x.size)


@pytest.mark.parametrize("kind", ["energy-like", "mixed-sign", "wide-range", "tiny"])
def test_permutation_invariance_and_accuracy(shim, kind):
    rng = np.random.default_rng(11)
    n = 43350                                                  # ne85
    if kind == "energy-like":
        x = 2.5e9 * (1.0 + 0.3 * rng.standard_normal(n))
    elif kind == "mixed-sign":
        x = rng.standard_normal(n) * 1e5
    elif kind == "wide-range":
        x = rng.standard_normal(n) * 10.0 ** rng.uniform(-12, 12, n)
    else:
        x = rng.standard_normal(n) * 1e-300
    ref = _sum(shim, x)
    for seed in range(5):
        perm = np.random.default_rng(seed).permutation(n)
        assert _sum(shim, x[perm]) == ref                      # bit for bit
    exact = math.fsum(x.tolist())
    scale = max(abs(exact), np.abs(x).max() * 2.0 ** -52)
    assert abs(ref - exact) <= 2.0 ** -52 * scale * 1.01 + n * np.abs(x).max() * 2.0 ** -95


def test_edge_cases(shim):
    assert _sum(shim, np.zeros(7)) == 0.0
    assert _sum(shim, np.array([3.0])) == 3.0
    assert _sum(shim, np.array([1.0, -1.0, 2.0 ** -60])) == 2.0 ** -60
    x = np.array([2.0 ** 40, 1.0, -2.0 ** 40, 2.0 ** -40])     # cancellation keeps the small terms
    assert _sum(shim, x) == 1.0 + 2.0 ** -40
    big = np.full(100000, 1.9999999)                           # many addends near the top of a binade
    assert abs(_sum(shim, big) - math.fsum(big.tolist())) <= 2.0 ** -52 * 2e5
