"""The arithmetic behind the pow-free EOS updates of the DIRK kernel (csrc/thetal_dirk_col.cuh, eos_level_tiny_step /
eos_level_small_step): pnh_new = pnh_old (1 + e)^gamma with e = dphi_old/dphi_new - 1 and (1 + e)^gamma as its binomial
series, degree 4 for |e| < 2^-10 and degree 11 for |e| < 2^-4.  The CUDA functions cannot run here; this restates their
Horner form in float64 (numpy evaluates the same IEEE operations, one rounding per operation instead of the fused
multiply-add of the device: at most one ulp more) and checks it against 40-digit arithmetic at and below the two
thresholds the kernel uses, i.e. it pins the coefficients, the degrees and the thresholds."""
import numpy as np
from mpmath import mp, mpf

RGAS, CP = 287.04, 1005.0          # physical_constants.F90 (Rgas, Cp): kappa = Rgas/Cp as in csrc/thetal_ctx.h


def coefficients(n):
    g = 1.0 / (1.0 - RGAS / CP)
    c = [g]
    for k in range(2, n + 1):
        c.append(c[-1] * (g - (k - 1)) / k)
    return c                         # G1 .. Gn exactly as the constexpr chain of the kernel builds them


def series(e, n):
    c = coefficients(n)
    s = np.full_like(e, c[-1])
    for ck in reversed(c[:-1]):
        s = s * e + ck
    return 1.0 * e * s + 1.0         # pnh = fma(pnh * e, s, pnh) with pnh = 1


def exact(e):
    mp.dps = 40
    g = mpf(1) / (1 - mpf(RGAS) / mpf(CP))
    return np.array([float((1 + mpf(float(x))) ** g) for x in e])


def test_kernel_constants_match():
    src = open(__file__.replace("tests/test_eos_series.py", "e3sm_b200/csrc/thetal_dirk_col.cuh")).read()
    assert "EOS_SERIES_TINY = 0.0009765625" in src and "#define DIRK_EOS_SERIES_MAX 0.0625" in src
    assert "G4 = G3 * (G1 - 3) / 4;" in src and "G11 = G10 * (G1 - 10) / 11;" in src
    ctx = open(__file__.replace("tests/test_eos_series.py", "e3sm_b200/csrc/thetal_ctx.h")).read()
    assert "KAPPA = RGAS / CP" in ctx and f"RGAS = {RGAS}" in ctx and f"CP = {CP}" in ctx


def test_degree_4_below_2_to_minus_10():
    rng = np.random.default_rng(1)
    e = np.concatenate([rng.uniform(-2.0 ** -10, 2.0 ** -10, 4000), [2.0 ** -10, -2.0 ** -10, 0.0, 1e-9, -1e-13]])
    err = np.abs(series(e, 4) / exact(e) - 1.0)
    assert err.max() <= 2.5e-16, err.max()


def test_degree_11_below_2_to_minus_4():
    rng = np.random.default_rng(2)
    e = np.concatenate([rng.uniform(-2.0 ** -4, 2.0 ** -4, 4000), [2.0 ** -4, -2.0 ** -4]])
    err = np.abs(series(e, 11) / exact(e) - 1.0)
    assert err.max() <= 3.5e-16, err.max()


def test_degree_4_is_not_enough_at_the_upper_threshold():
    """the tiers are needed: degree 4 at |e| = 2^-4 misses by 1e-9"""
    e = np.array([2.0 ** -4])
    assert abs(series(e, 4)[0] / exact(e)[0] - 1.0) > 1e-10
