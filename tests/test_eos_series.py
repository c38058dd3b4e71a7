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


def test_eos_at_the_hydrostatic_guess_returns_the_hydrostatic_pressure(case_c1):
    """IM:228-230 builds the Newton guess with phi_from_eos (EOS:286-302), the inverse of the EOS (EOS:175-183) at the
    hydrostatic mid-level pressure; the kernel therefore takes pnh = pmid at the guess instead of evaluating the EOS
    (IM:261).  What is neglected is the rounding of the interface sums: measured here on the JW state of C1."""
    p, grid, hv, st, phis, gradphis, refs = case_c1
    P0 = 1.0e5
    kappa = RGAS / CP
    dp, th = st.dp3d[:, 0], st.vtheta_dp[:, 0]                       # [e, nlev, 16]
    nlev = hv.nlev
    ptop = hv.hyai[0] * hv.ps0
    pint = ptop + np.concatenate([np.zeros_like(dp[:, :1]), np.cumsum(dp, axis=1)], axis=1)
    pmid = (pint[:, 1:] + pint[:, :-1]) / 2
    term = RGAS * th * (pmid / P0) ** (kappa - 1.0) / P0           # -dphi of the guess
    phi = np.empty((dp.shape[0], nlev + 1, 16))
    phi[:, nlev] = phis
    for k in range(nlev - 1, -1, -1):
        phi[:, k] = phi[:, k + 1] + term[:, k]
    dphi = phi[:, 1:] - phi[:, :-1]                                  # what the Newton iteration starts from
    pnh = P0 * (RGAS * th / (-dphi) / P0) ** (1.0 / (1.0 - kappa))
    rel = np.abs(pnh / pmid - 1.0).max()
    assert rel <= 2e-14, rel          # measured 2.4e-15 (L26), 4.7e-15 (L72)
