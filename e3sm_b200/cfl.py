"""HOMME's start-up stability report for the theta-l step (host side, numpy).

Restates print_cfl, H/src/share/global_norms_mod.F90:262-405 (H/ = components/homme/), for np=4 and
constant-coefficient hyperviscosity (hypervis_scaling = 0), with elem%normDinv = the largest singular
value of Dinv over the element's GLL points (cube_mod.F90:281-291, :453).  Every limit is quoted the
way the reference prints it: "dt < S * limit", S the stability region of the time integrator
(S = 2 for the forward-Euler viscosity subcycles).
"""
import numpy as np

LAMBDA_MAX_NP4 = 1.5        # global_norms_mod.F90:280
LAMBDA_VIS_NP4 = 12.0       # global_norms_mod.F90:281


def max_norm_dinv(grid):
    """max over elements and GLL points of ||Dinv||_2 (1/sqrt of the smaller eigenvalue of met = D^T D)."""
    D = grid.D                                                 # [nelem, b, a, 16] = D(i,j,a,b) in Fortran memory
    m11 = D[:, 0, 0] ** 2 + D[:, 0, 1] ** 2                    # met(1,1) = sum_a D(a,1)^2
    m22 = D[:, 1, 0] ** 2 + D[:, 1, 1] ** 2
    m12 = D[:, 0, 0] * D[:, 1, 0] + D[:, 0, 1] * D[:, 1, 1]
    disc = np.sqrt(4.0 * m12 * m12 + (m11 - m22) ** 2)
    l2 = 0.5 * (m11 + m22 - disc)                              # smaller eigenvalue of met
    return float(np.max(1.0 / np.sqrt(np.abs(l2))))


def cfl_report(grid, p):
    """dict of the limits print_cfl writes (seconds, to be multiplied by S) plus the step lengths in use."""
    nd = max_norm_dinv(grid) / grid.rearth                     # scale_factor_inv * max_normDinv
    hv = LAMBDA_VIS_NP4 ** 2 * nd ** 4                         # normDinv_hypervis, :339
    lap = LAMBDA_VIS_NP4 * nd ** 2                             # normDinv_laplace, :349
    out = {
        "max_normDinv": nd * grid.rearth,
        "gravity_wave_dt": 1.0 / (342.0 * nd * LAMBDA_MAX_NP4),                         # :361
        "nu_vor_dt": 1.0 / (p.nu * hv) if p.nu > 0 else np.inf,                          # :371
        "nu_div_dt": 1.0 / (p.nu_div * hv) if p.nu_div > 0 else np.inf,                  # :381 (MODEL_THETA_L)
        "nu_top_dt": 1.0 / (4.0 * p.nu_top * lap) if p.nu_top > 0 else np.inf,           # :387, max nu_scale_top = 4
        "dt_hypervis": p.tstep / max(p.hypervis_subcycle, 1),
        "dt_sponge": p.tstep / max(p.hypervis_subcycle_tom, 1) if p.hypervis_subcycle_tom > 0
        else p.tstep / max(p.hypervis_subcycle, 1),
    }
    return out


def hypervis_is_stable(grid, p, S=2.0):
    """Forward-Euler subcycles inside their stability region: dt_hv < S * min(nu, nu_div limits) and
    the sponge step inside S * nu_top limit."""
    r = cfl_report(grid, p)
    return (r["dt_hypervis"] < S * min(r["nu_vor_dt"], r["nu_div_dt"])) and (r["dt_sponge"] < S * r["nu_top_dt"])
