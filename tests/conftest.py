import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # the oracle is test infrastructure: build it if the prebuilt library did not travel
    if not os.path.exists(os.path.join(ROOT, "oracle", "libthetal_oracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "all"])


def rel_l2(a, b):
    """sqrt(sum (a-b)^2 / sum b^2), SURVEY.md 8(d)."""
    d = np.sqrt(np.sum((a - b) ** 2))
    n = np.sqrt(np.sum(b ** 2))
    return d / n if n > 0 else d


def assert_state_close(st_a, st_b, tl, tol, what=""):
    """per-prognostic relative L2 (u,v together) at time level tl."""
    out = {}
    for f in ("v", "w_i", "vtheta_dp", "phinh_i", "dp3d"):
        out[f] = rel_l2(getattr(st_a, f)[:, tl], getattr(st_b, f)[:, tl])
    bad = {k: v for k, v in out.items() if not (v <= tol)}
    assert not bad, f"{what}: rel-L2 above {tol:g}: {bad} (all: {out})"
    return out


@pytest.fixture(scope="session")
def case_c1():
    """C1: ne4 L26 JW baroclinic wave, host-side inputs."""
    from e3sm_b200.params import config
    from e3sm_b200.dycore import build_case
    p = config("C1")
    return (p,) + build_case(p)


def spun_up_state(case, nsteps=10, seed=20240611, perturb=1e-3):
    """Parity state of SURVEY 8(d): JW advanced nsteps by the oracle (so w != 0), every prognostic
    multiplied by 1+perturb*U(-1,1), made C0 again, phinh_i(nlevp)=phis, w b.c. reset."""
    from e3sm_b200.initial import make_c0
    from e3sm_b200.geometry import G
    from oracle.oracle import Oracle
    p, grid, hv, st0, phis, gradphis, refs = case
    st = st0.copy()
    o = Oracle(grid, hv, p, phis, gradphis, refs, st, nthreads=os.cpu_count() or 1)
    nm1, n0, np1 = 0, 1, 2
    for _ in range(nsteps):
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    assert o.error_flag == 0
    rng = np.random.default_rng(seed)
    nlev = hv.nlev
    for f in st.FIELDS:
        a = getattr(st, f)
        a[:, 0] = make_c0(grid, a[:, n0] * (1 + perturb * rng.uniform(-1, 1, a[:, n0].shape)))
    st.phinh_i[:, 0, nlev] = phis
    st.w_i[:, 0, nlev] = (st.v[:, 0, nlev - 1, 0] * gradphis[:, 0] + st.v[:, 0, nlev - 1, 1] * gradphis[:, 1]) / G
    st.set_all_timelevels(0)
    assert (st.dp3d > 0).all() and (np.diff(st.phinh_i[:, 0], axis=1) < 0).all() and (st.vtheta_dp > 0).all()
    return st


@pytest.fixture(scope="session")
def parity_state_c1(case_c1):
    return spun_up_state(case_c1)


def global_diagnostics(st, tl, grid, hv):
    """min surface pressure, dry mass and total energy (KE + w-KE + internal + potential) of one time
    level, in the spirit of prim_state_mod.F90:897-935 (same formula for both runs being compared)."""
    from e3sm_b200.geometry import CP, KAPPA, P0, RGAS
    dp, th = st.dp3d[:, tl], st.vtheta_dp[:, tl]
    v, w, phi = st.v[:, tl], st.w_i[:, tl], st.phinh_i[:, tl]
    ps = hv.hyai[0] * hv.ps0 + dp.sum(axis=1)                         # [e,16]
    area = grid.spheremp
    mass = float((ps * area).sum())
    dphi = phi[:, 1:] - phi[:, :-1]
    poe = RGAS * th / (-dphi)
    pnh = P0 * (poe / P0) ** (1.0 / (1.0 - KAPPA))
    exner = pnh / poe
    ke = 0.5 * dp * (v[:, :, 0] ** 2 + v[:, :, 1] ** 2)
    kew = 0.25 * dp * (w[:, :-1] ** 2 + w[:, 1:] ** 2)
    ie = CP * th * exner + pnh * dphi
    pe = 0.5 * dp * (phi[:, :-1] + phi[:, 1:])
    col = (ke + kew + ie + pe).sum(axis=1) + hv.hyai[0] * hv.ps0 * phi[:, 0]
    energy = float((col * area).sum())
    return float(ps.min()), mass, energy
