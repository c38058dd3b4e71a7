"""GPU-vs-oracle parity through the C ABI (libthetal_b200.so), in the differential style of the
reference's unit tests (H/test_execs/thetal_kokkos_ut/caar_ut.cpp, dirk_ut.cpp, hv_ut.cpp): the same
state goes through the oracle routine and the CUDA routine.  Tolerance: relative L2 <= 1e-10 per
prognostic (BASELINE.json north_star); the oracle is compiled without FMA contraction, the GPU uses
FMA, so bitwise equality is not expected."""
import os

import numpy as np
import pytest

from conftest import assert_state_close, rel_l2, spun_up_state

pytestmark = pytest.mark.gpu
TOL = 1e-10
NT = os.cpu_count() or 1


def distinct_levels(st, phis, gradphis):
    """make the three time levels different (still C0, still physical)"""
    from e3sm_b200.geometry import G
    s = st.copy()
    nlev = s.nlev
    for t, fac in ((1, 1.0 + 3e-4), (2, 1.0 - 2e-4)):
        for f in s.FIELDS:
            getattr(s, f)[:, t] *= fac
        s.phinh_i[:, t, nlev] = phis
        s.w_i[:, t, nlev] = (s.v[:, t, nlev - 1, 0] * gradphis[:, 0] + s.v[:, t, nlev - 1, 1] * gradphis[:, 1]) / G
    return s


def make_pair(case, st, **over):
    from e3sm_b200.dycore import ThetaLDycore
    from oracle.oracle import Oracle
    p, grid, hv, _, phis, gradphis, refs = case
    p = p.with_(**over) if over else p
    so = st.copy()
    o = Oracle(grid, hv, p, phis, gradphis, refs, so, nthreads=NT)
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    d.set_state(st)
    return p, o, so, d


@pytest.mark.parametrize("np1,nm1,n0", [(2, 0, 1), (2, 1, 1), (2, 1, 2), (0, 1, 1), (2, 2, 2)])
@pytest.mark.parametrize("form,pgrad", [(0, 0), (1, 1), (2, 0)])
def test_caar(case_c1, parity_state_c1, np1, nm1, n0, form, pgrad):
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = distinct_levels(parity_state_c1, phis, gradphis)
    p, o, so, d = make_pair(case_c1, st, theta_advect_form=form, pgrad_correction=pgrad)
    args = (np1, nm1, n0, 37.5, 0.25, 1.0, 0.0, 1.0)
    o.compute_andor_apply_rhs(*args)
    d.compute_andor_apply_rhs(*args)
    sg = d.get_state()
    assert_state_close(sg, so, np1, TOL, f"caar np1={np1} nm1={nm1} n0={n0} form={form}")
    vn0, eta, om = d.get_derived()
    assert rel_l2(vn0, o.vn0) <= TOL
    assert rel_l2(om, o.omega_p) <= 1e-9      # diagnostic only; cancellation in v.grad(p) - omega_i
    d.close()


def test_caar_scale2(case_c1, parity_state_c1):
    """fully explicit stages (tstep_type 5): scale2 = 1"""
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = distinct_levels(parity_state_c1, phis, gradphis)
    p, o, so, d = make_pair(case_c1, st)
    args = (2, 0, 1, 20.0, 0.0, 1.0, 1.0, 1.0)
    o.compute_andor_apply_rhs(*args)
    d.compute_andor_apply_rhs(*args)
    assert_state_close(d.get_state(), so, 2, TOL, "caar scale2=1")
    d.close()


@pytest.mark.parametrize("a_nm1,a_n0", [(0.0, 0.0), (8.0, 80.0)])
def test_dirk(case_c1, parity_state_c1, a_nm1, a_n0):
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = distinct_levels(parity_state_c1, phis, gradphis)
    p, o, so, d = make_pair(case_c1, st)
    # a CAAR stage first so that np1 is an un-converged explicit update, as in prim_advance_exp
    for x in (o, d):
        x.compute_andor_apply_rhs(2, 1, 1, 120.0, 0.0, 1.0, 0.0, 1.0)
    ito, _ = o.compute_stage_value_dirk(0, a_nm1, 1, a_n0, 2, 120.0)
    itg, _ = d.compute_stage_value_dirk(0, a_nm1, 1, a_n0, 2, 120.0)
    assert_state_close(d.get_state(), so, 2, TOL, "dirk")
    assert itg == ito
    d.close()


@pytest.mark.parametrize("hvs,tom,nu_div_fac", [(1, 1, 1.0), (3, 1, 2.5), (2, 0, 2.5)])
def test_hypervis(case_c1, parity_state_c1, hvs, tom, nu_div_fac):
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = parity_state_c1.copy()
    p, o, so, d = make_pair(case_c1, st, hypervis_subcycle=hvs, hypervis_subcycle_tom=tom,
                            nu_div=case_c1[0].nu * nu_div_fac)
    o.advance_hypervis(2, 600.0, 1.0)
    d.advance_hypervis(2, 600.0, 1.0)
    assert_state_close(d.get_state(), so, 2, TOL, f"hypervis hvs={hvs} tom={tom}")
    d.close()


@pytest.mark.parametrize("hvs", [1, 2])
def test_hypervis_tensor(hvs):
    """tensor hyperviscosity (hypervis_scaling = 3: tensorVisc from cube_mod.F90:313-428 + dss_hvtensor, var_coef
    on the second laplacian only, Cartesian vector laplacian; viscosity_theta.F90:147-161, derivative_mod.F90:
    1090-1117,1164-1208) on a perturbed spun-up state"""
    from e3sm_b200.dycore import build_case
    from e3sm_b200.params import config
    p = config("C1").with_(hypervis_scaling=3.0, nu=3.4e-8, nu_p=3.4e-8, nu_s=3.4e-8, nu_div=3.4e-8, hypervis_subcycle=hvs)
    case = (p,) + build_case(p)
    assert np.abs(case[1].tensorvisc).max() > 0
    st = spun_up_state(case, nsteps=4)
    p, o, so, d = make_pair(case, st)
    o.advance_hypervis(2, 600.0, 1.0)
    d.advance_hypervis(2, 600.0, 1.0)
    assert_state_close(d.get_state(), so, 2, TOL, f"tensor hypervis hvs={hvs}")
    # the operator really acted
    assert rel_l2(so.v[:, 2], st.v[:, 2]) > 1e-8
    d.close()


def test_tensor_hypervis_rejects_nu_div(case_c1):
    """derivative_mod.F90:1146-1150: tensorHV can not be used with nu_div /= nu"""
    import ctypes
    from e3sm_b200 import lib
    from e3sm_b200.dycore import ThetaLDycore
    p, grid, hv, st, phis, gradphis, refs = case_c1
    L = lib.load()
    L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(0)))
    try:
        with pytest.raises(RuntimeError, match="tensorHV can not be used"):
            ThetaLDycore(grid, hv, p.with_(hypervis_scaling=3.0, nu_div=2 * p.nu), phis, gradphis, refs)
    finally:
        L.thetal_b200_clear_error()
        L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(1)))
        L.finalize_hommexx_session()


@pytest.mark.parametrize("tt", [9, 10, 7, 5, 4, 1])
def test_one_step(case_c1, parity_state_c1, tt):
    """one full prim_advance_exp: rel-L2 <= 1e-10 per prognostic (north star)"""
    p, o, so, d = make_pair(case_c1, parity_state_c1.copy(), tstep_type=tt)
    dt = p.tstep if tt in (7, 9, 10) else 0.5   # explicit NH tables: vertical acoustic CFL
    o.prim_advance_exp(dt, 0, 1, 2)
    d.prim_advance_exp(dt, 0, 1, 2)
    assert o.error_flag == 0
    sg = d.get_state()
    res = assert_state_close(sg, so, 2, TOL, f"one step tstep_type={tt}")
    for t in (0, 1):  # the other time levels too (nm1 is scratch for tt 5,9,10; n0 gets its w b.c.)
        assert_state_close(sg, so, t, TOL, f"tstep_type={tt} level {t}")
    print(tt, res)
    d.close()


def test_ten_steps(case_c1):
    """JW from rest, 10 steps with time-level rotation: differences stay at rounding level."""
    p, grid, hv, st0, phis, gradphis, refs = case_c1
    p, o, so, d = make_pair(case_c1, st0.copy())
    nm1, n0, np1 = 0, 1, 2
    for _ in range(10):
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    ms, g_nm1, g_n0, g_np1 = d.time_steps(p.tstep, 10, 0, 1, 2)
    assert (g_nm1, g_n0, g_np1) == (nm1, n0, np1)
    sg = d.get_state()
    for f in ("v", "vtheta_dp", "phinh_i", "dp3d"):
        assert rel_l2(getattr(sg, f)[:, n0], getattr(so, f)[:, n0]) <= 1e-9, f
    # w is ~1e-3 m/s after 10 steps from rest: measure it against the velocity scale too
    assert rel_l2(sg.w_i[:, n0], so.w_i[:, n0]) <= 1e-7
    d.close()


@pytest.mark.parametrize("cfg,over", [("C2", {}), ("C5", {}), ("C4", {}), ("C3", {}), ("C2", {"tstep_type": 10}),
                                      ("C2", {"tstep_type": 7, "theta_advect_form": 2, "pgrad_correction": 1}),
                                      # the EAMxx option set short of the tensor hyperviscosity (namelist_defaults_eamxx.xml:925-997)
                                      ("C2", {"hv_ref_profiles": 6, "theta_advect_form": 1, "pgrad_correction": 1,
                                              "vert_remap_u_alg": 10}),
                                      # the full EAMxx set: tensor hyperviscosity hypervis_scaling = 3, nu = 3.4e-8, at L72 and L128
                                      ("C2", dict(EAMXX := {"hypervis_scaling": 3.0, "nu": 3.4e-8, "nu_p": 3.4e-8, "nu_s": 3.4e-8,
                                                            "nu_div": 3.4e-8, "hv_ref_profiles": 6, "theta_advect_form": 1,
                                                            "pgrad_correction": 1, "vert_remap_u_alg": 10})),
                                      ("C4", dict(EAMXX, nu_top=4.0e4, hypervis_subcycle_tom=1))])
def test_steps_at_full_vertical_resolution(cfg, over):
    """The bench/scaling configurations at their real nlev (72 / 128 levels, their dt, hyperviscosity
    and sponge settings) on the coarse ne4 sphere: three steps from the JW state, GPU vs oracle."""
    from e3sm_b200.dycore import build_case
    from e3sm_b200.params import config
    p = config(cfg).with_(ne=4, **over)
    grid, hv, st, phis, gradphis, refs = build_case(p)
    case = (p, grid, hv, st, phis, gradphis, refs)
    p, o, so, d = make_pair(case, st.copy())
    nm1, n0, np1 = 0, 1, 2
    for _ in range(3):
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        d.prim_advance_exp(p.tstep, nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    assert o.error_flag == 0
    sg = d.get_state()
    for f in ("v", "vtheta_dp", "phinh_i", "dp3d"):
        r = rel_l2(getattr(sg, f)[:, n0], getattr(so, f)[:, n0])
        assert r <= TOL, (cfg, over, f, r)
    # w starts from rest: measure it against its own max with an absolute floor
    dw = float(np.abs(sg.w_i[:, n0] - so.w_i[:, n0]).max())
    assert dw <= 1e-9 * max(1.0, float(np.abs(so.w_i[:, n0]).max())), dw
    # ... and the vertical remap that follows the rsplit = 3 steps (72 / 128 level columns)
    assert o.vertical_remap(n0, p.vert_remap_u_alg) == 0
    d.vertical_remap(n0)
    sg = d.get_state()
    for f in ("v", "vtheta_dp", "phinh_i", "dp3d"):
        r = rel_l2(getattr(sg, f)[:, n0], getattr(so, f)[:, n0])
        assert r <= TOL, ("remap", cfg, over, f, r)
    dw = float(np.abs(sg.w_i[:, n0] - so.w_i[:, n0]).max())
    assert dw <= 1e-9 * max(1.0, float(np.abs(so.w_i[:, n0]).max())), dw
    d.close()


@pytest.mark.parametrize("remap", [False, True])
def test_ten_day_baroclinic_wave_diagnostics(case_c1, remap):
    """North-star check: 10 simulated days of the JW baroclinic wave (C1: ne4 L26, dt=600 s, 1440
    dynamics steps, tstep_type 9 + hyperviscosity + sponge); min surface pressure, dry mass and total
    energy of the GPU run within 1e-6 relative of the oracle run.  remap=True is the reference's flow:
    prim_run_subcycle_c = rsplit Lagrangian steps + vertical_remap, 480 times; remap=False leaves the
    levels floating for the whole run (which the coarse ne4 wave tolerates)."""
    from conftest import global_diagnostics
    p, grid, hv, st0, phis, gradphis, refs = case_c1
    p, o, so, d = make_pair(case_c1, st0.copy())
    nsteps = 1440
    nm1, n0, np1 = 0, 1, 2
    for i in range(nsteps):
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
        if remap and i % p.rsplit == p.rsplit - 1:
            assert o.vertical_remap(n0, p.vert_remap_u_alg) == 0
    assert o.error_flag == 0
    if remap:
        for _ in range(nsteps // p.rsplit):
            nstep, g_nm1, g_n0, g_np1 = d.prim_run_subcycle(p.tstep)
        assert nstep == nsteps
    else:
        ms, g_nm1, g_n0, g_np1 = d.time_steps(p.tstep, nsteps, 0, 1, 2)
    assert g_n0 == n0
    sg = d.get_state()
    dg, do = global_diagnostics(sg, n0, grid, hv), global_diagnostics(so, n0, grid, hv)
    d0 = global_diagnostics(st0, 0, grid, hv)
    print("10-day diagnostics (min ps, mass, energy): gpu", dg, "oracle", do, "initial", d0)
    for a, b, name in zip(dg, do, ("min ps", "mass", "total energy")):
        assert abs(a - b) <= 1e-6 * abs(b), (name, a, b)
    assert abs(dg[1] - d0[1]) <= 1e-11 * d0[1]          # dry mass is conserved by the step (and the remap)
    dev = d.global_diagnostics(n0)                       # the same numbers computed on the device (N3)
    four_pi = 4.0 * np.pi
    assert abs(dev["ps_min"] - do[0]) <= 1e-6 * do[0]
    assert abs(dev["ps_mean"] * four_pi - do[1]) <= 1e-6 * do[1]
    assert abs(dev["energy"] * four_pi - do[2]) <= 1e-6 * abs(do[2])
    assert dg[0] < d0[0] - 100.0                         # and the wave really developed
    for f in ("v", "vtheta_dp", "phinh_i", "dp3d"):     # the fields themselves are still close
        assert rel_l2(getattr(sg, f)[:, n0], getattr(so, f)[:, n0]) <= 1e-6, f
    d.close()


def test_device_global_diagnostics_match_host_formulas(case_c1, parity_state_c1):
    """thetal_b200_global_diagnostics (device, order-independent sum) against the numpy restatement of
    prim_energy_halftimes + global_integral on the same state; repeated calls give the same bits."""
    from conftest import global_diagnostics
    p, grid, hv = case_c1[0], case_c1[1], case_c1[2]
    p, o, so, d = make_pair(case_c1, parity_state_c1.copy())
    d.prim_advance_exp(p.tstep, 0, 1, 2)
    sg = d.get_state()
    for tl in (1, 2):
        ps_min, mass, energy = global_diagnostics(sg, tl, grid, hv)
        dev = d.global_diagnostics(tl)
        four_pi = 4.0 * np.pi
        assert abs(dev["ps_min"] - ps_min) <= 1e-13 * ps_min
        assert abs(dev["ps_mean"] * four_pi - mass) <= 1e-13 * mass
        assert abs(dev["energy"] * four_pi - energy) <= 1e-12 * abs(energy)
        assert dev["ps_max"] >= dev["ps_mean"] >= dev["ps_min"] > 0
        assert abs(dev["KE"] + dev["PE"] + dev["IE"] - dev["energy"]) <= 1e-15 * abs(dev["energy"])
        assert d.global_diagnostics(tl) == dev
    d.close()


@pytest.mark.parametrize("alg", [1, 10, 11])
def test_vertical_remap(case_c1, parity_state_c1, alg):
    """vertical_remap (vertremap_mod.F90:23-171) after rsplit Lagrangian steps: GPU vs oracle <= 1e-10 per
    prognostic for the three PPM variants; dp3d lands exactly on the reference levels."""
    p0, grid, hv = case_c1[0], case_c1[1], case_c1[2]
    case = (p0.with_(vert_remap_u_alg=alg),) + tuple(case_c1[1:])
    p, o, so, d = make_pair(case, parity_state_c1.copy())
    nm1, n0, np1 = 0, 1, 2
    for _ in range(p.rsplit):
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        d.prim_advance_exp(p.tstep, nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    before = d.get_state().dp3d[:, n0].copy()
    assert o.vertical_remap(n0, alg) == 0
    d.vertical_remap(n0)
    sg = d.get_state()
    assert_state_close(sg, so, n0, 1e-10, f"vertical_remap alg {alg}")
    ps = hv.hyai[0] * hv.ps0 + before.sum(axis=1)
    dpref = ((hv.hyai[1:] - hv.hyai[:-1])[None, :, None] * hv.ps0
             + (hv.hybi[1:] - hv.hybi[:-1])[None, :, None] * ps[:, None, :])
    assert np.abs(sg.dp3d[:, n0] - dpref).max() <= 1e-12 * dpref.max()
    assert rel_l2(sg.dp3d[:, n0], before) > 1e-7            # the levels really had floated
    d.close()


def test_vertical_remap_negative_layer_is_fatal(case_c1, parity_state_c1):
    """vertremap_mod.F90:87-103: 'negative layer thickness' aborts; here error code 11."""
    import ctypes
    from e3sm_b200 import lib
    st = parity_state_c1.copy()
    st.dp3d[3, 1, 5, 7] = -1.0
    L = lib.load()
    L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(0)))      # latch instead of abort()
    try:
        p, o, so, d = make_pair(case_c1, st)
        with pytest.raises(RuntimeError, match="negative layer thickness"):
            d.vertical_remap(1)
        d.close()
    finally:
        L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(1)))


def test_pipelined_step_service(case_c1, parity_state_c1):
    """thetal_b200_pipeline_step: 7 overlapped calls on three different inputs reproduce, bit for bit, the
    plain prim_advance_exp of each input (staging buffers trade places with the time-level slots)."""
    import ctypes
    from e3sm_b200.dycore import _cd
    from e3sm_b200.initial import State
    FIELDS = ("v", "w_i", "vtheta_dp", "phinh_i", "dp3d")
    p, o, so, d = make_pair(case_c1, parity_state_c1.copy())
    base = parity_state_c1
    inputs, expect = [], []
    for m in range(3):
        st = base.copy()
        st.v[:, 1] *= (1.0 + 0.01 * m)
        d.set_state(st)
        d.prim_advance_exp(p.tstep, 0, 1, 2)
        sg = d.get_state()
        inputs.append({f: np.ascontiguousarray(getattr(st, f)[:, 1]) for f in FIELDS})
        expect.append({f: getattr(sg, f)[:, 2].copy() for f in FIELDS})
    outs = [{f: np.full_like(inputs[0][f], np.nan) for f in FIELDS} for _ in range(7)]
    cp = lambda a: ctypes.c_void_p(a.ctypes.data)
    for j in range(7):
        i = inputs[j % 3]
        d.L.thetal_b200_pipeline_step(_cd(p.tstep), *[cp(i[f]) for f in FIELDS], *[cp(outs[j][f]) for f in FIELDS])
    d.L.thetal_b200_pipeline_flush()
    d._check()
    for j in range(7):
        for f in FIELDS:
            assert np.array_equal(outs[j][f], expect[j % 3][f]), (j, f)
    d.close()


def test_prim_run_subcycle_against_committed_fixture(case_c1):
    """prim_run_subcycle_c from the JW initial state against tests/golden/oracle_c1_3steps_remap.json (a
    regression fixture written by the oracle, tools/make_golden.py): nothing on this path runs the oracle."""
    from conftest import global_diagnostics
    from test_oracle_invariants import _check_against_fixture
    from e3sm_b200.dycore import ThetaLDycore
    p, grid, hv, st0, phis, gradphis, refs = case_c1
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    d.set_state(st0.copy())
    nstep, c_nm1, c_n0, c_np1 = d.prim_run_subcycle(p.tstep)
    sg = d.get_state()
    _check_against_fixture(sg, c_n0, global_diagnostics(sg, c_n0, grid, hv), 1e-9)
    d.close()


def test_step_is_bitwise_repeatable():
    """The same dynamics step from the same device state gives the same bits every time (ne12 L72, 864
    elements = several waves of CTAs): no order-dependent reduction, no shared-memory race."""
    from e3sm_b200.dycore import ThetaLDycore, build_case
    from e3sm_b200.params import config
    p = config("C2").with_(ne=12, tstep=600.0, nu=2e16, nu_div=2e16, nu_p=2e16, nu_s=2e16)
    grid, hv, st, phis, gradphis, refs = build_case(p)
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    d.set_state(st)
    ms, nm1, n0, np1 = d.time_steps(p.tstep, 3, 0, 1, 2)
    base = d.get_state()
    outs = []
    for _ in range(4):
        d.set_state(base)
        d.prim_advance_exp(p.tstep, nm1, n0, np1)
        d.vertical_remap(np1)
        s = d.get_state()
        outs.append({f: getattr(s, f)[:, np1].copy() for f in s.FIELDS})
    for o in outs[1:]:
        for f in base.FIELDS:
            assert np.isfinite(o[f]).all() and np.array_equal(o[f], outs[0][f]), f
    d.close()


@pytest.mark.parametrize("tt", [9, 10, 7])
def test_dss_inside_the_dirk_kernel_equals_the_separate_dss_kernel(case_c1, parity_state_c1, tt):
    """THETAL_B200_FUSE_DSS=0 restores caar_dss_post_kernel between every CAAR stage and its DIRK solve; the default
    lets the DIRK kernel do that DSS while it stages an element.  Same arithmetic in the same order: same state."""
    import os
    from e3sm_b200.dycore import ThetaLDycore
    p, grid, hv, _, phis, gradphis, refs = case_c1
    p = p.with_(tstep_type=tt)
    outs, launches = [], []
    for knob in ("1", "0"):
        os.environ["THETAL_B200_FUSE_DSS"] = knob
        try:
            d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
            d.set_state(parity_state_c1.copy())
            n_before = d.L.thetal_b200_launch_count()
            d.prim_advance_exp(p.tstep, 0, 1, 2)
            launches.append(d.L.thetal_b200_launch_count() - n_before)
            outs.append(d.get_state())
            d.close()
        finally:
            del os.environ["THETAL_B200_FUSE_DSS"]
    assert launches[1] - launches[0] == (5 if tt == 7 else 4)      # one kernel less per fused stage
    for t in (0, 1, 2):
        assert_state_close(outs[0], outs[1], t, 1e-14, f"fused vs separate DSS, tstep_type={tt} level {t}")


@pytest.mark.parametrize("qsplit", [1, 2])
def test_prim_run_subcycle(case_c1, parity_state_c1, qsplit):
    """The reference entry point called as the Fortran does (prim_driver_mod.F90:554-560): four
    uninitialised integers in, the library's own 0-based TimeLevel out (prim_driver.cpp:174-178).  Each call =
    rsplit x prim_step; a prim_step zeroes vn0/omega_p/eta_dot_dpdn (prim_step.cpp:24-66) and runs qsplit
    dynamics steps; then vertical_remap and the last leapfrog update.  Two calls, so that accumulators that
    were not reset would show."""
    p, o, so, d = make_pair(case_c1, parity_state_c1.copy(), qsplit=qsplit)
    nm1, n0, np1 = 0, 1, 2
    for call in range(2):
        for r in range(p.rsplit):
            o.vn0[:] = 0; o.omega_p[:] = 0; o.eta_dot_dpdn[:] = 0
            for q in range(p.qsplit):
                o.prim_advance_exp(p.tstep, nm1, n0, np1)
                nm1, n0, np1 = n0, np1, nm1
        assert o.vertical_remap(n0, p.vert_remap_u_alg) == 0        # prim_driver_base.F90:1159
        nstep, c_nm1, c_n0, c_np1 = d.prim_run_subcycle(p.tstep)
        assert nstep == (call + 1) * p.rsplit * p.qsplit and (c_nm1, c_n0, c_np1) == (nm1, n0, np1)
        # what the Fortran would now hold: tl%n0 = n0_c + 1 etc.
        assert sorted((c_nm1 + 1, c_n0 + 1, c_np1 + 1)) == [1, 2, 3]
    assert_state_close(d.get_state(), so, n0, 1e-9, "prim_run_subcycle")
    vn0, eta, om = d.get_derived()
    assert rel_l2(vn0, o.vn0) <= 1e-9                                # mean flux over the LAST qsplit steps only
    assert rel_l2(om, o.omega_p) <= 1e-8
    d.close()


def test_prim_run_subcycle_follows_init_time_level(case_c1, parity_state_c1):
    """the library advances from the levels init_time_level_c gave it, whatever the integers hold on entry"""
    p, o, so, d = make_pair(case_c1, parity_state_c1.copy())
    d.init_time_level(3, 1, 2, nstep=7)          # Fortran 1-based: nm1=3, n0=1, np1=2
    nm1, n0, np1 = 2, 0, 1
    for _ in range(p.rsplit * p.qsplit):
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    assert o.vertical_remap(n0, p.vert_remap_u_alg) == 0
    nstep, c_nm1, c_n0, c_np1 = d.prim_run_subcycle(p.tstep)
    assert nstep == 7 + p.rsplit * p.qsplit and (c_nm1, c_n0, c_np1) == (nm1, n0, np1)
    assert_state_close(d.get_state(), so, n0, 1e-9, "prim_run_subcycle from rotated levels")
    d.close()


def test_latched_error_returns_before_touching_state(case_c1, parity_state_c1):
    """abort_on_error = 0: a failed precondition returns from the entry point (no out-of-range time level use)"""
    import ctypes
    from e3sm_b200 import lib
    L = lib.load()
    L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(0)))
    try:
        p, o, so, d = make_pair(case_c1, parity_state_c1.copy())
        before = d.get_state()
        with pytest.raises(RuntimeError, match="bad state/time level"):
            d.compute_andor_apply_rhs(7, 1, 1, 10.0, 0.0, 1.0, 0.0, 1.0)
        L.thetal_b200_clear_error()
        after = d.get_state()
        for f in before.FIELDS:
            assert np.array_equal(getattr(before, f), getattr(after, f))
        d.close()
    finally:
        L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(1)))


def test_limiters_random_state(case_c1):
    """unphysical random states trigger limiter_dp3d_k / vtheta limiter (caar_ut.cpp recipe)"""
    from e3sm_b200.initial import random_state, make_c0
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = random_state(grid, hv, phis, seed=7)
    for f in st.FIELDS:
        a = getattr(st, f)
        a[:, 0] = make_c0(grid, a[:, 0])
    st.phinh_i[:, 0, hv.nlev] = phis
    st.set_all_timelevels(0)
    # thin layers: force the dp3d limiter in some columns, and tiny theta in others
    st.dp3d[::7, :, 3, :] *= 0.02
    st.vtheta_dp[::5, :, 10, :] *= 1e-3
    p, o, so, d = make_pair(case_c1, st)
    args = (2, 1, 1, 1.0, 0.5, 1.0, 0.0, 1.0)
    o.compute_andor_apply_rhs(*args)
    d.compute_andor_apply_rhs(*args)
    assert_state_close(d.get_state(), so, 2, 1e-9, "caar on random state (limiters)")
    d.close()


@pytest.mark.parametrize("tt", [9, 10, 7])
def test_one_step_with_the_dp3d_limiter_active(case_c1, parity_state_c1, tt):
    """A physically valid state with one layer thinned to 5 % of its reference thickness inside a smooth blob (dp3d,
    vtheta_dp and the geopotential thickness scaled together): limiter_dp3d_k (prim_advance_mod.F90:1822-1931) acts in
    every stage of a full step, and the implicit solve runs on what it leaves.  Covers the limiter inside the DIRK
    kernel that does the DSS of the preceding CAAR stage (dirk_col_kernel<.., DSS = true>)."""
    from e3sm_b200.initial import make_c0
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = parity_state_c1.copy()
    nlev, K = hv.nlev, 8
    c = grid.cart.reshape(grid.nelem, 16, 3)
    fac = 1.0 - 0.95 * np.exp(-((c - c[17, 5]) ** 2).sum(-1) / 0.15)
    dphi = np.diff(st.phinh_i[:, 0], axis=1)
    st.dp3d[:, 0, K] *= fac
    st.vtheta_dp[:, 0, K] *= fac
    dphi[:, K] *= fac
    for k in range(nlev - 1, -1, -1):
        st.phinh_i[:, 0, k] = st.phinh_i[:, 0, k + 1] - dphi[:, k]
    for f in st.FIELDS:
        a = getattr(st, f)
        a[:, 0] = make_c0(grid, a[:, 0])
    st.phinh_i[:, 0, nlev] = phis
    st.set_all_timelevels(0)
    dp0 = (hv.hyai[1:] - hv.hyai[:-1]) * hv.ps0 + (hv.hybi[1:] - hv.hybi[:-1]) * hv.ps0
    assert (st.dp3d[:, 0, K] / dp0[K]).min() < 0.06 < p.dp3d_thresh
    p, o, so, d = make_pair(case_c1, st, tstep_type=tt)
    o.prim_advance_exp(p.tstep, 0, 1, 2)
    d.prim_advance_exp(p.tstep, 0, 1, 2)
    assert o.error_flag == 0
    assert (so.dp3d[:, 2, K] / dp0[K]).min() > 0.12          # the limiter has refilled the layer
    sg = d.get_state()
    for t in (2, 0, 1):
        assert_state_close(sg, so, t, 1e-9, f"limiter step tstep_type={tt} level {t}")
    d.close()


def test_bad_state_is_fatal(case_c1, parity_state_c1):
    """EOS bad state aborts in the reference (eos.F90:128-145); here: latched error, no abort in tests"""
    import ctypes
    from e3sm_b200 import lib
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = parity_state_c1.copy()
    st.dp3d[3, :, 5, 2] = -1.0
    L = lib.load()
    L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(0)))
    try:
        p, o, so, d = make_pair(case_c1, st)
        with pytest.raises(RuntimeError, match="EOS bad state"):
            d.compute_andor_apply_rhs(2, 1, 1, 10.0, 0.0, 1.0, 0.0, 1.0)
        d.close()
    finally:
        L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(1)))


# ---------------------------------------------------------------- rsplit = 0 (a13) and forcing (N2)
@pytest.mark.parametrize("np1,nm1,n0", [(2, 0, 1), (2, 1, 1), (0, 1, 1)])
@pytest.mark.parametrize("form,pgrad,eta_ave_w", [(0, 0, 0.25), (1, 1, 0.0), (2, 0, 1.0)])
def test_caar_vertically_eulerian(case_c1, parity_state_c1, np1, nm1, n0, form, pgrad, eta_ave_w):
    """rsplit = 0: eta_dot_dpdn, preq_vertadv_v1 and the theta/w/phi vertical advection terms
    (prim_advance_mod.F90:1227-1316, prim_si_mod.F90:94-148), accumulated derived%eta_dot_dpdn included."""
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = distinct_levels(parity_state_c1, phis, gradphis)
    p, o, so, d = make_pair(case_c1, st, theta_advect_form=form, pgrad_correction=pgrad, rsplit=0)
    args = (np1, nm1, n0, 37.5, eta_ave_w, 1.0, 0.0, 1.0)
    o.compute_andor_apply_rhs(*args)
    d.compute_andor_apply_rhs(*args)
    assert_state_close(d.get_state(), so, np1, TOL, f"caar rsplit=0 np1={np1} nm1={nm1} n0={n0} form={form}")
    vn0, eta, om = d.get_derived()
    assert rel_l2(vn0, o.vn0) <= TOL
    if eta_ave_w != 0:
        assert np.abs(o.eta_dot_dpdn).max() > 0
        assert rel_l2(eta, o.eta_dot_dpdn) <= 1e-9        # differences of prefix sums of div(v dp)
    d.close()


@pytest.mark.parametrize("tt,dt", [(9, 300.0), (10, 300.0), (5, 1.0)])     # tt5 is fully explicit: acoustic time step
def test_steps_vertically_eulerian(case_c1, parity_state_c1, tt, dt):
    """three dynamics steps with rsplit = 0 through prim_run_subcycle_c: one prim_step per call, no remap of the
    dynamics (prim_driver.cpp:79-83, vertical_remap with rsplit = 0)"""
    p, o, so, d = make_pair(case_c1, parity_state_c1.copy(), rsplit=0, tstep_type=tt, tstep=dt)
    nm1, n0, np1 = 0, 1, 2
    for call in range(3):
        o.vn0[:] = 0; o.omega_p[:] = 0; o.eta_dot_dpdn[:] = 0
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        assert o.error_flag == 0
        nm1, n0, np1 = n0, np1, nm1
        nstep, c_nm1, c_n0, c_np1 = d.prim_run_subcycle(p.tstep)
        assert nstep == call + 1 and (c_nm1, c_n0, c_np1) == (nm1, n0, np1)
    assert o.error_flag == 0
    assert_state_close(d.get_state(), so, n0, 1e-9, f"rsplit=0 steps tt{tt}")
    vn0, eta, om = d.get_derived()
    assert rel_l2(eta, o.eta_dot_dpdn) <= 1e-8
    d.close()


def test_forcing_through_prim_run_subcycle(case_c1, parity_state_c1):
    """push_forcing_to_c + ftype 0: prim_run_subcycle_c applies applyCAMforcing_dynamics to time level n0 with
    dt_remap = tstep*rsplit*qsplit before its dynamics steps (prim_driver.cpp:117-125, PA:517-541)"""
    p, grid, hv, _, phis, gradphis, refs = case_c1
    p, o, so, d = make_pair(case_c1, parity_state_c1.copy(), ftype=0)
    rng = np.random.default_rng(7)
    n, nlev = grid.nelem, hv.nlev
    from e3sm_b200.initial import make_c0
    fm = np.ascontiguousarray(1e-5 * make_c0(grid, rng.uniform(-1, 1, (n, nlev, 3, 16))))
    fvt = np.ascontiguousarray(1e-3 * so.vtheta_dp[:, 0] / 300.0 * make_c0(grid, rng.uniform(-1, 1, (n, nlev, 16))))
    fphi = np.ascontiguousarray(1e-3 * make_c0(grid, rng.uniform(-1, 1, (n, nlev + 1, 16))))
    d.push_forcing(fm, fvt, fphi)
    nm1, n0, np1 = 0, 1, 2
    for call in range(2):
        o.apply_forcing_dynamics(n0, p.tstep * p.rsplit * p.qsplit, fm, fvt, fphi)
        for r in range(p.rsplit * p.qsplit):
            o.prim_advance_exp(p.tstep, nm1, n0, np1)
            nm1, n0, np1 = n0, np1, nm1
        assert o.vertical_remap(n0, p.vert_remap_u_alg) == 0
        d.prim_run_subcycle(p.tstep)
    assert o.error_flag == 0
    assert_state_close(d.get_state(), so, n0, 1e-9, "forcing + prim_run_subcycle")
    # the forcing kernel alone
    o.apply_forcing_dynamics(n0, 123.0, fm, fvt, fphi)
    d.apply_forcing_dynamics(n0, 123.0)
    assert_state_close(d.get_state(), so, n0, 1e-9, "applyCAMforcing_dynamics")
    d.close()


# ---------------------------------------------------------------- boundary (b): the smaller entry points
def test_state_inside_array_of_element_structs(case_c1, parity_state_c1):
    """init_elements_states_strided_c: the plain theta-l build keeps elem(ie)%state inside element_t
    (element_mod.F90:32-144); base pointers of elem(1) + the struct size reach the same device state"""
    from e3sm_b200.dycore import ThetaLDycore
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = distinct_levels(parity_state_c1, phis, gradphis)
    n, nlev = grid.nelem, hv.nlev
    sizes = {"v": 3 * nlev * 32, "w_i": 3 * (nlev + 1) * 16, "vtheta_dp": 3 * nlev * 16,
             "phinh_i": 3 * (nlev + 1) * 16, "dp3d": 3 * nlev * 16}
    pad = 37                                     # other members of element_t between and after the state fields
    rec = sum(sizes.values()) + pad * (len(sizes) + 1)
    buf = np.full((n, rec), np.nan)
    off, o = {}, pad
    for f, sz in sizes.items():
        off[f] = o
        buf[:, o:o + sz] = getattr(st, f).reshape(n, sz)
        o += sz + pad
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    base = buf.ctypes.data
    d.set_state_strided(*[base + 8 * off[f] for f in ("v", "w_i", "vtheta_dp", "phinh_i", "dp3d")], rec * 8)
    out = d.get_state()
    for f in sizes:
        assert np.array_equal(getattr(out, f), getattr(st, f)), f
    d.close()


def test_ps_v_and_initialize_dp3d_from_ps(case_c1, parity_state_c1):
    """initialize_dp3d_from_ps_c (prim_driver.cpp:31-57) and the ps_v that cxx_push_results_to_f90 returns
    (hyai(1) ps0 + sum dp3d, vertremap_mod.F90:79-84)"""
    from e3sm_b200.dycore import ThetaLDycore
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = parity_state_c1.copy()
    n, nlev = grid.nelem, hv.nlev
    rng = np.random.default_rng(3)
    ps = np.ascontiguousarray(hv.ps0 * (1 + 0.02 * rng.uniform(-1, 1, (n, 3, 16))))
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    d.set_state(st, ps_v=ps)
    d.init_time_level(1, 2, 3)
    d.initialize_dp3d_from_ps()
    back = np.zeros_like(ps)
    out = d.get_state(ps_v=back)
    dA, dB = np.diff(hv.hyai), np.diff(hv.hybi)
    want = dA[None, :, None] * hv.ps0 + dB[None, :, None] * ps[:, 1][:, None, :]
    assert rel_l2(out.dp3d[:, 1], want) <= 1e-15
    assert np.array_equal(out.dp3d[:, 0], st.dp3d[:, 0])                 # only time level n0 is touched
    assert rel_l2(back[:, 1], ps[:, 1]) <= 1e-14                          # sum of the new layers = ps_v again
    assert rel_l2(back[:, 0], hv.hyai[0] * hv.ps0 + st.dp3d[:, 0].sum(axis=1)) <= 1e-15
    d.close()


# ---------------------------------------------------------------- the real C2 / C5 sphere
@pytest.mark.parametrize("cfg", ["C2", "C5"])
def test_one_step_on_the_real_ne30_sphere(cfg):
    """BASELINE config 2 (and 5: hypervis_subcycle 3, nu_div 2.5e15) at its real size, ne30np4 L72, 5400 elements:
    one full dynamics step from the JW state advanced two steps (so w != 0), GPU vs oracle, rel-L2 <= 1e-10."""
    from e3sm_b200.dycore import build_case
    from e3sm_b200.params import config
    p = config(cfg)
    assert p.ne == 30 and p.nlev == 72
    case = (p,) + build_case(p)
    p, o, so, d = make_pair(case, case[3].copy())
    nm1, n0, np1 = 0, 1, 2
    for _ in range(3):
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        d.prim_advance_exp(p.tstep, nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    assert o.error_flag == 0
    sg = d.get_state()
    for f in ("v", "vtheta_dp", "phinh_i", "dp3d"):
        r = rel_l2(getattr(sg, f)[:, n0], getattr(so, f)[:, n0])
        assert r <= TOL, (cfg, f, r)
    dw = float(np.abs(sg.w_i[:, n0] - so.w_i[:, n0]).max())
    assert dw <= 1e-10 * max(1.0, float(np.abs(so.w_i[:, n0]).max())), dw
    d.close()


def test_dss_c_of_a_user_field(case_c1):
    """dss_c: node sums of an arbitrary field in the Fortran unpack order = the reference DSS of geometry.dss_numpy"""
    from e3sm_b200.dycore import ThetaLDycore
    from e3sm_b200.geometry import dss_numpy
    p, grid, hv, st, phis, gradphis, refs = case_c1
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    rng = np.random.default_rng(11)
    for nl in (1, 7, 6 * hv.nlev + 1):
        f = rng.standard_normal((grid.nelem, nl, 16))
        want = dss_numpy(f, grid.dss_src)
        got = d.dss(f.copy())
        assert np.array_equal(got, want), nl
    d.close()


def test_field_extrema_match_numpy(case_c1, parity_state_c1):
    """thetal_b200_field_extrema: the min/max lines of prim_printstate (prim_state_mod.F90:219-330)"""
    from e3sm_b200.dycore import ThetaLDycore
    from e3sm_b200.geometry import G, KAPPA, P0, RGAS
    p, grid, hv, _, phis, gradphis, refs = case_c1
    st = distinct_levels(parity_state_c1, phis, gradphis)
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    d.set_state(st)
    tl = 2
    ex = d.field_extrema(tl)
    dp, th, phi = st.dp3d[:, tl], st.vtheta_dp[:, tl], st.phinh_i[:, tl]
    dphi = phi[:, 1:] - phi[:, :-1]
    poe = RGAS * th / (-dphi)
    exner = (P0 * (poe / P0) ** (1.0 / (1.0 - KAPPA))) / poe
    want = {"u": st.v[:, tl, :, 0], "v": st.v[:, tl, :, 1], "w": st.w_i[:, tl], "theta": th / dp, "T": th / dp * exner,
            "dz": -dphi / G, "dp3d": dp}
    for k, a in want.items():
        lo, hi = ex[k]
        assert abs(lo - a.min()) <= 1e-12 * max(1.0, abs(a.min())) and abs(hi - a.max()) <= 1e-12 * max(1.0, abs(a.max())), (k, ex[k], a.min(), a.max())
    assert ex["u"][0] < 0 < ex["u"][1]            # both signs go through the ordered keys
    d.close()
