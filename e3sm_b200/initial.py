"""Init-time producers of hot-path inputs: initial conditions, reference states, gradphis.

Host-side (numpy) restatement of the Fortran init code, H/ = components/homme/:
  jw_baroclinic            H/src/test_src/baroclinic_inst_mod.F90:57-197
  set_thermostate          H/src/theta-l/share/element_ops.F90:452-523
  tests_finalize           H/src/theta-l/share/element_ops.F90:716-760
  phi_from_eos             H/src/theta-l/share/eos.F90:228-303
  model_init2 (gradphis)   H/src/theta-l/share/model_init_mod.F90:36-101
  reference states         H/src/theta-l/share/element_ops.F90:763-864
  random physical states   H/src/theta-l_kokkos/cxx/ElementsState.cpp:112-241 (recipe of the unit tests)

State arrays use the flat Fortran layout of H/src/theta-l_kokkos/element_state.F90:35-67:
  v[nelem,3,nlev,2,16]  w_i,phinh_i[nelem,3,nlevp,16]  vtheta_dp,dp3d[nelem,3,nlev,16]
"""
import numpy as np

from .geometry import (CP, DD_PI, G, KAPPA, OMEGA, P0, REARTH, RGAS, TREF, dss_numpy)

TREF_LAPSE_RATE = 0.0065   # element_ops.F90:80


class State:
    """The six prognostics at three time levels + derived accumulators (element_state.F90:31-84)."""

    def __init__(self, nelem, nlev):
        self.nelem, self.nlev = nelem, nlev
        self.v = np.zeros((nelem, 3, nlev, 2, 16))
        self.w_i = np.zeros((nelem, 3, nlev + 1, 16))
        self.vtheta_dp = np.zeros((nelem, 3, nlev, 16))
        self.phinh_i = np.zeros((nelem, 3, nlev + 1, 16))
        self.dp3d = np.zeros((nelem, 3, nlev, 16))

    FIELDS = ("v", "w_i", "vtheta_dp", "phinh_i", "dp3d")

    def copy(self):
        s = State.__new__(State)
        s.nelem, s.nlev = self.nelem, self.nlev
        for f in self.FIELDS:
            setattr(s, f, getattr(self, f).copy())
        return s

    def set_all_timelevels(self, tl_src=0):
        for f in self.FIELDS:
            a = getattr(self, f)
            for t in range(3):
                if t != tl_src:
                    a[:, t] = a[:, tl_src]


def phi_from_eos(hv, phis, vtheta_dp, dp):
    """eos.F90:286-302. phis[e,16], vtheta_dp,dp[e,nlev,16] -> phi_i[e,nlevp,16]."""
    ne_, nlev = dp.shape[0], dp.shape[1]
    p_i = np.empty((ne_, nlev + 1, 16))
    p_i[:, 0] = hv.hyai[0] * hv.ps0
    for k in range(nlev):
        p_i[:, k + 1] = p_i[:, k] + dp[:, k]
    p = (p_i[:, 1:] + p_i[:, :-1]) / 2
    phi_i = np.empty((ne_, nlev + 1, 16))
    phi_i[:, nlev] = phis
    for k in range(nlev - 1, -1, -1):
        phi_i[:, k] = phi_i[:, k + 1] + (RGAS * vtheta_dp[:, k] * (p[:, k] / P0) ** (KAPPA - 1)) / P0
    return phi_i


def gradient_sphere(grid, s):
    """derivative_mod.F90:410-450 vectorised. s[e,...,16] -> ds[e,...,2,16]."""
    dvv = grid.dvv                       # [l][i] = Dvv(i,l)
    sh = s.shape
    s4 = s.reshape(sh[0], -1, 4, 4)      # [e,L,j,i]
    v1 = np.einsum("li,enji->enjl", dvv, s4) / grid.rearth      # v1(l,j)=sum_i Dvv(i,l) s(i,j)
    v2 = np.einsum("li,enij->enlj", dvv, s4) / grid.rearth      # v2(j,l)=sum_i Dvv(i,l) s(j,i)
    v1 = v1.reshape(sh[0], -1, 16)
    v2 = v2.reshape(sh[0], -1, 16)
    Dinv = grid.Dinv                     # [e,b,a,p] = Dinv(p,a,b)
    ds1 = Dinv[:, 0, 0, None, :] * v1 + Dinv[:, 0, 1, None, :] * v2
    ds2 = Dinv[:, 1, 0, None, :] * v1 + Dinv[:, 1, 1, None, :] * v2
    out = np.stack([ds1, ds2], axis=2)   # [e,L,2,16]
    return out.reshape(sh[:-1] + (2, 16))


def make_c0(grid, f):
    """make_C0 / make_C0_vector (H/src/share/viscosity_base.F90): spheremp-weighted DSS average."""
    sh = f.shape
    w = f.reshape(sh[0], -1, 16) * grid.spheremp[:, None, :]
    w = dss_numpy(w, grid.dss_src) * grid.rspheremp[:, None, :]
    return w.reshape(sh)


def reference_states(hv, phis, hv_ref_profiles=0):
    """initialize_reference_states + set_theta_ref (element_ops.F90:763-864) -> theta_ref, dp_ref, phi_ref."""
    nlev = hv.nlev
    ps_ref = hv.ps0 * np.exp(-phis / (RGAS * TREF))                                  # [e,16]
    dp_ref = ((hv.hyai[1:] - hv.hyai[:-1]) * hv.ps0)[None, :, None] + \
             (hv.hybi[1:] - hv.hybi[:-1])[None, :, None] * ps_ref[:, None, :]

    def theta_ref_of(dp, linear):
        T1 = TREF_LAPSE_RATE * TREF * CP / G
        T0 = TREF - T1
        p_i = np.empty((dp.shape[0], nlev + 1, 16))
        p_i[:, 0] = hv.hyai[0] * hv.ps0
        for k in range(nlev):
            p_i[:, k + 1] = p_i[:, k] + dp[:, k]
        exner = ((p_i[:, :-1] + p_i[:, 1:]) / (2 * P0)) ** KAPPA
        if linear:
            return T0 + T0 * (1 - exner) + T1
        return T0 / exner + T1

    theta_ref = theta_ref_of(dp_ref, 0)
    phi_ref = phi_from_eos(hv, phis, theta_ref * dp_ref, dp_ref)
    if hv_ref_profiles == 0:
        theta_ref = np.zeros_like(theta_ref)
        dp_ref = np.zeros_like(dp_ref)
    elif hv_ref_profiles == 6:
        theta_ref = theta_ref_of(dp_ref, 1)
    return theta_ref, dp_ref, phi_ref


def jw_phis(grid):
    """Surface geopotential of the JW test alone (baroclinic_inst_mod.F90:168-176) -> phis[e,16]."""
    u0, eta_s, eta_0 = 35.0, 1.0, 0.252
    snlat, cslat = np.sin(grid.lat), np.cos(grid.lat)
    trm2 = -2.0 * snlat ** 6 * (cslat ** 2 + 1.0 / 3.0) + 10.0 / 63.0
    trm4 = (1.60 * cslat ** 3 * (snlat ** 2 + 2.0 / 3.0) - DD_PI * 0.25) * grid.rearth * OMEGA
    s1 = u0 * (np.cos((eta_s - eta_0) * DD_PI * 0.5)) ** 1.5
    return s1 * (trm2 * s1 + trm4)


def jw_baroclinic(grid, hv, u_perturb=1.0):
    """Jablonowski-Williamson baroclinic wave (baroclinic_inst_mod.F90:57-197) -> (State, phis[e,16])."""
    nelem, nlev = grid.nelem, hv.nlev
    u0, t0, gama, ddt, eta_t, eta_s, eta_0 = 35.0, 288.0, 0.005, 4.8e5, 0.2, 1.0, 0.252
    eta = hv.etam
    etv = (eta - eta_0) * DD_PI * 0.5
    latc, lonc = DD_PI * (2.0 / 9.0), DD_PI * (1.0 / 9.0)
    lat, lon = grid.lat, grid.lon                                                     # [e,16]
    snlat, cslat = np.sin(lat), np.cos(lat)
    aa = np.sin(latc) * snlat + np.cos(latc) * cslat * np.cos(lon - lonc)
    rc = 10.0 * np.arccos(np.clip(aa, -1.0, 1.0))
    st = State(nelem, nlev)
    v1 = u0 * (np.cos(etv)[None, :, None]) ** 1.5 * (np.sin(2.0 * lat) ** 2)[:, None, :] \
        + u_perturb * np.exp(-rc * rc)[:, None, :]
    st.v[:, 0, :, 0, :] = v1
    tbar = np.where(eta <= eta_t, t0 * eta ** (RGAS * gama / G) + ddt * (np.maximum(eta_t - eta, 0.0)) ** 5,
                    t0 * eta ** (RGAS * gama / G))
    trm1 = 0.75 * (eta * DD_PI * u0 / RGAS) * np.sin(etv) * np.sqrt(np.cos(etv))      # [k]
    trm2 = -2.0 * snlat ** 6 * (cslat ** 2 + 1.0 / 3.0) + 10.0 / 63.0                 # [e,16]
    trm3 = 2.0 * u0 * np.cos(etv) ** 1.5                                              # [k]
    trm4 = (1.60 * cslat ** 3 * (snlat ** 2 + 2.0 / 3.0) - DD_PI * 0.25) * grid.rearth * OMEGA
    temperature = tbar[None, :, None] + trm1[None, :, None] * (trm2[:, None, :] * trm3[None, :, None] + trm4[:, None, :])
    s1 = u0 * (np.cos((eta_s - eta_0) * DD_PI * 0.5)) ** 1.5
    phis = s1 * (trm2 * s1 + trm4)
    ps = np.full((nelem, 16), P0)
    # set_thermostate
    dp = ((hv.hyai[1:] - hv.hyai[:-1]) * hv.ps0)[None, :, None] + (hv.hybi[1:] - hv.hybi[:-1])[None, :, None] * ps[:, None, :]
    pi_i = np.empty((nelem, nlev + 1, 16))
    pi_i[:, 0] = hv.hyai[0] * hv.ps0
    for k in range(nlev):
        pi_i[:, k + 1] = pi_i[:, k] + dp[:, k]
    p = pi_i[:, :-1] + dp / 2
    st.vtheta_dp[:, 0] = dp * temperature * (p / P0) ** (-KAPPA)
    st.dp3d[:, 0] = dp
    # tests_finalize: hydrostatic phi, w=0, copy to all time levels
    st.phinh_i[:, 0] = phi_from_eos(hv, phis, st.vtheta_dp[:, 0], st.dp3d[:, 0])
    st.w_i[:] = 0.0
    st.set_all_timelevels(0)
    return st, phis


def model_init2(grid, hv, st, phis, hv_ref_profiles=0, n0=0, gradphis=None):
    """model_init2 (model_init_mod.F90:36-101): gradphis, w b.c., phinh_i(nlevp)=phis, reference states.
    ``gradphis`` may be passed in (a rank's slice of the globally DSS'd field)."""
    if gradphis is None:
        gradphis = make_c0(grid, gradient_sphere(grid, phis))                         # [e,2,16]
    nlev = hv.nlev
    st.w_i[:, n0, nlev] = (st.v[:, n0, nlev - 1, 0] * gradphis[:, 0] + st.v[:, n0, nlev - 1, 1] * gradphis[:, 1]) / G
    for t in range(3):
        st.phinh_i[:, t, nlev] = phis
    theta_ref, dp_ref, phi_ref = reference_states(hv, phis, hv_ref_profiles)
    return gradphis, theta_ref, dp_ref, phi_ref


def random_state(grid, hv, phis, seed=1, all_levels=True):
    """Random but physical state in the spirit of ElementsState::randomize (ElementsState.cpp:112-241):
    v,w ~ U(1/64,64)-ish, vtheta_dp ~ U(100,1000), dp3d from sorted pressures, phinh_i decreasing.
    Used for kernel-vs-oracle tests only (triggers limiters); same values in all 3 time levels
    unless all_levels=False (then each level is drawn independently)."""
    rng = np.random.default_rng(seed)
    nelem, nlev = grid.nelem, hv.nlev
    st = State(nelem, nlev)
    for t in range(3):
        st.v[:, t] = rng.uniform(-30.0, 30.0, st.v[:, t].shape)
        st.w_i[:, t] = rng.uniform(-2.0, 2.0, st.w_i[:, t].shape)
        st.vtheta_dp[:, t] = rng.uniform(100.0, 1000.0, st.vtheta_dp[:, t].shape)
        ptop, pbot = hv.hyai[0] * hv.ps0, 1000.0 + hv.ps0
        pr = np.sort(rng.uniform(ptop, pbot, (nelem, nlev - 1, 16)), axis=1)
        pr = np.concatenate([np.full((nelem, 1, 16), ptop), pr, np.full((nelem, 1, 16), pbot)], axis=1)
        st.dp3d[:, t] = np.diff(pr, axis=1)
        ph = np.sort(rng.uniform(1.001, 100.0, (nelem, nlev, 16)), axis=1)[:, ::-1]
        base = np.maximum(np.abs(phis), 100.0)[:, None, :]
        st.phinh_i[:, t, :nlev] = phis[:, None, :] + base * (ph - 1.0) * 50.0
        st.phinh_i[:, t, nlev] = phis
        if not all_levels:
            continue
    if all_levels:
        st.set_all_timelevels(0)
    return st
