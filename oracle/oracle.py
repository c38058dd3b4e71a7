"""TEST INFRASTRUCTURE ONLY: ctypes driver of the CPU oracle (oracle/thetal_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Parity status: unpinned against the Fortran executable (see thetal_oracle.h).
"""
import ctypes
import os
from ctypes import POINTER, Structure, byref, c_double, c_int

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_DP = POINTER(c_double)
_IP = POINTER(c_int)


class OrcCtx(Structure):
    _fields_ = [
        ("nelem", c_int), ("nlev", c_int),
        ("rsplit", c_int), ("theta_advect_form", c_int), ("pgrad_correction", c_int),
        ("hypervis_subcycle", c_int), ("hypervis_subcycle_tom", c_int),
        ("nlev_tom", c_int), ("hypervis_scaling_on", c_int), ("hv_theta_correction", c_int),
        ("nu", c_double), ("nu_div", c_double), ("nu_p", c_double), ("nu_s", c_double),
        ("nu_top", c_double), ("dp3d_thresh", c_double), ("vtheta_thresh", c_double),
        ("rearth", c_double), ("ps0", c_double),
        ("hyai", _DP), ("hybi", _DP), ("hyam", _DP), ("hybm", _DP), ("dp0", _DP), ("nu_scale_top", _DP),
        ("dvv", _DP), ("mp", _DP),
        ("D", _DP), ("Dinv", _DP),
        ("metdet", _DP), ("rmetdet", _DP), ("spheremp", _DP), ("rspheremp", _DP), ("fcor", _DP),
        ("metinv", _DP), ("tensorvisc", _DP), ("vec_sph2cart", _DP), ("gradphis", _DP), ("phis", _DP),
        ("theta_ref", _DP), ("dp_ref", _DP), ("phi_ref", _DP),
        ("dss_src", _IP),
        ("v", _DP), ("w_i", _DP), ("vtheta_dp", _DP), ("phinh_i", _DP), ("dp3d", _DP),
        ("vn0", _DP), ("eta_dot_dpdn", _DP), ("omega_p", _DP),
        ("scratch1", _DP), ("scratch2", _DP),
        ("max_itercnt", c_int), ("max_deltaerr", c_double), ("max_reserr", c_double),
        ("nthreads", c_int), ("error_flag", c_int),
    ]


def _load(fast=False):
    name = "libthetal_oracle_fast.so" if fast else "libthetal_oracle.so"
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        raise RuntimeError(f"{path} not built: run `make -C oracle all`")
    lib = ctypes.CDLL(path)
    return lib


def _p(a):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_DP)


class Oracle:
    """CPU restatement of the theta-l step on flat Fortran-layout state arrays (time levels 0-based)."""

    def __init__(self, grid, hv, params, phis, gradphis, ref_states, state, nthreads=1, fast=False,
                 rsplit=None):
        self.lib = _load(fast)
        self.grid, self.hv, self.params = grid, hv, params
        self.state = state
        n, nlev = grid.nelem, hv.nlev
        self._keep = []
        k = self._keep.append
        c = OrcCtx()
        c.nelem, c.nlev = n, nlev
        c.rsplit = params.rsplit if rsplit is None else rsplit
        c.theta_advect_form = params.theta_advect_form
        c.pgrad_correction = params.pgrad_correction
        c.hypervis_subcycle = params.hypervis_subcycle
        c.hypervis_subcycle_tom = params.hypervis_subcycle_tom
        nst, nlev_tom = hv.nu_scale_top(params.tom_sponge_start)
        c.nlev_tom = nlev_tom
        c.hypervis_scaling_on = 1 if params.hypervis_scaling > 0 else 0
        c.hv_theta_correction = 0
        c.nu, c.nu_div, c.nu_p, c.nu_s, c.nu_top = params.nu, params.nu_div, params.nu_p, params.nu_s, params.nu_top
        c.dp3d_thresh, c.vtheta_thresh = params.dp3d_thresh, params.vtheta_thresh
        c.rearth, c.ps0 = grid.rearth, hv.ps0
        self.nst = np.ascontiguousarray(nst)
        self.dvv = np.ascontiguousarray(grid.dvv)
        self.mp = np.ascontiguousarray(grid.mp)
        c.hyai, c.hybi, c.hyam, c.hybm, c.dp0 = _p(hv.hyai), _p(hv.hybi), _p(hv.hyam), _p(hv.hybm), _p(hv.dp0)
        c.nu_scale_top, c.dvv, c.mp = _p(self.nst), _p(self.dvv), _p(self.mp)
        c.D, c.Dinv, c.metinv = _p(grid.D), _p(grid.Dinv), _p(grid.metinv)
        c.metdet, c.rmetdet, c.spheremp, c.rspheremp, c.fcor = (_p(grid.metdet), _p(grid.rmetdet),
                                                                _p(grid.spheremp), _p(grid.rspheremp), _p(grid.fcor))
        c.tensorvisc, c.vec_sph2cart = _p(grid.tensorvisc), _p(grid.vec_sph2cart)
        self.gradphis = np.ascontiguousarray(gradphis)
        self.phis = np.ascontiguousarray(phis)
        self.refs = [np.ascontiguousarray(a) for a in ref_states]
        c.gradphis, c.phis = _p(self.gradphis), _p(self.phis)
        c.theta_ref, c.dp_ref, c.phi_ref = _p(self.refs[0]), _p(self.refs[1]), _p(self.refs[2])
        self.dss = np.ascontiguousarray(grid.dss_src, dtype=np.int32)
        c.dss_src = self.dss.ctypes.data_as(_IP)
        c.v, c.w_i, c.vtheta_dp, c.phinh_i, c.dp3d = (_p(state.v), _p(state.w_i), _p(state.vtheta_dp),
                                                      _p(state.phinh_i), _p(state.dp3d))
        self.vn0 = np.zeros((n, nlev, 2, 16))
        self.eta_dot_dpdn = np.zeros((n, nlev + 1, 16))
        self.omega_p = np.zeros((n, nlev, 16))
        c.vn0, c.eta_dot_dpdn, c.omega_p = _p(self.vn0), _p(self.eta_dot_dpdn), _p(self.omega_p)
        self.s1 = np.zeros(n * 16 * (6 * nlev + 2))
        self.s2 = np.zeros(n * 16 * (6 * nlev + 2))
        c.scratch1, c.scratch2 = _p(self.s1), _p(self.s2)
        c.nthreads = nthreads
        self.c = c

    def compute_andor_apply_rhs(self, np1, nm1, n0, dt2, eta_ave_w, scale1, scale2, scale3):
        self.lib.orc_caar(byref(self.c), c_int(np1), c_int(nm1), c_int(n0), c_double(dt2), c_double(eta_ave_w),
                          c_double(scale1), c_double(scale2), c_double(scale3))

    def caar_pre(self, np1, nm1, n0, dt2, eta_ave_w, scale1, scale2, scale3):
        self.lib.orc_caar_pre(byref(self.c), c_int(np1), c_int(nm1), c_int(n0), c_double(dt2), c_double(eta_ave_w),
                              c_double(scale1), c_double(scale2), c_double(scale3))

    def compute_stage_value_dirk(self, nm1, alphadt_nm1, n0, alphadt_n0, np1, dt2):
        it, err = c_int(0), c_double(0)
        self.lib.orc_dirk(byref(self.c), c_int(nm1), c_double(alphadt_nm1), c_int(n0), c_double(alphadt_n0),
                          c_int(np1), c_double(dt2), byref(it), byref(err))
        return it.value, err.value

    def advance_hypervis(self, nt, dt2, eta_ave_w):
        self.lib.orc_advance_hypervis(byref(self.c), c_int(nt), c_double(dt2), c_double(eta_ave_w))

    def vertical_remap(self, np1, remap_alg=1):
        """vertremap_mod.F90:23-171 (rsplit>0, qsize=0) with remap_Q_ppm; returns 1 on a negative layer."""
        return int(self.lib.orc_vertical_remap(byref(self.c), c_int(np1), c_int(remap_alg)))

    def prim_advance_exp(self, dt, nm1, n0, np1):
        self.lib.orc_prim_advance_exp(byref(self.c), c_int(self.params.tstep_type), c_double(dt), c_int(nm1),
                                      c_int(n0), c_int(np1), c_int(self.params.qsplit))

    def apply_forcing_dynamics(self, np1, dt, fm, fvtheta, fphi):
        """applyCAMforcing_dynamics (PA:517-541); fm [ie][k][3][16], fvtheta [ie][k][16], fphi [ie][k<=nlev][16]."""
        self.lib.orc_apply_forcing_dynamics(byref(self.c), c_int(np1), c_double(dt), _p(fm), _p(fvtheta), _p(fphi))

    @property
    def error_flag(self):
        return self.c.error_flag
