"""Host-side mirror of the reference interface for the theta-l step, over the C ABI.

``ThetaLDycore`` drives libthetal_b200.so exactly as HOMME's Fortran driver for the
theta-l_kokkos target would (H/src/theta-l_kokkos/prim_driver_mod.F90:60-330 ->
prim_init_*: init_simulation_params_c, init_hvcoord_c, init_elements_c, connectivity,
init_reference_element_c, init_elements_2d_c, init_geopotential_c, init_functors_c,
init_elements_states_c, init_reference_states_c), then exposes the reference routine names:
prim_advance_exp, compute_andor_apply_rhs, compute_stage_value_dirk, advance_hypervis,
prim_run_subcycle.  Time levels are 0-based here.  There is no CPU fallback.
"""
import ctypes
from ctypes import POINTER, byref, c_char_p, c_double, c_int, c_void_p

import numpy as np

from . import lib as _libmod
from .initial import State

_DIR_OF = {"W": 0, "E": 1, "S": 2, "N": 3, "SW": 4, "SE": 5, "NW": 6, "NE": 7}


def _ci(x):
    return byref(c_int(int(x)))


def _cd(x):
    return byref(c_double(float(x)))


def _pp(a):
    """Fortran type(c_ptr) passed by reference -> pointer to pointer."""
    assert a.flags["C_CONTIGUOUS"] and a.dtype == np.float64
    return byref(c_void_p(a.ctypes.data))


def connections(grid):
    """GridEdge list of (e1, dir1, e2, dir2), 0-based elements, dir codes W,E,S,N,SW,SE,NW,NE = 0..7
    (what init_cxx_connectivity feeds add_connection, prim_cxx_driver_base.F90:137-197)."""
    src = grid.dss_src
    nelem = grid.nelem
    out = []

    def edge_of(p):
        i, j = p % 4, p // 4
        r = []
        if i == 0: r.append(0)
        if i == 3: r.append(1)
        if j == 0: r.append(2)
        if j == 3: r.append(3)
        return r

    corner_dir = {0: 4, 3: 5, 12: 6, 15: 7}
    # edge connections: use the two interior points of each edge
    probes = {0: (4, 8), 1: (7, 11), 2: (1, 2), 3: (13, 14)}
    for e in range(nelem):
        edge_nbr = {}
        for d, (pa, pb) in probes.items():
            sa = int(src[e, pa, 0])
            sb = int(src[e, pb, 0])
            e2 = sa // 16
            assert sb // 16 == e2
            d2 = (set(edge_of(sa % 16)) & set(edge_of(sb % 16))).pop()
            out.append((e, d, e2, d2))
            edge_nbr[d] = e2
        for p, d in corner_dir.items():
            adj = {edge_nbr[x] for x in edge_of(p)}
            for j in range(3):
                s = int(src[e, p, j])
                if s >= 0 and s // 16 not in adj:
                    out.append((e, d, s // 16, corner_dir[s % 16]))
    return out


class ThetaLDycore:
    """One rank <-> one GPU; state lives on the device between calls."""

    def __init__(self, grid, hv, params, phis, gradphis, ref_states, device=0):
        self.L = _libmod.load()
        L = self.L
        self.grid, self.hv, self.params = grid, hv, params
        self.nelem, self.nlev = grid.nelem, hv.nlev
        L.finalize_hommexx_session()
        L.thetal_b200_init_comm(_ci(0), _ci(1), _ci(device), None)
        L.initialize_hommexx_session()
        self._check()
        L.thetal_b200_set_nlev(_ci(hv.nlev))
        p = params
        name = c_char_p(b"jw_baroclinic")
        L.init_simulation_params_c(
            _ci(1), _ci(9), _ci(p.rsplit), _ci(p.qsplit), _ci(p.tstep_type), _ci(0), _ci(0),
            _cd(p.nu), _cd(p.nu_p), _cd(0.0), _cd(p.nu_s), _cd(p.nu_div), _cd(p.nu_top),
            _ci(p.hypervis_order), _ci(p.hypervis_subcycle), _ci(p.hypervis_subcycle_tom),
            _cd(p.hypervis_scaling), _cd(0.0), _cd(0.0),
            _ci(-1), _ci(p.theta_advect_form), _ci(0), _ci(0), _ci(1), _ci(0),
            _ci(0), _ci(0), byref(name),
            _ci(p.rsplit), _ci(p.qsplit), _cd(grid.rearth), _cd(1.0 / grid.rearth), _ci(1),
            _ci(p.pgrad_correction), _cd(p.dp3d_thresh), _cd(p.vtheta_thresh), _ci(0), _ci(0),
            _cd(p.tom_sponge_start))
        L.init_hvcoord_c(_cd(hv.ps0), _pp(hv.hyam), _pp(hv.hyai), _pp(hv.hybm), _pp(hv.hybi))
        L.init_elements_c(_ci(self.nelem))
        L.init_time_level_c(_ci(1), _ci(2), _ci(3), _ci(0), _ci(0))
        # connectivity (all on rank 1 / pid 1, max_corner_elem = 1)
        L.init_connectivity(_ci(self.nelem), _ci(1))
        add = L.add_connection
        one = _ci(1)
        for (e1, d1, e2, d2) in connections(grid):
            add(_ci(e1 + 1), _ci(int(grid.gnum[e1]) + 1), _ci(d1 + 1), one,
                _ci(e2 + 1), _ci(int(grid.gnum[e2]) + 1), _ci(d2 + 1), one)
        L.finalize_connectivity()
        dvv = np.ascontiguousarray(grid.dvv)
        mp = np.ascontiguousarray(grid.mp)
        L.init_reference_element_c(_pp(dvv), _pp(mp))
        cart = np.zeros(48)
        ll = np.zeros(32)
        gradphis = np.ascontiguousarray(gradphis)
        phis = np.ascontiguousarray(phis)
        for e in range(self.nelem):
            L.init_elements_2d_c(_ci(e), _pp(grid.D[e]), _pp(grid.Dinv[e]), _pp(grid.fcor[e]),
                                 _pp(grid.spheremp[e]), _pp(grid.rspheremp[e]), _pp(grid.metdet[e]),
                                 _pp(grid.metinv[e]), _pp(grid.tensorvisc[e]), _pp(grid.vec_sph2cart[e]),
                                 cart.ctypes.data_as(POINTER(c_double)), ll.ctypes.data_as(POINTER(c_double)),
                                 _pp(grid.tensorvisc[e]))
            L.init_geopotential_c(_ci(e), _pp(phis[e]), _pp(gradphis[e]))
        L.init_functors_c(_ci(1))
        L.init_boundary_exchanges_c()
        theta_ref, dp_ref, phi_ref = (np.ascontiguousarray(a) for a in ref_states)
        nst, self.nlev_tom = hv.nu_scale_top(p.tom_sponge_start)
        self.nu_scale_top = nst
        L.init_reference_states_c(_pp(theta_ref), _pp(dp_ref), _pp(phi_ref), _pp(nst))
        self._check()

    # ------------------------------------------------------------------ helpers
    def _check(self):
        if self.L.thetal_b200_last_error() != 0:
            raise RuntimeError("libthetal_b200: " + self.L.thetal_b200_last_error_msg().decode())

    def set_state(self, st):
        """init_elements_states_c: host (flat Fortran layout) -> device."""
        self.L.init_elements_states_c(_pp(st.v), _pp(st.w_i), _pp(st.vtheta_dp), _pp(st.phinh_i),
                                      _pp(st.dp3d), None, None)
        self._check()

    def get_state(self, st=None):
        """cxx_push_results_to_f90: device -> host."""
        if st is None:
            st = State(self.nelem, self.nlev)
        self.L.cxx_push_results_to_f90(_pp(st.v), _pp(st.w_i), _pp(st.vtheta_dp), _pp(st.phinh_i),
                                       _pp(st.dp3d), None, None, None, None)
        self._check()
        return st

    def get_derived(self):
        n, nlev = self.nelem, self.nlev
        vn0 = np.zeros((n, nlev, 2, 16))
        eta = np.zeros((n, nlev + 1, 16))
        om = np.zeros((n, nlev, 16))
        self.L.thetal_b200_get_derived(c_void_p(vn0.ctypes.data), c_void_p(eta.ctypes.data), c_void_p(om.ctypes.data))
        return vn0, eta, om

    def zero_derived(self):
        self.L.thetal_b200_zero_derived()

    # ------------------------------------------------------------------ reference routine names
    def compute_andor_apply_rhs(self, np1, nm1, n0, dt2, eta_ave_w, scale1, scale2, scale3):
        self.L.caar_c(_ci(np1), _ci(nm1), _ci(n0), _cd(dt2), _cd(eta_ave_w), _cd(scale1), _cd(scale2), _cd(scale3))
        self._check()

    def compute_stage_value_dirk(self, nm1, alphadt_nm1, n0, alphadt_n0, np1, dt2):
        it = c_int(0)
        err = c_double(0)
        self.L.dirk_c(_ci(nm1), _cd(alphadt_nm1), _ci(n0), _cd(alphadt_n0), _ci(np1), _cd(dt2), byref(it), byref(err))
        self._check()
        return it.value, err.value

    def advance_hypervis(self, nt, dt2, eta_ave_w):
        self.L.advance_hypervis_c(_ci(nt), _cd(dt2), _cd(eta_ave_w))
        self._check()

    def prim_advance_exp(self, dt, nm1, n0, np1):
        self.L.prim_advance_exp_c(_cd(dt), _ci(nm1), _ci(n0), _ci(np1), _ci(0), _ci(0))
        self._check()

    def prim_run_subcycle(self, dt, nstep, nm1, n0, np1):
        """1-based time levels like the Fortran caller; returns updated (nstep, nm1, n0, np1)."""
        a, b, c, d = c_int(nstep), c_int(nm1), c_int(n0), c_int(np1)
        self.L.prim_run_subcycle_c(_cd(dt), byref(a), byref(b), byref(c), byref(d), _ci(0), _ci(0))
        self._check()
        return a.value, b.value, c.value, d.value

    def time_steps(self, dt, nsteps, nm1, n0, np1):
        """Run nsteps dynamics steps back to back; returns (device ms, nm1, n0, np1)."""
        b, c, d = c_int(nm1), c_int(n0), c_int(np1)
        ms = self.L.thetal_b200_time_steps(_cd(dt), _ci(nsteps), byref(b), byref(c), byref(d))
        self._check()
        return ms, b.value, c.value, d.value

    def launch_count(self):
        return int(self.L.thetal_b200_launch_count())

    def close(self):
        self.L.finalize_hommexx_session()


def build_case(params, hv_ref_profiles=None):
    """Geometry + JW initial state + reference states for a SimParams config (host side)."""
    from .geometry import CubedSphere
    from .hvcoord import HVCoord
    from .initial import jw_baroclinic, model_init2
    grid = CubedSphere(params.ne)
    hv = HVCoord.from_table(params.nlev)
    st, phis = jw_baroclinic(grid, hv, params.u_perturb)
    hvr = params.hv_ref_profiles if hv_ref_profiles is None else hv_ref_profiles
    gradphis, theta_ref, dp_ref, phi_ref = model_init2(grid, hv, st, phis, hvr)
    return grid, hv, st, phis, gradphis, (theta_ref, dp_ref, phi_ref)
