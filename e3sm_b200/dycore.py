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


def connections(grid, e0=0, e1=None):
    """GridEdge list of (e1, dir1, e2, dir2) for the elements e0 <= e1 < e1 of the storage order,
    0-based, dir codes W,E,S,N,SW,SE,NW,NE = 0..7 (what init_cxx_connectivity feeds add_connection,
    prim_cxx_driver_base.F90:137-197)."""
    src = grid.dss_src
    e1 = grid.nelem if e1 is None else e1
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
    for e in range(e0, e1):
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
    """One rank <-> one GPU; state lives on the device between calls.

    Multi-GPU: pass the space-filling-curve ``part`` (partition.Partition; ``grid`` must be stored in
    ``part.order``), this rank and the 128-byte ncclUniqueId of rank 0.  ``phis``, ``gradphis``,
    ``ref_states`` and every State are then the LOCAL slices [e0:e1) of this rank."""

    def __init__(self, grid, hv, params, phis, gradphis, ref_states, device=0, rank=0, part=None,
                 nccl_id=None, host_only=False):
        self.L = _libmod.load()
        L = self.L
        self.grid, self.hv, self.params = grid, hv, params
        self.rank, self.part = rank, part
        nranks = 1 if part is None else part.npart
        self.e0, self.e1 = (0, grid.nelem) if part is None else part.local_range(rank)
        self.nelem, self.nlev = self.e1 - self.e0, hv.nlev
        L.finalize_hommexx_session()
        idbuf = None if nccl_id is None else ctypes.create_string_buffer(bytes(nccl_id), 128)
        L.thetal_b200_init_comm(_ci(rank), _ci(nranks), _ci(device), idbuf)
        if not host_only:
            L.initialize_hommexx_session()
        self._check()
        L.thetal_b200_set_nlev(_ci(hv.nlev))
        p = params
        name = c_char_p(b"jw_baroclinic")
        L.init_simulation_params_c(
            _ci(p.vert_remap_u_alg), _ci(9), _ci(p.rsplit), _ci(p.qsplit), _ci(p.tstep_type), _ci(0), _ci(0),
            _cd(p.nu), _cd(p.nu_p), _cd(0.0), _cd(p.nu_s), _cd(p.nu_div), _cd(p.nu_top),
            _ci(p.hypervis_order), _ci(p.hypervis_subcycle), _ci(p.hypervis_subcycle_tom),
            _cd(p.hypervis_scaling), _cd(0.0), _cd(0.0),
            _ci(getattr(p, 'ftype', -1)), _ci(p.theta_advect_form), _ci(0), _ci(0), _ci(1), _ci(0),
            _ci(0), _ci(0), byref(name),
            _ci(p.rsplit * p.qsplit), _ci(p.qsplit), _cd(grid.rearth), _cd(1.0 / grid.rearth), _ci(1),
            _ci(p.pgrad_correction), _cd(p.dp3d_thresh), _cd(p.vtheta_thresh), _ci(0), _ci(0),
            _cd(p.tom_sponge_start))
        L.init_hvcoord_c(_cd(hv.ps0), _pp(hv.hyam), _pp(hv.hyai), _pp(hv.hybm), _pp(hv.hybi))
        L.init_elements_c(_ci(self.nelem))
        L.init_time_level_c(_ci(1), _ci(2), _ci(3), _ci(0), _ci(0))
        # connectivity: local ids are offsets into the owner's slice of the curve, pids are ranks+1,
        # max_corner_elem = 1
        L.init_connectivity(_ci(self.nelem), _ci(1))
        add = L.add_connection
        owner = np.zeros(grid.nelem, dtype=np.int64) if part is None else part.owner
        start = np.zeros(2, dtype=np.int64) if part is None else part.start
        for (e1, d1, e2, d2) in connections(grid, self.e0, self.e1):
            p2 = int(owner[e2])
            add(_ci(e1 - self.e0 + 1), _ci(int(grid.gnum[e1]) + 1), _ci(d1 + 1), _ci(rank + 1),
                _ci(e2 - int(start[p2]) + 1), _ci(int(grid.gnum[e2]) + 1), _ci(d2 + 1), _ci(p2 + 1))
        L.finalize_connectivity()
        if host_only:
            return
        dvv = np.ascontiguousarray(grid.dvv)
        mp = np.ascontiguousarray(grid.mp)
        L.init_reference_element_c(_pp(dvv), _pp(mp))
        cart = np.zeros(48)
        ll = np.zeros(32)
        gradphis = np.ascontiguousarray(gradphis)
        phis = np.ascontiguousarray(phis)
        for e in range(self.nelem):
            g = e + self.e0
            L.init_elements_2d_c(_ci(e), _pp(grid.D[g]), _pp(grid.Dinv[g]), _pp(grid.fcor[g]),
                                 _pp(grid.spheremp[g]), _pp(grid.rspheremp[g]), _pp(grid.metdet[g]),
                                 _pp(grid.metinv[g]), _pp(grid.tensorvisc[g]), _pp(grid.vec_sph2cart[g]),
                                 cart.ctypes.data_as(POINTER(c_double)), ll.ctypes.data_as(POINTER(c_double)),
                                 _pp(grid.tensorvisc[g]))
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
        if self.L.thetal_b200_last_error() != 0:      # latched (abort_on_error = 0): report the first failure, then clear
            msg = self.L.thetal_b200_last_error_msg().decode()
            self.L.thetal_b200_clear_error()
            raise RuntimeError("libthetal_b200: " + msg)

    def set_state(self, st, ps_v=None):
        """init_elements_states_c: host (flat Fortran layout) -> device; ps_v [ie][tl][16] optional."""
        self.L.init_elements_states_c(_pp(st.v), _pp(st.w_i), _pp(st.vtheta_dp), _pp(st.phinh_i),
                                      _pp(st.dp3d), None, None if ps_v is None else _pp(ps_v))
        self._check()

    def set_state_strided(self, v, w_i, vtheta_dp, phinh_i, dp3d, stride_bytes):
        """init_elements_states_strided_c: the fields of elem(1) inside an array of element structs."""
        pp = lambda a: byref(c_void_p(a))
        self.L.init_elements_states_strided_c(pp(v), pp(w_i), pp(vtheta_dp), pp(phinh_i), pp(dp3d),
                                              byref(ctypes.c_long(int(stride_bytes))))
        self._check()

    def initialize_dp3d_from_ps(self):
        self.L.initialize_dp3d_from_ps_c()
        self._check()

    def get_state(self, st=None, ps_v=None):
        """cxx_push_results_to_f90: device -> host (ps_v [ie][tl][16] is filled when given)."""
        if st is None:
            st = State(self.nelem, self.nlev)
        self.L.cxx_push_results_to_f90(_pp(st.v), _pp(st.w_i), _pp(st.vtheta_dp), _pp(st.phinh_i),
                                       _pp(st.dp3d), None if ps_v is None else _pp(ps_v), None, None, None)
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

    def prim_run_subcycle(self, dt):
        """prim_run_subcycle_c called the way prim_driver_mod.F90:554-560 calls it: the four integers are
        uninitialised on entry (pure outputs).  Returns (nstep, nm1, n0, np1) as the library's TimeLevel
        holds them, 0-BASED (the Fortran caller adds 1)."""
        a, b, c, d = c_int(-12345), c_int(-12345), c_int(-12345), c_int(-12345)
        self.L.prim_run_subcycle_c(_cd(dt), byref(a), byref(b), byref(c), byref(d), _ci(0), _ci(0))
        self._check()
        return a.value, b.value, c.value, d.value

    def push_forcing(self, fm, fvtheta, fphi):
        """push_forcing_to_c: derived%FM [ie][k][3][16], FVTheta [ie][k][16], FPHI [ie][k<=nlev][16] (FT, FQ unused)."""
        for a in (fm, fvtheta, fphi):
            assert a.flags["C_CONTIGUOUS"] and a.dtype == np.float64
        self.L.push_forcing_to_c(c_void_p(fm.ctypes.data), c_void_p(fvtheta.ctypes.data), None,
                                 c_void_p(fphi.ctypes.data), None)
        self._check()

    def apply_forcing_dynamics(self, nt, dt):
        """applyCAMforcing_dynamics (prim_advance_mod.F90:517-541) on time level nt (0-based)."""
        self.L.apply_cam_forcing_dynamics_c(_ci(nt), _cd(dt))
        self._check()

    def dss(self, field):
        """dss_c: node sums of field[nelem, nlayers, 16] in place (edgeVpack / bndry_exchangeV / edgeVunpack)."""
        assert field.flags["C_CONTIGUOUS"] and field.dtype == np.float64 and field.shape[0] == self.nelem and field.shape[-1] == 16
        self.L.dss_c(_pp(field), _ci(field.size // (self.nelem * 16)))
        self._check()
        return field

    def init_time_level(self, nm1, n0, np1, nstep=0, nstep0=0):
        """init_time_level_c with 1-BASED levels like the Fortran caller (prim_driver_mod.F90:196)."""
        self.L.init_time_level_c(_ci(nm1), _ci(n0), _ci(np1), _ci(nstep), _ci(nstep0))

    def vertical_remap(self, np1):
        """vertremap_mod.F90:23-171 on time level np1 (0-based)."""
        self.L.vertical_remap_c(_ci(np1))
        self._check()

    def set_vertical_remap(self, on):
        """whether prim_run_subcycle remaps after its rsplit steps (default on, as the reference)."""
        self.L.thetal_b200_set_vertical_remap(_ci(1 if on else 0))

    def time_steps(self, dt, nsteps, nm1, n0, np1):
        """Run nsteps dynamics steps back to back; returns (device ms, nm1, n0, np1)."""
        b, c, d = c_int(nm1), c_int(n0), c_int(np1)
        ms = self.L.thetal_b200_time_steps(_cd(dt), _ci(nsteps), byref(b), byref(c), byref(d))
        self._check()
        return ms, b.value, c.value, d.value

    def launch_count(self):
        return int(self.L.thetal_b200_launch_count())

    def halo_tables(self):
        """Partition-boundary tables of this rank (host side): dict with nghost, peers
        [(rank, send_off, send_cnt, recv_off, recv_cnt)], send_idx, recv_idx, dss_src[nelem,16,3]."""
        L = self.L
        ng, npeer, ns, nr = c_int(0), c_int(0), c_int(0), c_int(0)
        L.thetal_b200_halo_counts(byref(ng), byref(npeer), byref(ns), byref(nr))
        arr = [np.zeros(max(npeer.value, 1), dtype=np.int32) for _ in range(5)]
        sidx = np.zeros(max(ns.value, 1), dtype=np.int32)
        ridx = np.zeros(max(nr.value, 1), dtype=np.int32)
        L.thetal_b200_halo_lists(*[c_void_p(a.ctypes.data) for a in arr], c_void_p(sidx.ctypes.data),
                                 c_void_p(ridx.ctypes.data))
        dss = np.zeros((self.nelem, 16, 3), dtype=np.int32)
        L.thetal_b200_dss_table(c_void_p(dss.ctypes.data))
        peers = [tuple(int(a[i]) for a in arr) for i in range(npeer.value)]
        return {"nghost": ng.value, "peers": peers, "send_idx": sidx[:ns.value], "recv_idx": ridx[:nr.value],
                "dss_src": dss}

    def global_diagnostics(self, tl):
        """prim_energy_halftimes + global_integral of time level tl on the device (prim_state_mod.F90:822-939,
        global_norms_mod.F90:47-90): dict ps_min, ps_max, ps_mean, KE, PE, IE, energy (integral / 4 pi).
        Order-independent sum: the same bits on every rank of every partition."""
        out = np.zeros(8)
        self.L.thetal_b200_global_diagnostics(_ci(tl), c_void_p(out.ctypes.data))
        self._check()
        keys = ("ps_min", "ps_max", "ps_mean", "KE", "PE", "IE", "energy")
        return {k: float(out[i]) for i, k in enumerate(keys)}

    def field_extrema(self, tl):
        """prim_printstate's extrema (prim_state_mod.F90:219-330): {name: (min, max)} for u, v, w, theta, T, dz, dp3d."""
        out = np.zeros(14)
        self.L.thetal_b200_field_extrema(_ci(tl), c_void_p(out.ctypes.data))
        self._check()
        return {k: (float(out[2 * i]), float(out[2 * i + 1])) for i, k in enumerate(("u", "v", "w", "theta", "T", "dz", "dp3d"))}

    def halo_stats(self):
        ex, by = ctypes.c_long(0), ctypes.c_long(0)
        self.L.thetal_b200_halo_stats(byref(ex), byref(by))
        return ex.value, by.value

    def close(self):
        self.L.finalize_hommexx_session()


def build_case(params, hv_ref_profiles=None, part=None, rank=0):
    """Geometry + JW initial state + reference states for a SimParams config (host side).

    With ``part`` the grid is stored along the space-filling curve and the state, phis, gradphis and
    reference states returned are this rank's slice only (the geometry stays global: it is small and
    the DSS of spheremp/gradphis needs the neighbours)."""
    from .geometry import CubedSphere
    from .hvcoord import HVCoord
    from .initial import jw_baroclinic, jw_phis, model_init2, make_c0, gradient_sphere
    hv = HVCoord.from_table(params.nlev)
    hvr = params.hv_ref_profiles if hv_ref_profiles is None else hv_ref_profiles
    if part is None:
        grid = CubedSphere(params.ne, hypervis_scaling=params.hypervis_scaling)
        st, phis = jw_baroclinic(grid, hv, params.u_perturb)
        gradphis, theta_ref, dp_ref, phi_ref = model_init2(grid, hv, st, phis, hvr)
        return grid, hv, st, phis, gradphis, (theta_ref, dp_ref, phi_ref)
    grid = CubedSphere(params.ne, order=part.order, hypervis_scaling=params.hypervis_scaling)
    e0, e1 = part.local_range(rank)
    gradphis_all = make_c0(grid, gradient_sphere(grid, jw_phis(grid)))
    loc = grid.local_view(e0, e1)
    st, phis = jw_baroclinic(loc, hv, params.u_perturb)
    gradphis, theta_ref, dp_ref, phi_ref = model_init2(loc, hv, st, phis, hvr,
                                                       gradphis=np.ascontiguousarray(gradphis_all[e0:e1]))
    return grid, hv, st, phis, gradphis, (theta_ref, dp_ref, phi_ref)


def broadcast_nccl_id(td, rank):
    """ncclUniqueId of rank 0 to every rank over the torch.distributed process group ``td``."""
    import torch
    buf = ctypes.create_string_buffer(128)
    if rank == 0:
        _libmod.load().thetal_b200_get_nccl_id(buf)
    dev = "cuda" if td.get_backend() == "nccl" else "cpu"
    t = torch.tensor(list(buf.raw), dtype=torch.uint8, device=dev)
    td.broadcast(t, src=0)
    return bytes(t.cpu().tolist())
