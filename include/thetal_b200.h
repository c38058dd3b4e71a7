/* C ABI of libthetal_b200.so -- B200-native theta-l nonhydrostatic dycore step.
 *
 * The first group re-exports, symbol for symbol, the C ABI that HOMME's Fortran driver for the
 * theta-l_kokkos target already binds (bind(c) interfaces), so that driver links against this
 * library unchanged.  Fortran bind(c) passes scalars and type(c_ptr) BY REFERENCE, hence every
 * scalar is a pointer and every array argument is a pointer to a pointer.  Arrays are Fortran
 * ordered (column major), element/time-level indices follow the reference (see each entry).
 * H/ = components/homme/ in the reference tree.
 *
 * The second group (thetal_b200_* and *_c additions) is the finer-grained surface used by the
 * parity tests and by drivers that keep prim_run_subcycle in Fortran.
 *
 * All functions are void like the reference's; fatal conditions print to stderr and abort()
 * unless thetal_b200_set_abort_on_error(0) was called, in which case the error is latched and
 * readable with thetal_b200_last_error().
 */
#ifndef THETAL_B200_H
#define THETAL_B200_H
#ifdef __cplusplus
extern "C" {
#endif

typedef const double *const *thetal_f90ptr;   /* Fortran "type(c_ptr), intent(in)" (by reference) */
typedef double *const *thetal_f90ptr_out;

/* ---- session / communicator ---------------------------------------------------------------- */
/* H/src/share/cxx/prim_cxx_driver_base.F90:46-57,125-128; impl Hommexx_Session.cpp:73,97,
 * mpi/mpi_cxx_f90_interface.cpp:20-28 */
void initialize_hommexx_session(void);
void finalize_hommexx_session(void);
void reset_cxx_comm(const int *f_comm);       /* MPI_Fint handle; ranks come from thetal_b200_init_comm */

/* ---- connectivity: H/src/share/cxx/prim_cxx_driver_base.F90:147-196; impl
 * mpi/mpi_cxx_f90_interface.cpp:30-85.  All indices 1-based; pos = (dir+1)*max_corner_elems+idx,
 * dir: W=0,E=1,S=2,N=3,SW=4,SE=5,NW=6,NE=7 (control_mod.F90 west=1..neast=8). */
void init_connectivity(const int *num_local_elems, const int *max_corner_elems);
void add_connection(const int *e1_lid, const int *e1_gid, const int *e1_pos, const int *e1_pid,
                    const int *e2_lid, const int *e2_gid, const int *e2_pos, const int *e2_pid);
void finalize_connectivity(void);

/* ---- initialisation: H/src/theta-l_kokkos/theta_f2c_mod.F90:10-186; impl
 * H/src/theta-l_kokkos/cxx/cxx_f90_interface_theta.cpp:41-686 */
void init_simulation_params_c(const int *remap_alg, const int *limiter_option, const int *rsplit,
    const int *qsplit, const int *time_step_type, const int *qsize, const int *state_frequency,
    const double *nu, const double *nu_p, const double *nu_q, const double *nu_s,
    const double *nu_div, const double *nu_top, const int *hypervis_order,
    const int *hypervis_subcycle, const int *hypervis_subcycle_tom,
    const double *hypervis_scaling, const double *laplace_scaling, const double *dcmip16_mu,
    const int *ftype, const int *theta_adv_form, const int *prescribed_wind,
    const int *use_moisture, const int *disable_diagnostics, const int *use_cpstar,
    const int *transport_alg, const int *theta_hydrostatic_mode, const char **test_case,
    const int *dt_remap_factor, const int *dt_tracer_factor, const double *scale_factor,
    const double *laplacian_rigid_factor, const int *nsplit, const int *pgrad_correction,
    const double *dp3d_thresh, const double *vtheta_thresh,
    const int *internal_diagnostics_level, const int *do_3d_turbulence,
    const double *tom_sponge_start);
void init_elements_c(const int *nelemd);
void init_hvcoord_c(const double *ps0, thetal_f90ptr hyam, thetal_f90ptr hyai,
                    thetal_f90ptr hybm, thetal_f90ptr hybi);
void init_time_level_c(const int *nm1, const int *n0, const int *np1, const int *nstep,
                       const int *nstep0);            /* 1-based time levels */
void init_reference_element_c(thetal_f90ptr deriv, thetal_f90ptr mass);  /* Dvv(np,np), mp(np,np) */
/* ie is 0-based here (the Fortran passes ie-1, H/src/theta-l_kokkos/prim_driver_mod.F90:242) */
void init_elements_2d_c(const int *ie, thetal_f90ptr D, thetal_f90ptr Dinv, thetal_f90ptr fcor,
    thetal_f90ptr spheremp, thetal_f90ptr rspheremp, thetal_f90ptr metdet, thetal_f90ptr metinv,
    thetal_f90ptr tensorvisc, thetal_f90ptr vec_sph2cart, const double *sphere_cart_vec,
    const double *sphere_latlon_vec, thetal_f90ptr tensorvisc2);
void init_tensorvisc_c(const int *ie, thetal_f90ptr tensorvisc);
void init_tensorvisc2_c(const int *ie, thetal_f90ptr tensorvisc2);
void init_geopotential_c(const int *ie, thetal_f90ptr phis, thetal_f90ptr gradphis);
void init_diagnostics_c(thetal_f90ptr q, thetal_f90ptr qvar, thetal_f90ptr qmass,
    thetal_f90ptr q1mass, thetal_f90ptr iener, thetal_f90ptr kener, thetal_f90ptr pener);
/* flat arrays (np,np,[2],nlev|nlevp,timelevels,nelemd), H/src/theta-l_kokkos/element_state.F90:35-67 */
void init_elements_states_c(thetal_f90ptr v, thetal_f90ptr w_i, thetal_f90ptr vtheta_dp,
    thetal_f90ptr phinh_i, thetal_f90ptr dp3d, thetal_f90ptr qdp, thetal_f90ptr ps_v);
void init_reference_states_c(thetal_f90ptr theta_ref, thetal_f90ptr dp_ref, thetal_f90ptr phi_ref,
    thetal_f90ptr nu_scale_top);
void init_functors_c(const int *allocate_buffer);
void init_boundary_exchanges_c(void);
void initialize_dp3d_from_ps_c(void);   /* dp3d(n0) = dA ps0 + dB ps_v(n0), prim_driver.cpp:31-57 */
/* addition: state embedded in the plain theta-l array of element structs (element_mod.F90:32-144):
 * base pointers of elem(1)%state%..., stride = storage size of element_t in bytes */
void init_elements_states_strided_c(thetal_f90ptr v, thetal_f90ptr w_i, thetal_f90ptr vtheta_dp,
    thetal_f90ptr phinh_i, thetal_f90ptr dp3d, const long *elem_stride_bytes);

/* ---- run time: theta_f2c_mod.F90:189-236; impl H/src/share/cxx/prim_driver.cpp:60-181 */
/* derived%FM(np,np,3,nlev,nelemd), FVTheta(np,np,nlev,nelemd), FT (unused), FPHI(np,np,nlevp,nelemd), FQ (qsize = 0:
 * unused).  prim_run_subcycle_c applies them to time level n0 with dt_remap when ftype is 0 or 2
 * (prim_driver.cpp:117-125 -> applyCAMforcing_dynamics, prim_advance_mod.F90:517-541, standalone variant). */
void push_forcing_to_c(const double *fm, const double *fvtheta, const double *ft,
                       const double *fphi, const double *fq);
/* prim_printstate's field extrema of time level *tl (0-based), prim_state_mod.F90:219-330: out[2f], out[2f+1] = global
 * min, max of f = u, v, w_i, theta, T, dz, dp3d (14 doubles); identical on every rank */
void thetal_b200_field_extrema(const int *tl, double *out);
/* addition: DSS of a caller's field(np,np,nlayers,nelemd) in place (node sums in the Fortran unpack order,
 * edge_mod.F90:500-629,862-944; halos over NCCL when partitioned); nlayers <= 6*nlev+1 */
void dss_c(thetal_f90ptr_out field, const int *nlayers);
/* addition: applyCAMforcing_dynamics alone on time level *nt (0-based), for parity tests */
void apply_cam_forcing_dynamics_c(const int *nt, const double *dt);
/* nstep, nm1, n0, np1 are pure OUTPUTS (never read): the library advances its own TimeLevel, set by
 * init_time_level_c, and reports it 0-BASED as prim_driver.cpp:174-178 does; the Fortran caller adds 1
 * (prim_driver_mod.F90:554-560).  One call = rsplit prim_steps (each: zero vn0/omega_p/eta_dot_dpdn, then
 * dt_tracer_factor dynamics steps), vertical_remap, last leapfrog update. */
void prim_run_subcycle_c(const double *tstep, int *nstep, int *nm1, int *n0, int *np1,
                         const int *next_output_step, const int *nsplit_iter);
void cxx_push_results_to_f90(thetal_f90ptr_out v, thetal_f90ptr_out w_i, thetal_f90ptr_out vtheta_dp,
    thetal_f90ptr_out phinh_i, thetal_f90ptr_out dp3d, thetal_f90ptr_out ps_v, thetal_f90ptr_out qdp,
    thetal_f90ptr_out q, thetal_f90ptr_out omega_p);
void push_test_state_to_c(thetal_f90ptr ps_v, thetal_f90ptr dp3d, thetal_f90ptr vtheta_dp,
    thetal_f90ptr phinh_i, thetal_f90ptr v, thetal_f90ptr w_i, thetal_f90ptr eta_dot_dpdn,
    thetal_f90ptr vn0);
void sync_diagnostics_to_host_c(void);

/* ================= additions (not in the reference ABI) ===================================== */
/* nlev is a compile-time constant in the reference (dimensions_mod.F90:30-36); here it is set
 * once before init_elements_c.  np is fixed at 4. */
void thetal_b200_set_nlev(const int *nlev);
/* multi-GPU: one rank per GPU.  nccl_id = 128-byte ncclUniqueId from rank 0 (NULL when nranks==1). */
void thetal_b200_init_comm(const int *rank, const int *nranks, const int *device, const void *nccl_id);
void thetal_b200_get_nccl_id(void *nccl_id_out);
void thetal_b200_set_abort_on_error(const int *flag);
int  thetal_b200_last_error(void);             /* 0 = ok; message via thetal_b200_last_error_msg */
const char *thetal_b200_last_error_msg(void);
void thetal_b200_clear_error(void);            /* reset the latched error (abort_on_error == 0 mode) */
/* extra controls the reference keeps in Fortran modules: control_mod.F90:101-103, model_init_mod */
void thetal_b200_set_extra_params(const int *hv_ref_profiles_unused, const int *nlev_tom);

/* the step and its pieces; time levels 0-based, argument order of the Fortran routines:
 * prim_advance_exp  H/src/theta-l/share/prim_advance_mod.F90:81
 * compute_andor_apply_rhs(np1,nm1,n0,dt2,...,eta_ave_w,scale1,scale2,scale3)  :1023-1024
 * compute_stage_value_dirk(nm1,alphadt_nm1,n0,alphadt_n0,np1,dt2,...,itercount,itererr) imex_mod.F90:60-61
 * advance_hypervis(...,nt,...,dt2,eta_ave_w)  prim_advance_mod.F90:546 */
void prim_advance_exp_c(const double *dt, const int *nm1, const int *n0, const int *np1,
                        const int *nstep, const int *compute_diagnostics);
void caar_c(const int *np1, const int *nm1, const int *n0, const double *dt2,
            const double *eta_ave_w, const double *scale1, const double *scale2,
            const double *scale3);
void dirk_c(const int *nm1, const double *alphadt_nm1, const int *n0, const double *alphadt_n0,
            const int *np1, const double *dt2, int *itercount, double *itererr);
void advance_hypervis_c(const int *nt, const double *dt2, const double *eta_ave_w);
/* host<->device transfers of single pieces (tests, bench e2e) */
void thetal_b200_get_derived(double *vn0, double *eta_dot_dpdn, double *omega_p);
void thetal_b200_zero_derived(void);
void thetal_b200_sync(void);
/* run `nsteps` dynamics steps back to back (time-level rotation of prim_run_subcycle without
 * vertical remap); returns elapsed device milliseconds measured with CUDA events */
double thetal_b200_time_steps(const double *dt, const int *nsteps, int *nm1, int *n0, int *np1);
/* average per-kernel-family device time (ms) of the last thetal_b200_time_steps call:
 * out[0]=caar_pre out[1]=dss_post out[2]=dirk out[3]=hypervis(all) out[4]=combine; returns launches */
long thetal_b200_launch_count(void);
/* partition boundary built by finalize_connectivity (edge_mod.F90 putmapP/getmapP + schedule_mod.F90:37-280
   equivalents): ghost element slots, per-peer send/recv entry lists (elem*16+point), the DSS gather table
   [nelem][16][3], and counters of the exchanges done so far.  Host only; no device needed. */
void thetal_b200_halo_counts(int *nghost, int *npeers, int *nsend, int *nrecv);
void thetal_b200_halo_lists(int *peer_rank, int *send_off, int *send_cnt, int *recv_off, int *recv_cnt,
                            int *send_idx, int *recv_idx);
void thetal_b200_dss_table(int *dss_src);
void thetal_b200_halo_stats(long *exchanges, long *bytes);

/* Vertical remap of time level *np1 (0-based) from the Lagrangian levels back to the reference levels
 * (SURVEY 8f N1).  Replaces vertical_remap, H/src/theta-l/share/vertremap_mod.F90:23-171 (rsplit > 0,
 * qsize = 0) with remap_Q_ppm, H/src/share/vertremap_base.F90:514-646; the algorithm is the remap_alg
 * given to init_simulation_params_c (1, 10 or 11 -- the PPM family HOMMEXX supports).  A negative
 * layer thickness is fatal (code 11) as in vertremap_mod.F90:87-103.  prim_run_subcycle_c calls it
 * after its rsplit dynamics steps unless thetal_b200_set_vertical_remap(0) was called. */
void vertical_remap_c(const int *np1);
void thetal_b200_set_vertical_remap(const int *on);

/* Global diagnostics of one time level, computed on the device.  Replaces the Fortran
 * side's prim_energy_halftimes + global_integral + wrap_repro_sum for the quantities prim_printstate
 * reports (H/src/theta-l/share/prim_state_mod.F90:822-939, :387-429; H/src/share/global_norms_mod.F90:47-90).
 * tl is 0-based.  out[8] = min ps, max ps, <ps>, <KE>, <PE>, <IE>, <KE+PE+IE> (global_integral's
 * convention: integral / 4 pi), elements owned by this rank.  The sum over elements is an
 * order-independent fixed-point sum: identical bits on every rank and for every partition. */
void thetal_b200_global_diagnostics(const int *tl, double *out);
/* The same diagnostics for callers that may not allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and the same bits in out[8] as thetal_b200_global_diagnostics; the
 * scratch (element integrals and accumulators) is a static device array inside the library, 512 KB held by
 * every GPU that loads it until exit.  Limits: at most THETAL_B200_NOALLOC_MAX_ELEMS elements on the rank and a
 * single-rank session; beyond either the call records an error (thetal_b200_last_error) and leaves out
 * untouched.  It calls no cudaMalloc/cudaFree and no allocating library routine.  All calls share the array:
 * calls that overlap on the GPU (several host threads or streams) race. */
#define THETAL_B200_NOALLOC_MAX_ELEMS 16384
void thetal_b200_global_diagnostics_noalloc(const int *tl, double *out);
void thetal_b200_profile_step(const double *dt, const int *nm1, const int *n0, const int *np1,
                              double *ms_by_family, int nfamily);

/* single-time-level transfers: host arrays [ie][k][(c)][16] of ONE time level (0-based tl),
 * asynchronous H2D on the library stream; pull synchronises.  Used by the bench's e2e leg and by
 * drivers that only need the new state (the reference pulls all three levels). */
void thetal_b200_push_level(const int *tl, const double *v, const double *w_i, const double *vtheta_dp,
                            const double *phinh_i, const double *dp3d);
void thetal_b200_pull_level(const int *tl, double *v, double *w_i, double *vtheta_dp,
                            double *phinh_i, double *dp3d);

/* Pipelined single-step service for independent inputs (bench e2e, ensemble-style drivers): one
 * prim_advance_exp per call, n0 state read from pinned host arrays of ONE time level
 * ([ie][k][(c)][16]), np1 state written to pinned host arrays.  Consecutive calls overlap their
 * host->device copy, kernels and device->host copy on three streams; the outputs are complete after
 * thetal_b200_pipeline_flush().  The device-resident state is scratch while this is in use. */
void thetal_b200_pipeline_step(const double *dt, const double *v_in, const double *w_in,
    const double *vtheta_dp_in, const double *phinh_in, const double *dp3d_in,
    double *v_out, double *w_out, double *vtheta_dp_out, double *phinh_out, double *dp3d_out);
void thetal_b200_pipeline_flush(void);
void thetal_b200_pin_host(void *ptr, const long *bytes, const int *pin);   /* cudaHostRegister */
void thetal_b200_enqueue_step(const double *dt, const int *nm1, const int *n0, const int *np1);
/* measured FP64 FMA rate of this GPU, TFLOP/s (independent DFMA chains on every SM, CUDA events) */
double thetal_b200_measure_fp64_peak(void);
/* device ms per launch of compute_stage_value_dirk alone, the same stage problem `reps` times
 * (w_i, phinh_i of np1 restored before every launch); kernel tuning aid (tools/bench_dirk.py) */
double thetal_b200_time_dirk(const int *nm1, const double *alphadt_nm1, const int *n0, const double *alphadt_n0,
                             const int *np1, const double *dt2, const int *reps, int *itercount);

#ifdef __cplusplus
}
#endif
#endif
