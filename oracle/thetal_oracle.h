/* TEST INFRASTRUCTURE ONLY -- CPU restatement (oracle) of HOMME's theta-l NH dycore step.
 *
 * Parity status: UNPINNED against the Fortran executable (no Fortran compiler, no MPI and no
 * golden vectors for this path exist in the reference tree).  The
 * oracle follows the Fortran source line by line (non-HOMMEXX_BFB_TESTING branches) and is
 * pinned only by the reference's own invariants (tests/test_oracle_invariants.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (e3sm_b200/csrc) never links or calls it.
 *
 * Layout: the flat Fortran arrays of H/src/theta-l_kokkos/element_state.F90:35-67 (column-major
 * (np,np,[2],nlev|nlevp,timelevels,nelemd)), i.e. C order [ie][tl][k][c][j][i].
 * H/ = components/homme/ in the reference tree.
 */
#ifndef THETAL_ORACLE_H
#define THETAL_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_NP 4
#define ORC_NPSQ 16
#define ORC_TL 3

typedef struct {
  /* dimensions */
  int nelem, nlev;
  /* controls (H/src/share/control_mod.F90:57-209) */
  int rsplit, theta_advect_form, pgrad_correction, hypervis_subcycle, hypervis_subcycle_tom;
  int nlev_tom, hypervis_scaling_on, hv_theta_correction;
  double nu, nu_div, nu_p, nu_s, nu_top, dp3d_thresh, vtheta_thresh;
  double rearth;            /* scale_factor; scale_factor_inv = laplacian_rigid_factor = 1/rearth */
  /* hvcoord (H/src/share/hybvcoord_mod.F90:19-32) */
  double ps0;
  const double *hyai, *hybi, *hyam, *hybm, *dp0, *nu_scale_top; /* nlevp, nlevp, nlev, nlev, nlev, nlev */
  /* reference element */
  const double *dvv;        /* Dvv(np,np) Fortran order: dvv[l*4+i] = Dvv(i,l) */
  const double *mp;         /* mp(np,np) */
  /* per element geometry, [ie][..][j][i] */
  const double *D, *Dinv;   /* (np,np,2,2): [ie][b][a][p]  = X(i,j,a,b) */
  const double *metdet, *rmetdet, *spheremp, *rspheremp, *fcor; /* [ie][p] */
  const double *metinv;     /* (np,np,2,2) */
  const double *tensorvisc; /* (np,np,2,2) */
  const double *vec_sph2cart; /* (np,np,3,2) */
  const double *gradphis;   /* (np,np,2): [ie][c][p] */
  const double *phis;       /* [ie][p] */
  const double *theta_ref, *dp_ref, *phi_ref; /* [ie][k][p]; phi_ref has nlevp */
  /* DSS gather table: for element ie, point p: up to 3 neighbour sources (ie2*16+p2), -1 = none,
     listed in the Fortran unpack order S,N,W,E,SW,SE,NE,NW (H/src/share/edge_mod.F90:888-941) */
  const int *dss_src;       /* [ie][16][3] */
  /* state, flat with 3 time levels */
  double *v, *w_i, *vtheta_dp, *phinh_i, *dp3d;
  /* derived accumulators [ie][k][..] */
  double *vn0, *eta_dot_dpdn, *omega_p;   /* (np,np,2,nlev), (np,np,nlevp), (np,np,nlev) */
  /* scratch owned by caller: nelem*16*(6*nlev+2) doubles each */
  double *scratch1, *scratch2;
  /* DIRK statistics (H/src/theta-l/element_state.F90:20-22) */
  int max_itercnt; double max_deltaerr, max_reserr;
  int nthreads;
  int error_flag;           /* set when the EOS sees a bad state (eos.F90:119-145) */
} orc_ctx;

void orc_caar(orc_ctx *c, int np1, int nm1, int n0, double dt2, double eta_ave_w,
              double scale1, double scale2, double scale3);
void orc_caar_pre(orc_ctx *c, int np1, int nm1, int n0, double dt2, double eta_ave_w,
              double scale1, double scale2, double scale3);
void orc_dirk(orc_ctx *c, int nm1, double alphadt_nm1, int n0, double alphadt_n0, int np1,
              double dt2, int *itercount, double *itererr);
void orc_advance_hypervis(orc_ctx *c, int nt, double dt2, double eta_ave_w);
int orc_vertical_remap(orc_ctx *c, int np1, int remap_alg);   /* vertremap_mod.F90:23-171, PPM remap_alg 1/10/11 */
void orc_prim_advance_exp(orc_ctx *c, int tstep_type, double dt, int nm1, int n0, int np1,
              int qsplit);
void orc_limiter_dp3d_k(const orc_ctx *c, double *dp3d, double *vtheta_dp, const double *spheremp);
/* helpers exposed for invariant tests */
void orc_eos(const orc_ctx *c, const double *vtheta_dp, const double *dp3d, const double *phi_i,
             double *pnh, double *exner, double *dpnh_dp_i, int *err);
void orc_phi_from_eos(const orc_ctx *c, const double *phis, const double *vtheta_dp,
             const double *dp, double *phi_i);
void orc_dirk_jacobian(const orc_ctx *c, double *JacL, double *JacD, double *JacU, double dt2,
             const double *dp3d, const double *dphi, const double *pnh);
void orc_tridiag(int nlev, const double *dl, double *d, const double *du, double *x);
void orc_dss_field(const orc_ctx *c, double *f, int nlayers, int elem_stride, double *tmp);
void orc_gradient_sphere(const orc_ctx *c, int ie, const double *s, double *ds);
void orc_divergence_sphere(const orc_ctx *c, int ie, const double *v, double *div);
void orc_vorticity_sphere(const orc_ctx *c, int ie, const double *v, double *vort);
void orc_laplace_sphere_wk(const orc_ctx *c, int ie, const double *s, int var_coef, double *lap);
void orc_vlaplace_sphere_wk(const orc_ctx *c, int ie, const double *v, int var_coef,
             double nu_ratio, double *lap);

#ifdef __cplusplus
}
#endif
/* PA:517-541 */
void orc_apply_forcing_dynamics(orc_ctx *c, int np1, double dt, const double *fm, const double *fvtheta,
                                const double *fphi);
#endif
