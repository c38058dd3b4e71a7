/* TEST INFRASTRUCTURE ONLY -- see thetal_oracle.h.  CPU restatement of the Fortran theta-l step.
 * H/ = components/homme/ ; PA = H/src/theta-l/share/prim_advance_mod.F90 ;
 * DM = H/src/share/derivative_mod.F90 ; IM = H/src/theta-l/share/imex_mod.F90 ;
 * EOS = H/src/theta-l/share/eos.F90 ; VT = H/src/theta-l/share/viscosity_theta.F90.
 * Compile with -ffp-contract=off so no FMA contraction changes the Fortran operation order.
 */
#include "thetal_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NP 4
#define NPSQ 16
/* physical_constants.F90:55-73,86 */
static const double g_ = 9.80616, Rgas = 287.04, Cp = 1005.0, p0 = 100000.0, TREF = 288.0;
static const double tref_lapse_rate = 0.0065; /* element_ops.F90:80 */
#define KAPPA (Rgas / Cp)

/* X(i,j,a,b), a,b 0-based, p = i+4j */
#define G22(X, ie, a, b, p) ((X)[(size_t)(ie)*64 + ((b)*2 + (a))*16 + (p)])
#define DVV(i, l) (c->dvv[(l)*4 + (i)])

static inline size_t off_v(const orc_ctx *c, int ie, int tl) { return ((size_t)ie*ORC_TL + tl)*32*c->nlev; }
static inline size_t off_m(const orc_ctx *c, int ie, int tl) { return ((size_t)ie*ORC_TL + tl)*16*c->nlev; }
static inline size_t off_i(const orc_ctx *c, int ie, int tl) { return ((size_t)ie*ORC_TL + tl)*16*(c->nlev+1); }

/* ---------------- sphere operators ---------------- */
/* DM:410-450 */
void orc_gradient_sphere(const orc_ctx *c, int ie, const double *s, double *ds)
{
  double v1[NPSQ], v2[NPSQ];
  const double sfi = 1.0 / c->rearth;
  for (int j = 0; j < NP; j++)
    for (int l = 0; l < NP; l++) {
      double dsdx00 = 0.0, dsdy00 = 0.0;
      for (int i = 0; i < NP; i++) {
        dsdx00 = dsdx00 + DVV(i, l) * s[i + 4*j];
        dsdy00 = dsdy00 + DVV(i, l) * s[j + 4*i];
      }
      v1[l + 4*j] = dsdx00 * sfi;
      v2[j + 4*l] = dsdy00 * sfi;
    }
  for (int p = 0; p < NPSQ; p++) {
    ds[p]      = G22(c->Dinv, ie, 0, 0, p) * v1[p] + G22(c->Dinv, ie, 1, 0, p) * v2[p];
    ds[16 + p] = G22(c->Dinv, ie, 0, 1, p) * v1[p] + G22(c->Dinv, ie, 1, 1, p) * v2[p];
  }
}

/* DM:1021-1068 ; v is (np,np,2) */
void orc_divergence_sphere(const orc_ctx *c, int ie, const double *v, double *div)
{
  double gv1[NPSQ], gv2[NPSQ], vvtemp[NPSQ];
  const double sfi = 1.0 / c->rearth;
  const double *metdet = c->metdet + (size_t)ie*16, *rmetdet = c->rmetdet + (size_t)ie*16;
  for (int p = 0; p < NPSQ; p++) {
    gv1[p] = metdet[p] * (G22(c->Dinv, ie, 0, 0, p) * v[p] + G22(c->Dinv, ie, 0, 1, p) * v[16 + p]);
    gv2[p] = metdet[p] * (G22(c->Dinv, ie, 1, 0, p) * v[p] + G22(c->Dinv, ie, 1, 1, p) * v[16 + p]);
  }
  for (int j = 0; j < NP; j++)
    for (int l = 0; l < NP; l++) {
      double dudx00 = 0.0, dvdy00 = 0.0;
      for (int i = 0; i < NP; i++) {
        dudx00 = dudx00 + DVV(i, l) * gv1[i + 4*j];
        dvdy00 = dvdy00 + DVV(i, l) * gv2[j + 4*i];
      }
      div[l + 4*j] = dudx00;
      vvtemp[j + 4*l] = dvdy00;
    }
  for (int p = 0; p < NPSQ; p++) div[p] = (div[p] + vvtemp[p]) * (rmetdet[p] * sfi);
}

/* DM:915-966 */
void orc_vorticity_sphere(const orc_ctx *c, int ie, const double *v, double *vort)
{
  double vco1[NPSQ], vco2[NPSQ], vtemp[NPSQ];
  const double sfi = 1.0 / c->rearth;
  const double *rmetdet = c->rmetdet + (size_t)ie*16;
  for (int p = 0; p < NPSQ; p++) {
    vco1[p] = G22(c->D, ie, 0, 0, p) * v[p] + G22(c->D, ie, 1, 0, p) * v[16 + p];
    vco2[p] = G22(c->D, ie, 0, 1, p) * v[p] + G22(c->D, ie, 1, 1, p) * v[16 + p];
  }
  for (int j = 0; j < NP; j++)
    for (int l = 0; l < NP; l++) {
      double dudy00 = 0.0, dvdx00 = 0.0;
      for (int i = 0; i < NP; i++) {
        dvdx00 = dvdx00 + DVV(i, l) * vco2[i + 4*j];
        dudy00 = dudy00 + DVV(i, l) * vco1[j + 4*i];
      }
      vort[l + 4*j] = dvdx00;
      vtemp[j + 4*l] = dudy00;
    }
  for (int p = 0; p < NPSQ; p++) vort[p] = (vort[p] - vtemp[p]) * (rmetdet[p] * sfi);
}

/* DM:784-858 */
static void divergence_sphere_wk(const orc_ctx *c, int ie, const double *v, double *div)
{
  double vt1[NPSQ], vt2[NPSQ];
  const double sfi = 1.0 / c->rearth;
  const double *spheremp = c->spheremp + (size_t)ie*16;
  for (int p = 0; p < NPSQ; p++) {
    vt1[p] = G22(c->Dinv, ie, 0, 0, p) * v[p] + G22(c->Dinv, ie, 0, 1, p) * v[16 + p];
    vt2[p] = G22(c->Dinv, ie, 1, 0, p) * v[p] + G22(c->Dinv, ie, 1, 1, p) * v[16 + p];
  }
  for (int n = 0; n < NP; n++)
    for (int m = 0; m < NP; m++) {
      double d = 0.0;
      for (int j = 0; j < NP; j++)
        d = d - (spheremp[j + 4*n] * vt1[j + 4*n] * DVV(m, j)
               + spheremp[m + 4*j] * vt2[m + 4*j] * DVV(n, j)) * sfi;
      div[m + 4*n] = d;
    }
}

/* DM:1072-1119 ; var_coef uses elem%tensorVisc */
void orc_laplace_sphere_wk(const orc_ctx *c, int ie, const double *s, int var_coef, double *lap)
{
  double grads[32], old[32];
  orc_gradient_sphere(c, ie, s, grads);
  if (var_coef) {
    memcpy(old, grads, sizeof old);
    for (int p = 0; p < NPSQ; p++) {
      grads[p]      = old[p] * G22(c->tensorvisc, ie, 0, 0, p) + old[16 + p] * G22(c->tensorvisc, ie, 0, 1, p);
      grads[16 + p] = old[p] * G22(c->tensorvisc, ie, 1, 0, p) + old[16 + p] * G22(c->tensorvisc, ie, 1, 1, p);
    }
  }
  divergence_sphere_wk(c, ie, grads, lap);
}

/* DM:516-580 */
static void gradient_sphere_wk_testcov(const orc_ctx *c, int ie, const double *s, double *ds)
{
  double dc1[NPSQ], dc2[NPSQ];
  const double sfi = 1.0 / c->rearth;
  const double *mp = c->mp, *metdet = c->metdet + (size_t)ie*16;
  for (int n = 0; n < NP; n++)
    for (int m = 0; m < NP; m++) {
      int mn = m + 4*n;
      double a1 = 0.0, a2 = 0.0;
      for (int j = 0; j < NP; j++) {
        a1 = a1 - ((mp[j + 4*n] * G22(c->metinv, ie, 0, 0, mn) * metdet[mn] * s[j + 4*n] * DVV(m, j))
                 + (mp[m + 4*j] * G22(c->metinv, ie, 1, 0, mn) * metdet[mn] * s[m + 4*j] * DVV(n, j)));
        a2 = a2 - ((mp[j + 4*n] * G22(c->metinv, ie, 0, 1, mn) * metdet[mn] * s[j + 4*n] * DVV(m, j))
                 + (mp[m + 4*j] * G22(c->metinv, ie, 1, 1, mn) * metdet[mn] * s[m + 4*j] * DVV(n, j)));
      }
      dc1[mn] = a1; dc2[mn] = a2;
    }
  for (int p = 0; p < NPSQ; p++) {
    ds[p]      = (G22(c->D, ie, 0, 0, p) * dc1[p] + G22(c->D, ie, 0, 1, p) * dc2[p]) * sfi;
    ds[16 + p] = (G22(c->D, ie, 1, 0, p) * dc1[p] + G22(c->D, ie, 1, 1, p) * dc2[p]) * sfi;
  }
}

/* DM:453-513 */
static void curl_sphere_wk_testcov(const orc_ctx *c, int ie, const double *s, double *ds)
{
  double dc1[NPSQ], dc2[NPSQ];
  const double sfi = 1.0 / c->rearth;
  const double *mp = c->mp;
  for (int n = 0; n < NP; n++)
    for (int m = 0; m < NP; m++) {
      double a1 = 0.0, a2 = 0.0;
      for (int j = 0; j < NP; j++) {
        a1 = a1 - (mp[m + 4*j] * s[m + 4*j] * DVV(n, j));
        a2 = a2 + (mp[j + 4*n] * s[j + 4*n] * DVV(m, j));
      }
      dc1[m + 4*n] = a1; dc2[m + 4*n] = a2;
    }
  for (int p = 0; p < NPSQ; p++) {
    ds[p]      = (G22(c->D, ie, 0, 0, p) * dc1[p] + G22(c->D, ie, 0, 1, p) * dc2[p]) * sfi;
    ds[16 + p] = (G22(c->D, ie, 1, 0, p) * dc1[p] + G22(c->D, ie, 1, 1, p) * dc2[p]) * sfi;
  }
}

/* DM:1122-1251 (contra form, and cartesian form when var_coef) */
void orc_vlaplace_sphere_wk(const orc_ctx *c, int ie, const double *v, int var_coef,
                            double nu_ratio, double *lap)
{
  const double lrf = 1.0 / c->rearth; /* laplacian_rigid_factor, namelist_mod.F90:1047 */
  const double *spheremp = c->spheremp + (size_t)ie*16;
  if (var_coef) { /* DM:1164-1208 */
    const double *s2c = c->vec_sph2cart + (size_t)ie*96; /* (np,np,3,2): [comp2][comp3][p] */
    double dum[3][NPSQ], out[3][NPSQ];
    for (int comp = 0; comp < 3; comp++)
      for (int p = 0; p < NPSQ; p++)
        dum[comp][p] = s2c[(0*3 + comp)*16 + p] * v[p] + s2c[(1*3 + comp)*16 + p] * v[16 + p];
    for (int comp = 0; comp < 3; comp++) orc_laplace_sphere_wk(c, ie, dum[comp], 1, out[comp]);
    for (int comp = 0; comp < 2; comp++)
      for (int p = 0; p < NPSQ; p++)
        lap[comp*16 + p] = out[0][p] * s2c[(comp*3 + 0)*16 + p] + out[1][p] * s2c[(comp*3 + 1)*16 + p]
                         + out[2][p] * s2c[(comp*3 + 2)*16 + p];
  } else {
    double div[NPSQ], vor[NPSQ], gr[32], cu[32];
    orc_divergence_sphere(c, ie, v, div);
    orc_vorticity_sphere(c, ie, v, vor);
    for (int p = 0; p < NPSQ; p++) div[p] = nu_ratio * div[p];
    gradient_sphere_wk_testcov(c, ie, div, gr);
    curl_sphere_wk_testcov(c, ie, vor, cu);
    for (int p = 0; p < 32; p++) lap[p] = gr[p] - cu[p];
  }
  for (int p = 0; p < NPSQ; p++) {
    lap[p]      = lap[p]      + 2 * spheremp[p] * v[p]      * (lrf * lrf);
    lap[16 + p] = lap[16 + p] + 2 * spheremp[p] * v[16 + p] * (lrf * lrf);
  }
}

/* ---------------- equation of state ---------------- */
/* EOS:80-223 (NH branch) with dphi given; arrays are (np,np,nlev|nlevp) */
static void eos2(const orc_ctx *c, const double *vtheta_dp, const double *dp3d, const double *dphi,
                 double *pnh, double *exner, double *dpnh_dp_i, int *err)
{
  const int nlev = c->nlev, nlevp = nlev + 1;
  for (int q = 0; q < 16*nlev; q++)
    if (vtheta_dp[q] < 0 || dp3d[q] < 0 || dphi[q] > 0) { *err = 1; }
  for (int k = 0; k < nlev; k++)
    for (int p = 0; p < NPSQ; p++) {
      int q = k*16 + p;
      double p_over_exner = Rgas * vtheta_dp[q] / (-dphi[q]);
      pnh[q] = p0 * pow(p_over_exner / p0, 1 / (1 - KAPPA));
      exner[q] = pnh[q] / p_over_exner;
    }
  for (int p = 0; p < NPSQ; p++) {
    double pnh_i_top = c->hyai[0] * c->ps0;
    double pnh_i_bot = pnh[(nlev-1)*16 + p] + dp3d[(nlev-1)*16 + p] / 2;
    double dp3d_i_top = dp3d[p], dp3d_i_bot = dp3d[(nlev-1)*16 + p];
    dpnh_dp_i[p] = 2 * (pnh[p] - pnh_i_top) / dp3d_i_top;
    dpnh_dp_i[(nlevp-1)*16 + p] = 2 * (pnh_i_bot - pnh[(nlev-1)*16 + p]) / dp3d_i_bot;
    for (int k = 1; k < nlev; k++) {
      double dp3d_i = (dp3d[k*16 + p] + dp3d[(k-1)*16 + p]) / 2;
      dpnh_dp_i[k*16 + p] = (pnh[k*16 + p] - pnh[(k-1)*16 + p]) / dp3d_i;
    }
  }
}

/* EOS:35-76 */
void orc_eos(const orc_ctx *c, const double *vtheta_dp, const double *dp3d, const double *phi_i,
             double *pnh, double *exner, double *dpnh_dp_i, int *err)
{
  const int nlev = c->nlev;
  double *dphi = (double *)malloc(sizeof(double)*16*nlev);
  for (int k = 0; k < nlev; k++)
    for (int p = 0; p < NPSQ; p++) dphi[k*16 + p] = phi_i[(k+1)*16 + p] - phi_i[k*16 + p];
  eos2(c, vtheta_dp, dp3d, dphi, pnh, exner, dpnh_dp_i, err);
  free(dphi);
}

/* EOS:228-303 */
void orc_phi_from_eos(const orc_ctx *c, const double *phis, const double *vtheta_dp,
                      const double *dp, double *phi_i)
{
  const int nlev = c->nlev;
  double *p_i = (double *)malloc(sizeof(double)*16*(nlev+1));
  for (int p = 0; p < NPSQ; p++) {
    p_i[p] = c->hyai[0] * c->ps0;
    for (int k = 0; k < nlev; k++) p_i[(k+1)*16 + p] = p_i[k*16 + p] + dp[k*16 + p];
    phi_i[nlev*16 + p] = phis[p];
    for (int k = nlev - 1; k >= 0; k--) {
      double pm = (p_i[(k+1)*16 + p] + p_i[k*16 + p]) / 2;
      phi_i[k*16 + p] = phi_i[(k+1)*16 + p] + (Rgas * vtheta_dp[k*16 + p] * pow(pm / p0, KAPPA - 1)) / p0;
    }
  }
  free(p_i);
}

/* ---------------- DSS ---------------- */
/* edge_mod.F90:500-629,862-944 + bndry_mod.F90:51-169 as a node sum: every point receives its
 * own value plus the neighbours' values in the order S,N,W,E,SW,SE,NE,NW (dss_src lists them so).
 * f: per element nlayers*16 doubles at stride elem_stride; tmp: nelem*nlayers*16 scratch. */
void orc_dss_field(const orc_ctx *c, double *f, int nlayers, int elem_stride, double *tmp)
{
  const int nelem = c->nelem;
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < nelem; ie++)
    for (int l = 0; l < nlayers; l++)
      for (int p = 0; p < NPSQ; p++) {
        double s = f[(size_t)ie*elem_stride + l*16 + p];
        const int *src = c->dss_src + ((size_t)ie*16 + p)*3;
        for (int j = 0; j < 3; j++)
          if (src[j] >= 0) s = s + f[(size_t)(src[j] >> 4)*elem_stride + l*16 + (src[j] & 15)];
        tmp[((size_t)ie*nlayers + l)*16 + p] = s;
      }
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < nelem; ie++)
    memcpy(f + (size_t)ie*elem_stride, tmp + (size_t)ie*nlayers*16, sizeof(double)*nlayers*16);
}

/* ---------------- limiter ---------------- */
/* PA:1822-1931 ; dp3d, vtheta_dp (np,np,nlev) of one element */
void orc_limiter_dp3d_k(const orc_ctx *c, double *dp3d, double *vtheta_dp, const double *spheremp)
{
  const int nlev = c->nlev;
  const double *dp0 = c->dp0;
  int warn = 0;
  for (int k = 0; k < nlev; k++) {
    double mn = dp3d[k*16];
    for (int p = 1; p < NPSQ; p++) mn = fmin(mn, dp3d[k*16 + p]);
    if (mn < c->dp3d_thresh * dp0[k]) warn = 1;
  }
  if (warn) {
    double *Qcol = (double *)malloc(sizeof(double)*nlev);
    for (int q = 0; q < 16*nlev; q++) vtheta_dp[q] = vtheta_dp[q] / dp3d[q];
    for (int p = 0; p < NPSQ; p++) {
      double mn = dp3d[p] - c->dp3d_thresh * dp0[0];
      for (int k = 1; k < nlev; k++) mn = fmin(mn, dp3d[k*16 + p] - c->dp3d_thresh * dp0[k]);
      if (mn < 0) {
        double mass = 0, mass_new = 0;
        for (int k = 0; k < nlev; k++) {
          Qcol[k] = (dp3d[k*16 + p] - c->dp3d_thresh * dp0[k]) * spheremp[p];
          mass = mass + Qcol[k];
        }
        if (mass < 0) for (int k = 0; k < nlev; k++) Qcol[k] = -Qcol[k];
        for (int k = 0; k < nlev; k++) {
          if (Qcol[k] < 0) Qcol[k] = 0; else mass_new = mass_new + Qcol[k];
        }
        if (mass_new > 0) for (int k = 0; k < nlev; k++) Qcol[k] = Qcol[k] * (fabs(mass) / mass_new);
        if (mass < 0) for (int k = 0; k < nlev; k++) Qcol[k] = -Qcol[k];
        for (int k = 0; k < nlev; k++) dp3d[k*16 + p] = Qcol[k] / spheremp[p] + c->dp3d_thresh * dp0[k];
      }
    }
    for (int q = 0; q < 16*nlev; q++) vtheta_dp[q] = vtheta_dp[q] * dp3d[q];
    free(Qcol);
  }
  warn = 0;
  for (int q = 0; q < 16*nlev; q++)
    if (vtheta_dp[q] - c->vtheta_thresh * dp3d[q] < 0) warn = 1;
  if (warn)
    for (int q = 0; q < 16*nlev; q++)
      if ((vtheta_dp[q] - c->vtheta_thresh * dp3d[q]) < 0) vtheta_dp[q] = c->vtheta_thresh * dp3d[q];
}

/* ---------------- compute_andor_apply_rhs, element-local part (PA:1113-1702) ---------------- */
typedef struct {
  double *vtheta, *vtheta_i, *omega_i, *omega, *vort, *divdp, *phi, *pnh, *dp3d_i, *exner,
         *dpnh_dp_i, *eta_dot_dpdn, *KE, *gradexner, *gradpterm, *gradphinh_i, *mgrad, *gradKE,
         *wvor, *gradw_i, *v_gradw_i, *v_theta, *div_v_theta, *v_gradphinh_i, *v_i, *v_vadv,
         *theta_vadv, *w_vadv_i, *phi_vadv_i, *vtens1, *vtens2, *w_tens, *theta_tens, *phi_tens,
         *pi, *pi_i, *vgrad_p, *temp, *vtemp;
  double *base;
} caar_ws;

static void ws_alloc(caar_ws *w, int nlev)
{
  const size_t L = 16*(size_t)(nlev + 1);
  const int n1 = 28, n2 = 11; /* scalar-sized and vector-sized arrays */
  w->base = (double *)calloc((n1 + 2*n2)*L, sizeof(double));
  double *q = w->base;
#define A1(x) w->x = q; q += L
#define A2(x) w->x = q; q += 2*L
  A1(vtheta); A1(vtheta_i); A1(omega_i); A1(omega); A1(vort); A1(divdp); A1(phi); A1(pnh);
  A1(dp3d_i); A1(exner); A1(dpnh_dp_i); A1(eta_dot_dpdn); A1(KE); A1(v_gradw_i); A1(div_v_theta);
  A1(v_gradphinh_i); A1(theta_vadv); A1(w_vadv_i); A1(phi_vadv_i); A1(vtens1); A1(vtens2);
  A1(w_tens); A1(theta_tens); A1(phi_tens); A1(pi); A1(pi_i); A1(vgrad_p); A1(temp);
  A2(gradexner); A2(gradpterm); A2(gradphinh_i); A2(mgrad); A2(gradKE); A2(wvor); A2(gradw_i);
  A2(v_theta); A2(v_i); A2(v_vadv); A2(vtemp);
#undef A1
#undef A2
}
static void ws_free(caar_ws *w) { free(w->base); }

/* H/src/share/prim_si_mod.F90:94-148 */
static void preq_vertadv_v1(int nlev, const double *v, const double *eta_dot, const double *dp, double *v_vadv)
{
  for (int p = 0; p < NPSQ; p++) {
    int k = 0;
    double facp = 0.5 * eta_dot[(k+1)*16 + p] / dp[k*16 + p], facm;
    v_vadv[k*32 + p]      = facp * (v[(k+1)*32 + p] - v[k*32 + p]);
    v_vadv[k*32 + 16 + p] = facp * (v[(k+1)*32 + 16 + p] - v[k*32 + 16 + p]);
    for (k = 1; k < nlev - 1; k++) {
      facp = 0.5 * eta_dot[(k+1)*16 + p] / dp[k*16 + p];
      facm = 0.5 * eta_dot[k*16 + p] / dp[k*16 + p];
      v_vadv[k*32 + p]      = facp * (v[(k+1)*32 + p] - v[k*32 + p]) + facm * (v[k*32 + p] - v[(k-1)*32 + p]);
      v_vadv[k*32 + 16 + p] = facp * (v[(k+1)*32 + 16 + p] - v[k*32 + 16 + p]) + facm * (v[k*32 + 16 + p] - v[(k-1)*32 + 16 + p]);
    }
    k = nlev - 1;
    facm = 0.5 * eta_dot[k*16 + p] / dp[k*16 + p];
    v_vadv[k*32 + p]      = facm * (v[k*32 + p] - v[(k-1)*32 + p]);
    v_vadv[k*32 + 16 + p] = facm * (v[k*32 + 16 + p] - v[(k-1)*32 + 16 + p]);
  }
}

static void caar_element(orc_ctx *c, caar_ws *w, int ie, int np1, int nm1, int n0, double dt2,
                         double eta_ave_w, double scale1, double scale2, double scale3)
{
  const int nlev = c->nlev, nlevp = nlev + 1;
  const double *dp3d = c->dp3d + off_m(c, ie, n0);
  const double *vtheta_dp = c->vtheta_dp + off_m(c, ie, n0);
  const double *phi_i = c->phinh_i + off_i(c, ie, n0);
  const double *v = c->v + off_v(c, ie, n0);
  const double *w_i = c->w_i + off_i(c, ie, n0);
  const double *gradphis = c->gradphis + (size_t)ie*32;
  const double *fcor = c->fcor + (size_t)ie*16, *spheremp = c->spheremp + (size_t)ie*16;
  int err = 0;

  for (int q = 0; q < 16*nlev; q++) w->vtheta[q] = vtheta_dp[q] / dp3d[q];                 /* PA:1116 */
  orc_eos(c, vtheta_dp, dp3d, phi_i, w->pnh, w->exner, w->dpnh_dp_i, &err);                /* PA:1154 */
  if (err) c->error_flag = 1;

  for (int p = 0; p < NPSQ; p++) {                                                         /* PA:1156-1170 */
    w->dp3d_i[p] = dp3d[p];
    w->dp3d_i[nlev*16 + p] = dp3d[(nlev-1)*16 + p];
    for (int k = 1; k < nlev; k++) w->dp3d_i[k*16 + p] = (dp3d[k*16 + p] + dp3d[(k-1)*16 + p]) / 2;
    for (int cc = 0; cc < 2; cc++) {
      w->v_i[cc*16 + p] = v[cc*16 + p];
      w->v_i[nlev*32 + cc*16 + p] = v[(nlev-1)*32 + cc*16 + p];
      for (int k = 1; k < nlev; k++)
        w->v_i[k*32 + cc*16 + p] = (dp3d[k*16 + p] * v[k*32 + cc*16 + p] + dp3d[(k-1)*16 + p] * v[(k-1)*32 + cc*16 + p])
                                   / (2 * w->dp3d_i[k*16 + p]);
    }
  }

  for (int k = 0; k < nlev; k++) {                                                         /* PA:1182-1194 */
    for (int p = 0; p < NPSQ; p++) {
      w->phi[k*16 + p] = (phi_i[k*16 + p] + phi_i[(k+1)*16 + p]) / 2;
      w->vtemp[k*32 + p]      = v[k*32 + p] * dp3d[k*16 + p];
      w->vtemp[k*32 + 16 + p] = v[k*32 + 16 + p] * dp3d[k*16 + p];
    }
    if (c->vn0) {
      double *vn0 = c->vn0 + (size_t)ie*32*nlev;
      for (int p = 0; p < 32; p++) vn0[k*32 + p] = vn0[k*32 + p] + eta_ave_w * w->vtemp[k*32 + p];
    }
    orc_divergence_sphere(c, ie, w->vtemp + k*32, w->divdp + k*16);
    orc_vorticity_sphere(c, ie, v + k*32, w->vort + k*16);
  }

  for (int p = 0; p < NPSQ; p++) {                                                         /* PA:1197-1202 */
    w->pi_i[p] = c->hyai[0] * c->ps0;
    w->omega_i[p] = 0;
    for (int k = 0; k < nlev; k++) {
      w->pi_i[(k+1)*16 + p] = w->pi_i[k*16 + p] + dp3d[k*16 + p];
      w->omega_i[(k+1)*16 + p] = w->omega_i[k*16 + p] + w->divdp[k*16 + p];
    }
  }
  for (int k = 0; k < nlev; k++) {                                                         /* PA:1203-1213 */
    for (int p = 0; p < NPSQ; p++) w->pi[k*16 + p] = w->pi_i[k*16 + p] + dp3d[k*16 + p] / 2;
    orc_gradient_sphere(c, ie, w->pi + k*16, w->vtemp + k*32);
    for (int p = 0; p < NPSQ; p++) {
      w->vgrad_p[k*16 + p] = v[k*32 + p] * w->vtemp[k*32 + p] + v[k*32 + 16 + p] * w->vtemp[k*32 + 16 + p];
      w->omega[k*16 + p] = (w->vgrad_p[k*16 + p] - (w->omega_i[k*16 + p] + w->omega_i[(k+1)*16 + p]) / 2);
    }
  }

  if (c->rsplit > 0) {                                                                     /* PA:1219-1225 */
    memset(w->eta_dot_dpdn, 0, sizeof(double)*16*nlevp);
    memset(w->w_vadv_i, 0, sizeof(double)*16*nlevp);
    memset(w->phi_vadv_i, 0, sizeof(double)*16*nlevp);
    memset(w->theta_vadv, 0, sizeof(double)*16*nlev);
    memset(w->v_vadv, 0, sizeof(double)*32*nlev);
  } else {                                                                                 /* PA:1227-1302 */
    double sdot_sum[NPSQ];
    for (int p = 0; p < NPSQ; p++) {
      sdot_sum[p] = 0;
      for (int k = 0; k < nlev; k++) {
        sdot_sum[p] = sdot_sum[p] + w->divdp[k*16 + p];
        w->eta_dot_dpdn[(k+1)*16 + p] = sdot_sum[p];
      }
      for (int k = 0; k < nlev - 1; k++)
        w->eta_dot_dpdn[(k+1)*16 + p] = c->hybi[k+1] * sdot_sum[p] - w->eta_dot_dpdn[(k+1)*16 + p];
      w->eta_dot_dpdn[p] = 0;
      w->eta_dot_dpdn[nlev*16 + p] = 0;
      w->vtheta_i[p] = 0;
      w->vtheta_i[nlev*16 + p] = 0;
    }
    preq_vertadv_v1(nlev, v, w->eta_dot_dpdn, dp3d, w->v_vadv);
    for (int k = 1; k < nlev; k++)
      for (int p = 0; p < NPSQ; p++)
        w->vtheta_i[k*16 + p] = -w->dpnh_dp_i[k*16 + p] * (w->phi[k*16 + p] - w->phi[(k-1)*16 + p])
                                / (w->exner[k*16 + p] - w->exner[(k-1)*16 + p]) / Cp;
    for (int k = 0; k < nlev; k++)
      for (int p = 0; p < NPSQ; p++) {
        w->temp[k*16 + p] = ((w->eta_dot_dpdn[k*16 + p] + w->eta_dot_dpdn[(k+1)*16 + p]) / 2)
                            * (w_i[(k+1)*16 + p] - w_i[k*16 + p]);
        w->theta_vadv[k*16 + p] = w->eta_dot_dpdn[(k+1)*16 + p] * w->vtheta_i[(k+1)*16 + p]
                                - w->eta_dot_dpdn[k*16 + p] * w->vtheta_i[k*16 + p];
      }
    for (int p = 0; p < NPSQ; p++) {
      for (int k = 1; k < nlev; k++) {
        w->w_vadv_i[k*16 + p] = (w->temp[(k-1)*16 + p] + w->temp[k*16 + p]) / 2;
        w->phi_vadv_i[k*16 + p] = w->eta_dot_dpdn[k*16 + p] * (w->phi[k*16 + p] - w->phi[(k-1)*16 + p]);
      }
      w->w_vadv_i[p] = w->temp[p];
      w->w_vadv_i[nlev*16 + p] = w->temp[(nlev-1)*16 + p];
      w->phi_vadv_i[p] = 0;
      w->phi_vadv_i[nlev*16 + p] = 0;
    }
    for (int q = 0; q < 16*nlevp; q++) {
      w->w_vadv_i[q] = w->w_vadv_i[q] / w->dp3d_i[q];
      w->phi_vadv_i[q] = w->phi_vadv_i[q] / w->dp3d_i[q];
    }
  }

  if (c->eta_dot_dpdn) {                                                                   /* PA:1309-1316 */
    double *ed = c->eta_dot_dpdn + (size_t)ie*16*nlevp, *om = c->omega_p + (size_t)ie*16*nlev;
    for (int q = 0; q < 16*nlev; q++) {
      ed[q] = ed[q] + eta_ave_w * w->eta_dot_dpdn[q];
      om[q] = om[q] + eta_ave_w * w->omega[q];
    }
    for (int p = 0; p < NPSQ; p++)
      ed[nlev*16 + p] = ed[nlev*16 + p] + eta_ave_w * w->eta_dot_dpdn[nlev*16 + p];
  }

  for (int k = 0; k < nlevp; k++) {                                                        /* PA:1321-1361 */
    const double s2 = (k < nlev) ? scale2 : scale1;
    orc_gradient_sphere(c, ie, phi_i + k*16, w->gradphinh_i + k*32);
    orc_gradient_sphere(c, ie, w_i + k*16, w->gradw_i + k*32);
    for (int p = 0; p < NPSQ; p++) {
      w->v_gradw_i[k*16 + p] = w->v_i[k*32 + p] * w->gradw_i[k*32 + p] + w->v_i[k*32 + 16 + p] * w->gradw_i[k*32 + 16 + p];
      w->w_tens[k*16 + p] = (-w->w_vadv_i[k*16 + p] - w->v_gradw_i[k*16 + p]) * scale1
                            - s2 * g_ * (1 - w->dpnh_dp_i[k*16 + p]);
      w->v_gradphinh_i[k*16 + p] = w->v_i[k*32 + p] * w->gradphinh_i[k*32 + p]
                                 + w->v_i[k*32 + 16 + p] * w->gradphinh_i[k*32 + 16 + p];
      w->phi_tens[k*16 + p] = (-w->phi_vadv_i[k*16 + p] - w->v_gradphinh_i[k*16 + p]) * scale1
                              + s2 * g_ * w_i[k*16 + p];
      if (k < nlev && scale1 != scale2)
        w->phi_tens[k*16 + p] = w->phi_tens[k*16 + p] + (scale1 - scale2) *
            (w->v_i[k*32 + p] * gradphis[p] + w->v_i[k*32 + 16 + p] * gradphis[16 + p]) * c->hybi[k];
    }
  }

  for (int k = 0; k < nlev; k++) {                                                         /* PA:1370-1510 */
    const double *vk = v + k*32;
    if (c->theta_advect_form == 0) {
      for (int p = 0; p < NPSQ; p++) {
        w->v_theta[k*32 + p] = vk[p] * vtheta_dp[k*16 + p];
        w->v_theta[k*32 + 16 + p] = vk[16 + p] * vtheta_dp[k*16 + p];
      }
      orc_divergence_sphere(c, ie, w->v_theta + k*32, w->div_v_theta + k*16);
    } else if (c->theta_advect_form == 1) {
      orc_gradient_sphere(c, ie, w->vtheta + k*16, w->v_theta + k*32);
      for (int p = 0; p < NPSQ; p++)
        w->div_v_theta[k*16 + p] = w->vtheta[k*16 + p] * w->divdp[k*16 + p]
            + dp3d[k*16 + p] * vk[p] * w->v_theta[k*32 + p]
            + dp3d[k*16 + p] * vk[16 + p] * w->v_theta[k*32 + 16 + p];
    } else {
      for (int p = 0; p < NPSQ; p++) {
        w->v_theta[k*32 + p] = vk[p] * vtheta_dp[k*16 + p];
        w->v_theta[k*32 + 16 + p] = vk[16 + p] * vtheta_dp[k*16 + p];
      }
      orc_divergence_sphere(c, ie, w->v_theta + k*32, w->div_v_theta + k*16);
      for (int p = 0; p < NPSQ; p++) w->div_v_theta[k*16 + p] = w->div_v_theta[k*16 + p] / 2;
      orc_gradient_sphere(c, ie, w->vtheta + k*16, w->v_theta + k*32);
      for (int p = 0; p < NPSQ; p++)
        w->div_v_theta[k*16 + p] = w->div_v_theta[k*16 + p] + (
            w->vtheta[k*16 + p] * w->divdp[k*16 + p]
            + dp3d[k*16 + p] * vk[p] * w->v_theta[k*32 + p]
            + dp3d[k*16 + p] * vk[16 + p] * w->v_theta[k*32 + 16 + p]) / 2;
    }
    for (int p = 0; p < NPSQ; p++)
      w->theta_tens[k*16 + p] = (-w->theta_vadv[k*16 + p] - w->div_v_theta[k*16 + p]) * scale1;

    for (int p = 0; p < NPSQ; p++)                                                         /* PA:1401-1407 */
      w->temp[k*16 + p] = (w_i[k*16 + p] * w_i[k*16 + p] + w_i[(k+1)*16 + p] * w_i[(k+1)*16 + p]) / 4;
    orc_gradient_sphere(c, ie, w->temp + k*16, w->wvor + k*32);
    for (int cc = 0; cc < 2; cc++)
      for (int p = 0; p < NPSQ; p++)
        w->wvor[k*32 + cc*16 + p] = w->wvor[k*32 + cc*16 + p]
            - (w_i[k*16 + p] * w->gradw_i[k*32 + cc*16 + p] + w_i[(k+1)*16 + p] * w->gradw_i[(k+1)*32 + cc*16 + p]) / 2;

    for (int p = 0; p < NPSQ; p++) w->KE[k*16 + p] = (vk[p] * vk[p] + vk[16 + p] * vk[16 + p]) / 2;
    orc_gradient_sphere(c, ie, w->KE + k*16, w->gradKE + k*32);
    orc_gradient_sphere(c, ie, w->exner + k*16, w->gradexner + k*32);
    for (int cc = 0; cc < 2; cc++)
      for (int p = 0; p < NPSQ; p++)
        w->gradpterm[k*32 + cc*16 + p] = Cp * w->vtheta[k*16 + p] * w->gradexner[k*32 + cc*16 + p];

    if (c->theta_advect_form == 2) {                                                       /* PA:1446-1452 */
      for (int p = 0; p < NPSQ; p++) w->temp[k*16 + p] = w->vtheta[k*16 + p] * w->exner[k*16 + p];
      orc_gradient_sphere(c, ie, w->temp + k*16, w->vtemp + k*32);
      orc_gradient_sphere(c, ie, w->vtheta + k*16, w->v_theta + k*32);
      for (int cc = 0; cc < 2; cc++)
        for (int p = 0; p < NPSQ; p++)
          w->gradpterm[k*32 + cc*16 + p] = (w->gradpterm[k*32 + cc*16 + p]
              + Cp * (w->vtemp[k*32 + cc*16 + p] - w->exner[k*16 + p] * w->v_theta[k*32 + cc*16 + p])) / 2;
    }

    for (int cc = 0; cc < 2; cc++)                                                         /* PA:1455-1458 */
      for (int p = 0; p < NPSQ; p++)
        w->mgrad[k*32 + cc*16 + p] = (w->dpnh_dp_i[k*16 + p] * w->gradphinh_i[k*32 + cc*16 + p]
            + w->dpnh_dp_i[(k+1)*16 + p] * w->gradphinh_i[(k+1)*32 + cc*16 + p]) / 2;

    if (c->pgrad_correction == 1) {                                                        /* PA:1460-1477 */
      const double T0 = TREF - tref_lapse_rate * TREF * Cp / g_;
      for (int p = 0; p < NPSQ; p++) w->temp[k*16 + p] = log(w->exner[k*16 + p]);
      orc_gradient_sphere(c, ie, w->temp + k*16, w->vtemp + k*32);
      for (int cc = 0; cc < 2; cc++)
        for (int p = 0; p < NPSQ; p++)
          w->mgrad[k*32 + cc*16 + p] = w->mgrad[k*32 + cc*16 + p]
              + Cp * T0 * (w->vtemp[k*32 + cc*16 + p] - w->gradexner[k*32 + cc*16 + p] / w->exner[k*16 + p]);
    }

    for (int p = 0; p < NPSQ; p++) {                                                       /* PA:1479-1509 */
      double v1 = vk[p], v2 = vk[16 + p];
      w->vtens1[k*16 + p] = (-w->v_vadv[k*32 + p]
          + v2 * (fcor[p] + w->vort[k*16 + p])
          - w->gradKE[k*32 + p] - w->mgrad[k*32 + p]
          - w->gradpterm[k*32 + p]
          - w->wvor[k*32 + p]) * scale1;
      w->vtens2[k*16 + p] = (-w->v_vadv[k*32 + 16 + p]
          - v1 * (fcor[p] + w->vort[k*16 + p])
          - w->gradKE[k*32 + 16 + p] - w->mgrad[k*32 + 16 + p]
          - w->gradpterm[k*32 + 16 + p]
          - w->wvor[k*32 + 16 + p]) * scale1;
    }
  }

  /* PA:1674-1702 ; np1 may alias n0 or nm1: every tendency is already in the workspace */
  {
    double *v1_ = c->v + off_v(c, ie, np1), *th1 = c->vtheta_dp + off_m(c, ie, np1);
    double *w1 = c->w_i + off_i(c, ie, np1), *ph1 = c->phinh_i + off_i(c, ie, np1);
    double *dp1 = c->dp3d + off_m(c, ie, np1);
    const double *vm = c->v + off_v(c, ie, nm1), *thm = c->vtheta_dp + off_m(c, ie, nm1);
    const double *wm = c->w_i + off_i(c, ie, nm1), *phm = c->phinh_i + off_i(c, ie, nm1);
    const double *dpm = c->dp3d + off_m(c, ie, nm1);
    for (int k = 0; k < nlev; k++)
      for (int p = 0; p < NPSQ; p++) {
        v1_[k*32 + p]      = spheremp[p] * (scale3 * vm[k*32 + p] + dt2 * w->vtens1[k*16 + p]);
        v1_[k*32 + 16 + p] = spheremp[p] * (scale3 * vm[k*32 + 16 + p] + dt2 * w->vtens2[k*16 + p]);
        th1[k*16 + p] = spheremp[p] * (scale3 * thm[k*16 + p] + dt2 * w->theta_tens[k*16 + p]);
        w1[k*16 + p]  = spheremp[p] * (scale3 * wm[k*16 + p] + dt2 * w->w_tens[k*16 + p]);
        ph1[k*16 + p] = spheremp[p] * (scale3 * phm[k*16 + p] + dt2 * w->phi_tens[k*16 + p]);
        dp1[k*16 + p] = spheremp[p] * (scale3 * dpm[k*16 + p]
            - scale1 * dt2 * (w->divdp[k*16 + p] + w->eta_dot_dpdn[(k+1)*16 + p] - w->eta_dot_dpdn[k*16 + p]));
      }
    for (int p = 0; p < NPSQ; p++)
      w1[nlev*16 + p] = spheremp[p] * (scale3 * wm[nlev*16 + p] + dt2 * w->w_tens[nlev*16 + p]);
  }
}

/* PA:1113-1718: all elements up to (not including) the boundary exchange */
void orc_caar_pre(orc_ctx *c, int np1, int nm1, int n0, double dt2, double eta_ave_w,
                  double scale1, double scale2, double scale3)
{
#pragma omp parallel num_threads(c->nthreads)
  {
    caar_ws w;
    ws_alloc(&w, c->nlev);
#pragma omp for schedule(static)
    for (int ie = 0; ie < c->nelem; ie++)
      caar_element(c, &w, ie, np1, nm1, n0, dt2, eta_ave_w, scale1, scale2, scale3);
    ws_free(&w);
  }
}

/* PA:1023-1819 */
void orc_caar(orc_ctx *c, int np1, int nm1, int n0, double dt2, double eta_ave_w,
              double scale1, double scale2, double scale3)
{
  const int nlev = c->nlev, nlevp = nlev + 1;
  orc_caar_pre(c, np1, nm1, n0, dt2, eta_ave_w, scale1, scale2, scale3);

  /* PA:1705-1736: pack, bndry_exchangeV, unpack == node sums */
  orc_dss_field(c, c->v + off_v(c, 0, np1), 2*nlev, ORC_TL*32*nlev, c->scratch1);
  orc_dss_field(c, c->dp3d + off_m(c, 0, np1), nlev, ORC_TL*16*nlev, c->scratch1);
  orc_dss_field(c, c->vtheta_dp + off_m(c, 0, np1), nlev, ORC_TL*16*nlev, c->scratch1);
  orc_dss_field(c, c->w_i + off_i(c, 0, np1), nlevp, ORC_TL*16*nlevp, c->scratch1);
  orc_dss_field(c, c->phinh_i + off_i(c, 0, np1), nlev, ORC_TL*16*nlevp, c->scratch1);

#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < c->nelem; ie++) {
    double *v1 = c->v + off_v(c, ie, np1), *th1 = c->vtheta_dp + off_m(c, ie, np1);
    double *w1 = c->w_i + off_i(c, ie, np1), *ph1 = c->phinh_i + off_i(c, ie, np1);
    double *dp1 = c->dp3d + off_m(c, ie, np1);
    const double *rspheremp = c->rspheremp + (size_t)ie*16, *gradphis = c->gradphis + (size_t)ie*32;
    for (int k = 0; k < nlev; k++)                                                         /* PA:1741-1753 */
      for (int p = 0; p < NPSQ; p++) {
        dp1[k*16 + p] = rspheremp[p] * dp1[k*16 + p];
        th1[k*16 + p] = rspheremp[p] * th1[k*16 + p];
        w1[k*16 + p]  = rspheremp[p] * w1[k*16 + p];
        ph1[k*16 + p] = rspheremp[p] * ph1[k*16 + p];
        v1[k*32 + p]      = rspheremp[p] * v1[k*32 + p];
        v1[k*32 + 16 + p] = rspheremp[p] * v1[k*32 + 16 + p];
      }
    for (int p = 0; p < NPSQ; p++) w1[nlev*16 + p] = rspheremp[p] * w1[nlev*16 + p];
    for (int p = 0; p < NPSQ; p++) {                                                       /* PA:1759-1772 */
      double *vb = v1 + (nlev-1)*32;
      double dpnh = 1 + (((vb[p] * gradphis[p] + vb[16 + p] * gradphis[16 + p]) / g_ - w1[nlev*16 + p])
                         / (g_ + (gradphis[p] * gradphis[p] + gradphis[16 + p] * gradphis[16 + p]) / (2 * g_))) / dt2;
      w1[nlev*16 + p] = w1[nlev*16 + p] + scale1 * dt2 * g_ * (dpnh - 1);
      vb[p]      = vb[p]      - scale1 * dt2 * (dpnh - 1) * gradphis[p] / 2;
      vb[16 + p] = vb[16 + p] - scale1 * dt2 * (dpnh - 1) * gradphis[16 + p] / 2;
    }
    if (scale3 != 0) orc_limiter_dp3d_k(c, dp1, th1, c->spheremp + (size_t)ie*16);         /* PA:1812-1815 */
  }
}

/* ---------------- DIRK ---------------- */
/* IM:26-54 ; gwh_i (np,np,nlevp) */
static void compute_gwphis(const orc_ctx *c, double *gwh_i, const double *dp3d, const double *v,
                           const double *gradphis)
{
  const int nlev = c->nlev;
  for (int p = 0; p < NPSQ; p++) {
    gwh_i[p] = 0;
    gwh_i[nlev*16 + p] = 0;
    for (int k = 1; k < nlev; k++) {
      double vi1 = (dp3d[k*16 + p] * v[k*32 + p] + dp3d[(k-1)*16 + p] * v[(k-1)*32 + p]) / (dp3d[k*16 + p] + dp3d[(k-1)*16 + p]);
      double vi2 = (dp3d[k*16 + p] * v[k*32 + 16 + p] + dp3d[(k-1)*16 + p] * v[(k-1)*32 + 16 + p]) / (dp3d[k*16 + p] + dp3d[(k-1)*16 + p]);
      gwh_i[k*16 + p] = (vi1 * gradphis[p] + vi2 * gradphis[16 + p]) * c->hybi[k];
    }
  }
}

/* IM:410-433 (analytic). JacL, JacU: (np,np,nlev-1); JacD: (np,np,nlev) */
void orc_dirk_jacobian(const orc_ctx *c, double *JacL, double *JacD, double *JacU, double dt2,
                       const double *dp3d, const double *dphi, const double *pnh)
{
  const int nlev = c->nlev;
  const double a = (dt2 * g_) * (dt2 * g_) / (1 - KAPPA);
  for (int p = 0; p < NPSQ; p++) {
    int k = 0;
    double b = a / dp3d[k*16 + p];
    double ck = pnh[k*16 + p] / dphi[k*16 + p], ckm1;
    JacU[k*16 + p] = 2 * b * ck;
    JacD[k*16 + p] = 1 - JacU[k*16 + p];
    ckm1 = ck;
    for (k = 1; k < nlev - 1; k++) {
      b = 2 * a / (dp3d[(k-1)*16 + p] + dp3d[k*16 + p]);
      ck = pnh[k*16 + p] / dphi[k*16 + p];
      JacL[(k-1)*16 + p] = b * ckm1;
      JacU[k*16 + p] = b * ck;
      JacD[k*16 + p] = 1 - JacL[(k-1)*16 + p] - JacU[k*16 + p];
      ckm1 = ck;
    }
    k = nlev - 1;
    b = 2 * a / (dp3d[k*16 + p] + dp3d[(k-1)*16 + p]);
    ck = pnh[k*16 + p] / dphi[k*16 + p];
    JacL[(k-1)*16 + p] = b * ckm1;
    JacD[k*16 + p] = 1 - JacL[(k-1)*16 + p] - b * ck;
  }
}

/* IM:576-599 ; arrays (npsq, nlev) */
void orc_tridiag(int nlev, const double *dl, double *d, const double *du, double *x)
{
  for (int p = 0; p < NPSQ; p++) {
    for (int k = 0; k < nlev - 1; k++) {
      double dlk = dl[k*16 + p] / d[k*16 + p];
      d[(k+1)*16 + p] = d[(k+1)*16 + p] - dlk * du[k*16 + p];
      x[(k+1)*16 + p] = x[(k+1)*16 + p] - dlk * x[k*16 + p];
    }
    x[(nlev-1)*16 + p] = x[(nlev-1)*16 + p] / d[(nlev-1)*16 + p];
    for (int k = nlev - 2; k >= 0; k--)
      x[k*16 + p] = (x[k*16 + p] - du[k*16 + p] * x[k*16 + p + 16]) / d[k*16 + p];
  }
}

/* IM:167-358, one element. Returns itercount; *deltaerr_out, *reserr_out per element. */
static int dirk_element(orc_ctx *c, int ie, int nm1, double alphadt_nm1, int n0, double alphadt_n0,
                        int np1, double dt2, double *ws, double *deltaerr_out, double *reserr_out)
{
  const int nlev = c->nlev, nlevp = nlev + 1, maxiter = 20;
  const double deltatol = 1.0e-11;
  const size_t L = 16*(size_t)nlevp;
  double *JacD = ws, *JacL = ws + L, *JacU = ws + 2*L, *pnh = ws + 3*L, *dpnh_dp_i = ws + 4*L,
         *exner = ws + 5*L, *w_n0 = ws + 6*L, *dphi = ws + 7*L, *dphi_n0 = ws + 8*L,
         *phi_n0 = ws + 9*L, *Fn = ws + 10*L, *x = ws + 11*L, *gwh_i = ws + 12*L;
  double *phi_np1 = c->phinh_i + off_i(c, ie, np1), *w_np1 = c->w_i + off_i(c, ie, np1);
  const double *gradphis = c->gradphis + (size_t)ie*32;
  const double *th1 = c->vtheta_dp + off_m(c, ie, np1), *dp1 = c->dp3d + off_m(c, ie, np1);
  int err = 0;

  memcpy(phi_n0, phi_np1, sizeof(double)*L);                                               /* IM:168-170 */
  memcpy(w_n0, w_np1, sizeof(double)*L);
  double wmax = 0;
  for (size_t q = 0; q < L; q++) wmax = fmax(wmax, fabs(w_n0[q]));
  wmax = fmax(1.0, wmax);

  for (int pass = 0; pass < 2; pass++) {                                                   /* IM:176-205 */
    const double dt3 = pass == 0 ? alphadt_n0 : alphadt_nm1;
    const int nt = pass == 0 ? n0 : nm1;
    if (dt3 == 0.0) continue;
    orc_eos(c, c->vtheta_dp + off_m(c, ie, nt), c->dp3d + off_m(c, ie, nt), c->phinh_i + off_i(c, ie, nt),
            pnh, exner, dpnh_dp_i, &err);
    for (int q = 0; q < 16*nlev; q++) w_n0[q] = w_n0[q] + dt3 * g_ * (dpnh_dp_i[q] - 1);
    compute_gwphis(c, gwh_i, c->dp3d + off_m(c, ie, nt), c->v + off_v(c, ie, nt), gradphis);
    const double *wnt = c->w_i + off_i(c, ie, nt);
    for (int q = 0; q < 16*nlev; q++) phi_n0[q] = phi_n0[q] + dt3 * g_ * wnt[q] - dt3 * gwh_i[q];
  }
  compute_gwphis(c, gwh_i, dp1, c->v + off_v(c, ie, np1), gradphis);                       /* IM:208-211 */
  for (int q = 0; q < 16*nlev; q++) phi_n0[q] = phi_n0[q] - dt2 * gwh_i[q];

  orc_phi_from_eos(c, c->phis + (size_t)ie*16, th1, dp1, phi_np1);                         /* IM:228-230 */
  for (int q = 0; q < 16*nlev; q++) w_np1[q] = (phi_np1[q] - phi_n0[q]) / (dt2 * g_);

  int nsafe = 0;                                                                           /* IM:234-260 */
  for (int q = 0; q < 16*nlev; q++) {
    dphi[q] = phi_np1[q + 16] - phi_np1[q];
    dphi_n0[q] = phi_n0[q + 16] - phi_n0[q];
    if (dphi[q] > -g_) { dphi[q] = -g_; nsafe = 1; }
  }
  if (nsafe == 1) {
    for (int k = nlev - 1; k >= 0; k--)
      for (int p = 0; p < NPSQ; p++) phi_np1[k*16 + p] = phi_np1[(k+1)*16 + p] - dphi[k*16 + p];
    for (int q = 0; q < 16*nlev; q++) w_np1[q] = (phi_np1[q] - phi_n0[q]) / (dt2 * g_);
  }
  eos2(c, th1, dp1, dphi, pnh, exner, dpnh_dp_i, &err);                                    /* IM:261-264 */
  for (int q = 0; q < 16*nlev; q++) Fn[q] = w_np1[q] - (w_n0[q] + g_ * dt2 * (dpnh_dp_i[q] - 1));

  int itercount = 0;
  double deltaerr = 0;
  while (itercount < maxiter) {                                                            /* IM:268-336 */
    orc_dirk_jacobian(c, JacL, JacD, JacU, dt2, dp1, dphi, pnh);
    for (int q = 0; q < 16*nlev; q++) x[q] = -Fn[q];
    orc_tridiag(nlev, JacL, JacD, JacU, x);
    int anypos = 0;
    for (int p = 0; p < NPSQ; p++) {
      for (int k = 0; k < nlev - 1; k++)
        dphi[k*16 + p] = dphi_n0[k*16 + p] + dt2 * g_ * ((w_np1[(k+1)*16 + p] - w_np1[k*16 + p]) + (x[(k+1)*16 + p] - x[k*16 + p]));
      dphi[(nlev-1)*16 + p] = dphi_n0[(nlev-1)*16 + p] - dt2 * g_ * (w_np1[(nlev-1)*16 + p] + x[(nlev-1)*16 + p]);
    }
    for (int q = 0; q < 16*nlev; q++) if (dphi[q] >= 0) anypos = 1;
    double alphas[NPSQ];
    for (int p = 0; p < NPSQ; p++) alphas[p] = 1;
    if (anypos) {                                                                          /* IM:285-319 */
      for (int p = 0; p < NPSQ; p++) {
        int colpos = 0;
        for (int k = 0; k < nlev; k++) if (dphi[k*16 + p] >= 0) colpos = 1;
        if (!colpos) continue;
        double alpha = 1;
        for (int k = 0; k < nlev; k++) {
          double dx, dw;
          if (k < nlev - 1) { dx = x[(k+1)*16 + p] - x[k*16 + p]; dw = w_np1[(k+1)*16 + p] - w_np1[k*16 + p]; }
          else { dx = -x[k*16 + p]; dw = -w_np1[k*16 + p]; }
          if (dx != 0) {
            double alpha_k = -(dphi_n0[k*16 + p] + dt2 * g_ * dw) / (dt2 * g_ * dx);
            if (alpha_k >= 0) alpha = fmin(alpha, alpha_k);
          }
        }
        alpha = alpha / 2;
        alphas[p] = alpha;
        for (int k = 0; k < nlev - 1; k++)
          dphi[k*16 + p] = dphi_n0[k*16 + p] + dt2 * g_ * ((w_np1[(k+1)*16 + p] - w_np1[k*16 + p]) + alpha * (x[(k+1)*16 + p] - x[k*16 + p]));
        dphi[(nlev-1)*16 + p] = dphi_n0[(nlev-1)*16 + p] - dt2 * g_ * (w_np1[(nlev-1)*16 + p] + alpha * x[(nlev-1)*16 + p]);
      }
    }
    for (int k = 0; k < nlev; k++)
      for (int p = 0; p < NPSQ; p++) w_np1[k*16 + p] = w_np1[k*16 + p] + alphas[p] * x[k*16 + p];
    eos2(c, th1, dp1, dphi, pnh, exner, dpnh_dp_i, &err);
    for (int q = 0; q < 16*nlev; q++) Fn[q] = w_np1[q] - (w_n0[q] + g_ * dt2 * (dpnh_dp_i[q] - 1));
    double xmax = 0;
    for (int q = 0; q < 16*nlev; q++) xmax = fmax(xmax, fabs(x[q]));
    deltaerr = xmax / wmax;
    itercount = itercount + 1;
    if (deltaerr < deltatol) break;
  }
  for (int q = 0; q < 16*nlev; q++) phi_np1[q] = phi_n0[q] + dt2 * g_ * w_np1[q];           /* IM:339 */
  double fmaxv = 0;
  for (int q = 0; q < 16*nlev; q++) fmaxv = fmax(fmaxv, fabs(Fn[q]));
  *reserr_out = fmaxv / (wmax * fabs(dt2));
  *deltaerr_out = deltaerr;
  if (err) c->error_flag = 1;
  return itercount;
}

/* IM:60-364 */
void orc_dirk(orc_ctx *c, int nm1, double alphadt_nm1, int n0, double alphadt_n0, int np1,
              double dt2, int *itercount, double *itererr)
{
  int max_it = c->max_itercnt;
  double max_de = c->max_deltaerr, max_re = c->max_reserr;
  int last_it = 0;
#pragma omp parallel num_threads(c->nthreads) reduction(max : max_it, max_de, max_re, last_it)
  {
    double *ws = (double *)malloc(sizeof(double)*13*16*(size_t)(c->nlev + 1));
#pragma omp for schedule(static)
    for (int ie = 0; ie < c->nelem; ie++) {
      double de, re;
      int it = dirk_element(c, ie, nm1, alphadt_nm1, n0, alphadt_n0, np1, dt2, ws, &de, &re);
      if (it > max_it) max_it = it;
      if (de > max_de) max_de = de;
      if (re > max_re) max_re = re;
      if (it > last_it) last_it = it;
    }
    free(ws);
  }
  c->max_itercnt = max_it; c->max_deltaerr = max_de; c->max_reserr = max_re;
  if (itercount) *itercount = last_it;  /* Fortran returns the last element's count; we return the max */
  if (itererr) *itererr = fmin(max_re, max_de);
}

/* applyCAMforcing_dynamics, PA:517-541 (standalone HOMME: the #ifndef CAM branch forces w_i with FM(:,:,3,:)).
 * fm = derived%FM(np,np,3,nlev,ie), fvtheta = FVTheta(np,np,nlev,ie), fphi = FPHI(np,np,nlevp,ie). */
void orc_apply_forcing_dynamics(orc_ctx *c, int np1, double dt, const double *fm, const double *fvtheta,
                                const double *fphi)
{
  const int nlev = c->nlev, nlevp = nlev + 1;
  for (int ie = 0; ie < c->nelem; ie++) {
    double *v = c->v + off_v(c, ie, np1), *th = c->vtheta_dp + off_m(c, ie, np1);
    double *w = c->w_i + off_i(c, ie, np1), *ph = c->phinh_i + off_i(c, ie, np1);
    const double *f = fm + (size_t)ie*48*nlev, *ft = fvtheta + (size_t)ie*16*nlev, *fp = fphi + (size_t)ie*16*nlevp;
    const double *gradphis = c->gradphis + (size_t)ie*32;
    for (int k = 0; k < nlev; k++)
      for (int p = 0; p < NPSQ; p++) {
        th[k*16 + p] = th[k*16 + p] + dt * ft[k*16 + p];
        ph[k*16 + p] = ph[k*16 + p] + dt * fp[k*16 + p];
        v[k*32 + p]      = v[k*32 + p]      + dt * f[k*48 + p];
        v[k*32 + 16 + p] = v[k*32 + 16 + p] + dt * f[k*48 + 16 + p];
        w[k*16 + p] = w[k*16 + p] + dt * f[k*48 + 32 + p];
      }
    for (int p = 0; p < NPSQ; p++)
      w[nlev*16 + p] = (v[(nlev-1)*32 + p] * gradphis[p] + v[(nlev-1)*32 + 16 + p] * gradphis[16 + p]) / g_;
  }
}

/* ---------------- hyperviscosity ---------------- */
/* stens layout (np,np,nlev,4,ie): dp3d, theta, w, phi ; vtens (np,np,2,nlev,ie).
 * scratch1 = [ie][ vtens 32*nlev | stens 4*16*nlev ] so that one DSS call covers 6*nlev layers. */
static void hv_laplace_pass(orc_ctx *c, int nt, int second, double nu_ratio, double *buf)
{
  const int nlev = c->nlev;
  const size_t es = 96*(size_t)nlev;
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < c->nelem; ie++) {
    double *vt = buf + ie*es, *st = vt + 32*nlev;
    const double *rs = c->rspheremp + (size_t)ie*16;
    double tmp[32];
    for (int k = 0; k < nlev; k++) {
      if (!second) {                                                                       /* VT:93-107 */
        orc_laplace_sphere_wk(c, ie, c->dp3d + off_m(c, ie, nt) + k*16, 0, st + (0*nlev + k)*16);
        orc_laplace_sphere_wk(c, ie, c->vtheta_dp + off_m(c, ie, nt) + k*16, 0, st + (1*nlev + k)*16);
        orc_laplace_sphere_wk(c, ie, c->w_i + off_i(c, ie, nt) + k*16, 0, st + (2*nlev + k)*16);
        orc_laplace_sphere_wk(c, ie, c->phinh_i + off_i(c, ie, nt) + k*16, 0, st + (3*nlev + k)*16);
        orc_vlaplace_sphere_wk(c, ie, c->v + off_v(c, ie, nt) + k*32, 0, nu_ratio, vt + k*32);
      } else {                                                                             /* VT:145-163 */
        for (int f = 0; f < 4; f++) {
          for (int p = 0; p < NPSQ; p++) tmp[p] = rs[p] * st[(f*nlev + k)*16 + p];
          orc_laplace_sphere_wk(c, ie, tmp, c->hypervis_scaling_on, st + (f*nlev + k)*16);
        }
        for (int p = 0; p < NPSQ; p++) { tmp[p] = rs[p] * vt[k*32 + p]; tmp[16 + p] = rs[p] * vt[k*32 + 16 + p]; }
        orc_vlaplace_sphere_wk(c, ie, tmp, c->hypervis_scaling_on, nu_ratio, vt + k*32);
      }
    }
  }
}

/* PA:546-862 */
void orc_advance_hypervis(orc_ctx *c, int nt, double dt2, double eta_ave_w)
{
  const int nlev = c->nlev, nlevp = nlev + 1, nelem = c->nelem;
  const size_t es = 96*(size_t)nlev;
  double *buf = c->scratch1;
  (void)eta_ave_w; /* dpdiss_ave / dpdiss_biharmonic (PA:651-656) are diagnostics outside the prognostic state */

#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < nelem; ie++) {                                                     /* PA:615-621 */
    double *th = c->vtheta_dp + off_m(c, ie, nt); const double *dp = c->dp3d + off_m(c, ie, nt);
    for (int q = 0; q < 16*nlev; q++) th[q] = th[q] / dp[q];
  }

  double dt = dt2 / c->hypervis_subcycle;
  for (int ic = 0; ic < c->hypervis_subcycle; ic++) {
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
    for (int ie = 0; ie < nelem; ie++) {                                                   /* PA:629-636 */
      double *th = c->vtheta_dp + off_m(c, ie, nt), *dp = c->dp3d + off_m(c, ie, nt), *ph = c->phinh_i + off_i(c, ie, nt);
      for (int q = 0; q < 16*nlev; q++) { th[q] = th[q] - c->theta_ref[(size_t)ie*16*nlev + q]; dp[q] = dp[q] - c->dp_ref[(size_t)ie*16*nlev + q]; }
      for (int q = 0; q < 16*nlevp; q++) ph[q] = ph[q] - c->phi_ref[(size_t)ie*16*nlevp + q];
    }
    /* biharmonic_wk_theta VT:39-169 */
    double nu_ratio1 = 1, nu_ratio2 = 1;
    if (c->nu_div != c->nu) { nu_ratio1 = c->nu_div / c->nu; nu_ratio2 = 1; }
    hv_laplace_pass(c, nt, 0, nu_ratio1, buf);
    orc_dss_field(c, buf, 6*nlev, (int)es, c->scratch2);
    hv_laplace_pass(c, nt, 1, nu_ratio2, buf);

#pragma omp parallel for num_threads(c->nthreads) schedule(static)
    for (int ie = 0; ie < nelem; ie++) {                                                   /* PA:642-663 */
      double *th = c->vtheta_dp + off_m(c, ie, nt), *dp = c->dp3d + off_m(c, ie, nt), *ph = c->phinh_i + off_i(c, ie, nt);
      for (int q = 0; q < 16*nlev; q++) { th[q] = th[q] + c->theta_ref[(size_t)ie*16*nlev + q]; dp[q] = dp[q] + c->dp_ref[(size_t)ie*16*nlev + q]; }
      for (int q = 0; q < 16*nlevp; q++) ph[q] = ph[q] + c->phi_ref[(size_t)ie*16*nlevp + q];
      double *vt = buf + ie*es, *st = vt + 32*nlev;
      for (int q = 0; q < 32*nlev; q++) vt[q] = -c->nu * vt[q];
      for (int q = 0; q < 16*nlev; q++) {
        st[0*16*nlev + q] = -c->nu_p * st[0*16*nlev + q];
        st[1*16*nlev + q] = -c->nu   * st[1*16*nlev + q];
        st[2*16*nlev + q] = -c->nu   * st[2*16*nlev + q];
        st[3*16*nlev + q] = -c->nu_s * st[3*16*nlev + q];
      }
    }
    orc_dss_field(c, buf, 6*nlev, (int)es, c->scratch2);                                   /* PA:665-679 */
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
    for (int ie = 0; ie < nelem; ie++) {                                                   /* PA:692-758 */
      double *vt = buf + ie*es, *st = vt + 32*nlev;
      const double *rs = c->rspheremp + (size_t)ie*16;
      double *v = c->v + off_v(c, ie, nt), *th = c->vtheta_dp + off_m(c, ie, nt), *dp = c->dp3d + off_m(c, ie, nt);
      double *w = c->w_i + off_i(c, ie, nt), *ph = c->phinh_i + off_i(c, ie, nt);
      for (int k = 0; k < nlev; k++)
        for (int p = 0; p < NPSQ; p++) {
          vt[k*32 + p] = dt * vt[k*32 + p] * rs[p];
          vt[k*32 + 16 + p] = dt * vt[k*32 + 16 + p] * rs[p];
          for (int f = 0; f < 4; f++) st[(f*nlev + k)*16 + p] = dt * st[(f*nlev + k)*16 + p] * rs[p];
          v[k*32 + p] = v[k*32 + p] + vt[k*32 + p];
          v[k*32 + 16 + p] = v[k*32 + 16 + p] + vt[k*32 + 16 + p];
          w[k*16 + p] = w[k*16 + p] + st[(2*nlev + k)*16 + p];
          dp[k*16 + p] = dp[k*16 + p] + st[(0*nlev + k)*16 + p];
          ph[k*16 + p] = ph[k*16 + p] + st[(3*nlev + k)*16 + p];
          th[k*16 + p] = th[k*16 + p] + st[(1*nlev + k)*16 + p];
        }
    }
  }

#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < nelem; ie++) {                                                     /* PA:767-774 */
    double *th = c->vtheta_dp + off_m(c, ie, nt); const double *dp = c->dp3d + off_m(c, ie, nt);
    const double *v = c->v + off_v(c, ie, nt), *gp = c->gradphis + (size_t)ie*32;
    double *w = c->w_i + off_i(c, ie, nt);
    for (int q = 0; q < 16*nlev; q++) th[q] = th[q] * dp[q];
    for (int p = 0; p < NPSQ; p++)
      w[nlev*16 + p] = (v[(nlev-1)*32 + p] * gp[p] + v[(nlev-1)*32 + 16 + p] * gp[16 + p]) / g_;
  }

  if (c->nu_top > 0 && c->hypervis_subcycle_tom > 0) {                                     /* PA:780-856 */
    const int nt_ = c->nlev_tom;
    const size_t est = 96*(size_t)nt_;
    dt = dt2 / c->hypervis_subcycle_tom;
    for (int ic = 0; ic < c->hypervis_subcycle_tom; ic++) {
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
      for (int ie = 0; ie < nelem; ie++) {
        double *vt = buf + ie*est, *st = vt + 32*nt_;
        double lap_s[16], lap_v[32];
        for (int k = 0; k < nt_; k++) {
          const double xfac = dt * c->nu_scale_top[k] * c->nu_top;
          const double *fld[4] = { c->dp3d + off_m(c, ie, nt) + k*16, c->vtheta_dp + off_m(c, ie, nt) + k*16,
                                   c->w_i + off_i(c, ie, nt) + k*16, c->phinh_i + off_i(c, ie, nt) + k*16 };
          for (int f = 0; f < 4; f++) {
            orc_laplace_sphere_wk(c, ie, fld[f], 0, lap_s); /* laplace_scaling == 0 */
            for (int p = 0; p < NPSQ; p++) st[(f*nt_ + k)*16 + p] = xfac * lap_s[p];
          }
          orc_vlaplace_sphere_wk(c, ie, c->v + off_v(c, ie, nt) + k*32, 0, 1.0, lap_v);
          for (int p = 0; p < 32; p++) vt[k*32 + p] = xfac * lap_v[p];
        }
      }
      orc_dss_field(c, buf, 6*nt_, (int)est, c->scratch2);
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
      for (int ie = 0; ie < nelem; ie++) {                                                 /* PA:837-852 */
        double *vt = buf + ie*est, *st = vt + 32*nt_;
        const double *rs = c->rspheremp + (size_t)ie*16;
        double *v = c->v + off_v(c, ie, nt), *th = c->vtheta_dp + off_m(c, ie, nt), *dp = c->dp3d + off_m(c, ie, nt);
        double *w = c->w_i + off_i(c, ie, nt), *ph = c->phinh_i + off_i(c, ie, nt);
        for (int k = 0; k < nt_; k++)
          for (int p = 0; p < NPSQ; p++) {
            v[k*32 + p] = v[k*32 + p] + vt[k*32 + p] * rs[p];
            v[k*32 + 16 + p] = v[k*32 + 16 + p] + vt[k*32 + 16 + p] * rs[p];
            w[k*16 + p] = w[k*16 + p] + st[(2*nt_ + k)*16 + p] * rs[p];
            dp[k*16 + p] = dp[k*16 + p] + st[(0*nt_ + k)*16 + p] * rs[p];
            ph[k*16 + p] = ph[k*16 + p] + st[(3*nt_ + k)*16 + p] * rs[p];
            th[k*16 + p] = th[k*16 + p] + st[(1*nt_ + k)*16 + p] * rs[p];
          }
      }
    }
  }
}

/* ---------------- prim_advance_exp (PA:81-513) ---------------- */
/* X(tl_dst) = a*X(tl_dst) + b*X(tl_src) style combos are written out as in the Fortran */
static void combo_tt5(orc_ctx *c, int nm1, int n0)
{ /* PA:197-208: u(nm1) = (5*u(nm1) - u(n0))/4 ; phinh_i only 1:nlev */
  const int nlev = c->nlev, nlevp = nlev + 1;
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < c->nelem; ie++) {
    double *a; const double *b;
    a = c->v + off_v(c, ie, nm1); b = c->v + off_v(c, ie, n0);
    for (int q = 0; q < 32*nlev; q++) a[q] = (5 * a[q] - b[q]) / 4;
    a = c->vtheta_dp + off_m(c, ie, nm1); b = c->vtheta_dp + off_m(c, ie, n0);
    for (int q = 0; q < 16*nlev; q++) a[q] = (5 * a[q] - b[q]) / 4;
    a = c->dp3d + off_m(c, ie, nm1); b = c->dp3d + off_m(c, ie, n0);
    for (int q = 0; q < 16*nlev; q++) a[q] = (5 * a[q] - b[q]) / 4;
    a = c->w_i + off_i(c, ie, nm1); b = c->w_i + off_i(c, ie, n0);
    for (int q = 0; q < 16*nlevp; q++) a[q] = (5 * a[q] - b[q]) / 4;
    a = c->phinh_i + off_i(c, ie, nm1); b = c->phinh_i + off_i(c, ie, n0);
    for (int q = 0; q < 16*nlev; q++) a[q] = (5 * a[q] - b[q]) / 4;
  }
}

static void combo_tt9(orc_ctx *c, int nm1, int n0, int np1)
{ /* PA:313-326: u(np1) += (u(nm1)-u(n0))/4, then limiter */
  const int nlev = c->nlev, nlevp = nlev + 1;
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < c->nelem; ie++) {
    double *a; const double *b, *d;
    a = c->v + off_v(c, ie, np1); b = c->v + off_v(c, ie, nm1); d = c->v + off_v(c, ie, n0);
    for (int q = 0; q < 32*nlev; q++) a[q] = a[q] + (b[q] - d[q]) / 4;
    a = c->vtheta_dp + off_m(c, ie, np1); b = c->vtheta_dp + off_m(c, ie, nm1); d = c->vtheta_dp + off_m(c, ie, n0);
    for (int q = 0; q < 16*nlev; q++) a[q] = a[q] + (b[q] - d[q]) / 4;
    a = c->dp3d + off_m(c, ie, np1); b = c->dp3d + off_m(c, ie, nm1); d = c->dp3d + off_m(c, ie, n0);
    for (int q = 0; q < 16*nlev; q++) a[q] = a[q] + (b[q] - d[q]) / 4;
    a = c->w_i + off_i(c, ie, np1); b = c->w_i + off_i(c, ie, nm1); d = c->w_i + off_i(c, ie, n0);
    for (int q = 0; q < 16*nlevp; q++) a[q] = a[q] + (b[q] - d[q]) / 4;
    a = c->phinh_i + off_i(c, ie, np1); b = c->phinh_i + off_i(c, ie, nm1); d = c->phinh_i + off_i(c, ie, n0);
    for (int q = 0; q < 16*nlev; q++) a[q] = a[q] + (b[q] - d[q]) / 4;
    orc_limiter_dp3d_k(c, c->dp3d + off_m(c, ie, np1), c->vtheta_dp + off_m(c, ie, np1), c->spheremp + (size_t)ie*16);
  }
}

void orc_prim_advance_exp(orc_ctx *c, int tstep_type, double dt, int nm1, int n0, int np1, int qsplit)
{
  const int nlev = c->nlev;
  const double eta_ave_w = 1.0 / qsplit;                                                   /* PA:137 */
  int it; double ie_;
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int ie = 0; ie < c->nelem; ie++) {                                                  /* PA:140-143 */
    const double *v = c->v + off_v(c, ie, n0) + (nlev-1)*32, *gp = c->gradphis + (size_t)ie*32;
    double *w = c->w_i + off_i(c, ie, n0) + nlev*16;
    for (int p = 0; p < NPSQ; p++) w[p] = (v[p] * gp[p] + v[16 + p] * gp[16 + p]) / g_;
  }
  if (tstep_type == 1) {                                                                   /* PA:158-165 */
    orc_caar(c, np1, n0, n0, dt/2, 0.0, 1.0, 1.0, 1.0);
    orc_caar(c, np1, n0, np1, dt, eta_ave_w, 1.0, 1.0, 1.0);
  } else if (tstep_type == 4) {                                                            /* PA:168-178 */
    orc_caar(c, np1, n0, n0, dt/4, 0.0, 1.0, 1.0, 1.0);
    orc_caar(c, np1, n0, np1, dt/6, 0.0, 1.0, 1.0, 1.0);
    orc_caar(c, np1, n0, np1, 3*dt/8, 0.0, 1.0, 1.0, 1.0);
    orc_caar(c, np1, n0, np1, dt/2, 0.0, 1.0, 1.0, 1.0);
    orc_caar(c, np1, n0, np1, dt, eta_ave_w*1.0, 1.0, 1.0, 1.0);
  } else if (tstep_type == 5) {                                                            /* PA:182-211 */
    orc_caar(c, nm1, n0, n0, dt/5, eta_ave_w/4, 1.0, 1.0, 1.0);
    orc_caar(c, np1, n0, nm1, dt/5, 0.0, 1.0, 1.0, 1.0);
    orc_caar(c, np1, n0, np1, dt/3, 0.0, 1.0, 1.0, 1.0);
    orc_caar(c, np1, n0, np1, 2*dt/3, 0.0, 1.0, 1.0, 1.0);
    combo_tt5(c, nm1, n0);
    orc_caar(c, np1, nm1, np1, 3*dt/4, 3*eta_ave_w/4, 1.0, 1.0, 1.0);
  } else if (tstep_type == 7) {                                                            /* PA:215-246 */
    const double a1 = 0.0, a2 = 1 - a1;
    double dt2 = dt/4;
    orc_caar(c, np1, n0, n0, dt2, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, a1*dt2, np1, a2*dt2, &it, &ie_);
    dt2 = dt/6;
    orc_caar(c, nm1, n0, np1, dt2, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, a1*dt2, nm1, a2*dt2, &it, &ie_);
    dt2 = 3*dt/8;
    orc_caar(c, np1, n0, nm1, dt2, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, a1*dt2, np1, a2*dt2, &it, &ie_);
    dt2 = dt/2;
    orc_caar(c, np1, n0, np1, dt2, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, a1*dt2, np1, a2*dt2, &it, &ie_);
    orc_caar(c, np1, n0, np1, dt, eta_ave_w*1.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, a1*dt, np1, a2*dt, &it, &ie_);
  } else if (tstep_type == 9) {                                                            /* PA:281-334 */
    orc_caar(c, nm1, n0, n0, dt/5, eta_ave_w/4, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, 0.0, nm1, dt/5, &it, &ie_);
    orc_caar(c, np1, n0, nm1, dt/5, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, 0.0, np1, dt/5, &it, &ie_);
    orc_caar(c, np1, n0, np1, dt/3, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, 0.0, np1, dt/3, &it, &ie_);
    orc_caar(c, np1, n0, np1, 2*dt/3, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, 0.0, np1, 2*dt/3, &it, &ie_);
    orc_caar(c, np1, nm1, np1, 3*dt/4, 3*eta_ave_w/4, 1.0, 0.0, 1.0);
    combo_tt9(c, nm1, n0, np1);
    const double a1 = 5*dt/18, a2 = dt/36, a3 = 8*dt/18;
    orc_dirk(c, nm1, a2, n0, a1, np1, a3, &it, &ie_);
  } else if (tstep_type == 10) {                                                           /* PA:336-370 */
    double dt2 = dt/4;
    orc_caar(c, nm1, n0, n0, dt2, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, 0.0, nm1, dt2, &it, &ie_);
    dt2 = dt/6;
    orc_caar(c, np1, n0, nm1, dt2, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, 0.0, np1, dt2, &it, &ie_);
    dt2 = 3*dt/8;
    orc_caar(c, np1, n0, np1, dt2, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, 0.0, np1, dt2, &it, &ie_);
    dt2 = dt/2;
    orc_caar(c, np1, n0, np1, dt2, 0.0, 1.0, 0.0, 1.0);
    orc_dirk(c, nm1, 0.0, n0, 0.0, np1, dt2, &it, &ie_);
    orc_caar(c, np1, n0, np1, dt, eta_ave_w*1.0, 1.0, 0.0, 1.0);
    const double a1 = .24362, a2 = .34184, a3 = 1 - (a1 + a2);
    orc_dirk(c, nm1, a2*dt, n0, a1*dt, np1, a3*dt, &it, &ie_);
  } else {
    fprintf(stderr, "orc_prim_advance_exp: bad tstep_type %d\n", tstep_type);
    c->error_flag = 2;
    return;
  }
  if (c->nu > 0) orc_advance_hypervis(c, np1, dt, eta_ave_w);                              /* PA:499-500 */
}

/* ---------------- vertical remap ---------------- */
/* linextrap, H/src/share/vertremap_base.F90:768-800 */
static void linextrap(double dx1, double dx2, double dx3, double dx4, double y1, double y2,
                      double *y3, double *y4, int limit, double lo, double hi)
{
  double den = (dx1 + dx2) / 2;
  double num = den + (dx2 + dx3) / 2;
  double a = num / den;
  *y3 = (1 - a) * y1 + a * y2;
  num = num + (dx3 + dx4) / 2;
  a = num / den;
  *y4 = (1 - a) * y1 + a * y2;
  if (limit) {
    *y3 = fmax(lo, fmin(hi, *y3));
    *y4 = fmax(lo, fmin(hi, *y4));
  }
}

/* remap_Q_ppm for one column, H/src/share/vertremap_base.F90:514-646 with compute_ppm_grids :650-673,
 * compute_ppm :680-723, integrate_parabola :730-737, binary_search :744-765.
 * q[f][k] at stride (fstride, kstride); dp1/dp2 at stride kstride.  remap_alg 1, 10 or 11. */
static void remap_q_ppm_column(int nlev, int nf, double *q, size_t fstride, size_t kstride,
                               const double *dp1, const double *dp2, int remap_alg, double *work)
{
  /* arrays with Fortran bounds mapped to offsets: X(1-gs:nlev+gs) -> X[i+1] for index i>=-1 */
  double *pio = work;                     /* pio(1:nlev+2)  -> pio[k]  (k 1-based, slot 0 unused) */
  double *pin = pio + (nlev + 3);         /* pin(1:nlev+1) */
  double *masso = pin + (nlev + 2);       /* masso(1:nlev+1) */
  double *ao = masso + (nlev + 2);        /* ao(-1:nlev+2) -> ao[i+1] */
  double *dpo = ao + (nlev + 4);          /* dpo(-1:nlev+2) -> dpo[i+1] */
  double *z2 = dpo + (nlev + 4);          /* z2(1:nlev) */
  double *dma = z2 + (nlev + 1);          /* dma(0:nlev+1) */
  double *ai = dma + (nlev + 2);          /* ai(0:nlev) */
  double *coefs = ai + (nlev + 1);        /* coefs(0:2,1:nlev) -> coefs[3*j+m] */
  double *ppmdx = coefs + 3 * (nlev + 1); /* ppmdx(1:10,0:nlev+1) -> ppmdx[10*j + (m-1)] */
  int *kid = (int *)(ppmdx + 10 * (nlev + 2));
#define AO(i) ao[(i) + 1]
#define DPO(i) dpo[(i) + 1]
#define DX(m, j) ppmdx[10 * (j) + (m) - 1]
  pin[1] = 0; pio[1] = 0;
  for (int k = 1; k <= nlev; k++) {
    DPO(k) = dp1[(k - 1) * kstride];
    pin[k + 1] = pin[k] + dp2[(k - 1) * kstride];
    pio[k + 1] = pio[k] + DPO(k);
  }
  pio[nlev + 2] = pio[nlev + 1] + 1.;
  pin[nlev + 1] = pio[nlev + 1];
  for (int k = 1; k <= 2; k++) { DPO(1 - k) = DPO(k); DPO(nlev + k) = DPO(nlev + 1 - k); }
  for (int k = 1; k <= nlev; k++) {
    int kk = k;
    if (pio[kk] <= pin[k + 1]) {
      while (pio[kk] <= pin[k + 1]) kk = kk + 1;
      kk = kk - 1;
    } else {                                   /* binary_search */
      int lo, hi;
      if (pio[kk] > pin[k + 1]) { lo = 1; hi = kk; } else { lo = kk; hi = nlev + 2; }
      while (hi > lo + 1) {
        kk = (lo + hi) / 2;
        if (pio[kk] > pin[k + 1]) hi = kk; else lo = kk;
      }
      kk = lo;
    }
    if (kk == nlev + 1) kk = nlev;
    kid[k] = kk;
    z2[k] = (pin[k + 1] - (pio[kk] + pio[kk + 1]) * 0.5) / DPO(kk);
  }
  for (int j = 0; j <= nlev + 1; j++) {        /* compute_ppm_grids */
    DX(1, j) = DPO(j) / (DPO(j - 1) + DPO(j) + DPO(j + 1));
    DX(2, j) = (2. * DPO(j - 1) + DPO(j)) / (DPO(j + 1) + DPO(j));
    DX(3, j) = (DPO(j) + 2. * DPO(j + 1)) / (DPO(j - 1) + DPO(j));
  }
  for (int j = 0; j <= nlev; j++) {
    DX(4, j) = DPO(j) / (DPO(j) + DPO(j + 1));
    DX(5, j) = 1. / (((DPO(j - 1) + DPO(j)) + DPO(j + 1)) + DPO(j + 2));
    DX(6, j) = (2. * DPO(j + 1) * DPO(j)) / (DPO(j) + DPO(j + 1));
    DX(7, j) = (DPO(j - 1) + DPO(j)) / (2. * DPO(j) + DPO(j + 1));
    DX(8, j) = (DPO(j + 2) + DPO(j + 1)) / (2. * DPO(j + 1) + DPO(j));
    DX(9, j) = DPO(j) * (DPO(j - 1) + DPO(j)) / (2. * DPO(j) + DPO(j + 1));
    DX(10, j) = DPO(j + 1) * (DPO(j + 1) + DPO(j + 2)) / (DPO(j) + 2. * DPO(j + 1));
  }
  for (int f = 0; f < nf; f++) {
    double *Q = q + f * fstride;
    masso[1] = 0.;
    for (int k = 1; k <= nlev; k++) {
      AO(k) = Q[(k - 1) * kstride];
      masso[k + 1] = masso[k] + AO(k);
      AO(k) = AO(k) / DPO(k);
    }
    if (remap_alg == 10 || remap_alg == 11) {
      const int limit = (remap_alg == 10);
      double lo = 0, hi = 0;
      if (limit) {
        lo = hi = AO(1);
        for (int k = 2; k <= nlev; k++) { lo = fmin(lo, AO(k)); hi = fmax(hi, AO(k)); }
      }
      linextrap(DPO(2), DPO(1), DPO(0), DPO(-1), AO(2), AO(1), &AO(0), &AO(-1), limit, lo, hi);
      linextrap(DPO(nlev - 1), DPO(nlev), DPO(nlev + 1), DPO(nlev + 2), AO(nlev - 1), AO(nlev),
                &AO(nlev + 1), &AO(nlev + 2), limit, lo, hi);
    } else {
      for (int k = 1; k <= 2; k++) { AO(1 - k) = AO(1); AO(nlev + k) = AO(nlev); }
    }
    for (int j = 0; j <= nlev + 1; j++) {      /* compute_ppm stage 1 */
      double da = DX(1, j) * (DX(2, j) * (AO(j + 1) - AO(j)) + DX(3, j) * (AO(j) - AO(j - 1)));
      double m = fmin(fabs(da), fmin(2. * fabs(AO(j) - AO(j - 1)), 2. * fabs(AO(j + 1) - AO(j))));
      dma[j] = copysign(m, da);
      if ((AO(j + 1) - AO(j)) * (AO(j) - AO(j - 1)) <= 0.) dma[j] = 0.;
    }
    for (int j = 0; j <= nlev; j++)            /* stage 2 */
      ai[j] = AO(j) + DX(4, j) * (AO(j + 1) - AO(j)) + DX(5, j) * (DX(6, j) * (DX(7, j) - DX(8, j))
              * (AO(j + 1) - AO(j)) - DX(9, j) * dma[j + 1] + DX(10, j) * dma[j]);
    for (int j = 1; j <= nlev; j++) {          /* stage 3 */
      double al = ai[j - 1], ar = ai[j];
      if ((ar - AO(j)) * (AO(j) - al) <= 0.) { al = AO(j); ar = AO(j); }
      if ((ar - al) * (AO(j) - (al + ar) / 2.) > (ar - al) * (ar - al) / 6.) al = 3. * AO(j) - 2. * ar;
      if ((ar - al) * (AO(j) - (al + ar) / 2.) < -((ar - al) * (ar - al)) / 6.) ar = 3. * AO(j) - 2. * al;
      coefs[3 * j + 0] = 1.5 * AO(j) - (al + ar) / 4.;
      coefs[3 * j + 1] = ar - al;
      coefs[3 * j + 2] = 3. * (-2. * AO(j) + (al + ar));
    }
    double massn1 = 0.;
    for (int k = 1; k <= nlev; k++) {
      const int kk = kid[k];
      const double x1 = -0.5, x2 = z2[k];
      const double *a = coefs + 3 * kk;
      const double par = a[0] * (x2 - x1) + a[1] * (x2 * x2 - x1 * x1) / 0.2e1 + a[2] * (x2 * x2 * x2 - x1 * x1 * x1) / 0.3e1;
      const double massn2 = masso[kk] + par * DPO(kk);
      Q[(k - 1) * kstride] = massn2 - massn1;
      massn1 = massn2;
    }
  }
#undef AO
#undef DPO
#undef DX
}

/* vertical_remap, H/src/theta-l/share/vertremap_mod.F90:23-171, rsplit>0 branch, qsize = 0.
 * Returns 1 (and sets error_flag) on a negative layer thickness (:87-103). */
int orc_vertical_remap(orc_ctx *c, int np1, int remap_alg)
{
  const int nlev = c->nlev, nlevp = nlev + 1;
  int bad = 0;
#pragma omp parallel for num_threads(c->nthreads) schedule(static) reduction(|:bad)
  for (int ie = 0; ie < c->nelem; ie++) {
    double *v = c->v + off_v(c, ie, np1), *w = c->w_i + off_i(c, ie, np1);
    double *th = c->vtheta_dp + off_m(c, ie, np1), *phi = c->phinh_i + off_i(c, ie, np1);
    double *dp3d = c->dp3d + off_m(c, ie, np1);
    double *dp = (double *)malloc(sizeof(double) * 16 * (7 * nlev + nlevp));
    double *ttmp = dp + 16 * nlev, *phi_ref = ttmp + 5 * 16 * nlev;      /* ttmp[f][k][p] */
    double *work = (double *)malloc(sizeof(double) * (24 * (nlev + 4) + 64));
    double ps[NPSQ];
    for (int p = 0; p < NPSQ; p++) {
      double s = 0;                                                      /* sum(dp3d,3) */
      for (int k = 0; k < nlev; k++) s = s + dp3d[k * 16 + p];
      ps[p] = c->hyai[0] * c->ps0 + s;
    }
    for (int k = 0; k < nlev; k++)
      for (int p = 0; p < NPSQ; p++) {
        dp[k * 16 + p] = (c->hyai[k + 1] - c->hyai[k]) * c->ps0 + (c->hybi[k + 1] - c->hybi[k]) * ps[p];
        if (dp3d[k * 16 + p] < 0) bad |= 1;
      }
    orc_phi_from_eos(c, c->phis + (size_t)ie * 16, th, dp3d, phi_ref);
    for (int i = 0; i < 16 * nlevp; i++) phi[i] = phi[i] - phi_ref[i];
    for (int k = 0; k < nlev; k++)
      for (int p = 0; p < NPSQ; p++) {
        const double ds = dp3d[k * 16 + p];
        ttmp[(0 * nlev + k) * 16 + p] = v[k * 32 + p] * ds;
        ttmp[(1 * nlev + k) * 16 + p] = v[k * 32 + 16 + p] * ds;
        ttmp[(2 * nlev + k) * 16 + p] = th[k * 16 + p];
        ttmp[(3 * nlev + k) * 16 + p] = phi[(k + 1) * 16 + p] - phi[k * 16 + p];
        ttmp[(4 * nlev + k) * 16 + p] = w[(k + 1) * 16 + p] - w[k * 16 + p];
      }
    for (int p = 0; p < NPSQ; p++)
      remap_q_ppm_column(nlev, 5, ttmp + p, (size_t)16 * nlev, 16, dp3d + p, dp + p, remap_alg, work);
    for (int k = 0; k < nlev; k++)
      for (int p = 0; p < NPSQ; p++) {
        v[k * 32 + p] = ttmp[(0 * nlev + k) * 16 + p] / dp[k * 16 + p];
        v[k * 32 + 16 + p] = ttmp[(1 * nlev + k) * 16 + p] / dp[k * 16 + p];
        th[k * 16 + p] = ttmp[(2 * nlev + k) * 16 + p];
      }
    for (int k = nlev - 1; k >= 0; k--)
      for (int p = 0; p < NPSQ; p++) {
        phi[k * 16 + p] = phi[(k + 1) * 16 + p] - ttmp[(3 * nlev + k) * 16 + p];
        w[k * 16 + p] = w[(k + 1) * 16 + p] - ttmp[(4 * nlev + k) * 16 + p];
      }
    orc_phi_from_eos(c, c->phis + (size_t)ie * 16, th, dp, phi_ref);
    for (int i = 0; i < 16 * nlevp; i++) phi[i] = phi[i] + phi_ref[i];
    const double *gp = c->gradphis + (size_t)ie * 32;
    for (int p = 0; p < NPSQ; p++)
      w[nlev * 16 + p] = (v[(nlev - 1) * 32 + p] * gp[p] + v[(nlev - 1) * 32 + 16 + p] * gp[16 + p]) / g_;
    for (int i = 0; i < 16 * nlev; i++) dp3d[i] = dp[i];
    free(work); free(dp);
  }
  if (bad) c->error_flag |= 4;
  return bad;
}
