// Shared host/device definitions for libthetal_b200.so (sm_100a only, FP64).
// Device data layout (HBM):
//   every prognostic field of one time level is its own allocation, element-major, Fortran order
//   inside the element:  v[ie][k][c][16]  w_i,phinh_i[ie][k<=nlev][16]  vtheta_dp,dp3d[ie][k][16]
//   (16 = np*np points, i fastest).  Time levels are pointer slots, never copied.
//   One extra "tens" state (pre-DSS, mass weighted) per exchange: a pull-style DSS reads the
//   neighbours' tens values and writes the finished state, so no edge buffers exist.
#pragma once
#include <cstddef>
#include <cstdint>

namespace thetal {

constexpr int NP = 4;
constexpr int NPSQ = 16;
constexpr int NTL = 3;

// physical_constants.F90:55-73,86 ; element_ops.F90:80
constexpr double G_GRAV = 9.80616;
constexpr double RGAS = 287.04;
constexpr double CP = 1005.0;
constexpr double P0 = 100000.0;
constexpr double TREF = 288.0;
constexpr double KAPPA = RGAS / CP;
constexpr double TREF_LAPSE = 0.0065;

struct StatePtrs {          // one time level (or one tens buffer)
  double *v, *w, *th, *phi, *dp;
};

struct Geo {                // per-element geometry, Fortran memory per element
  const double *D, *Dinv;             // [ie][b][a][16]
  const double *metdet, *rmetdet, *spheremp, *rspheremp, *fcor, *phis;  // [ie][16]
  const double *metinv, *tensorvisc;  // [ie][b][a][16]
  const double *vec_sph2cart;         // [ie][2][3][16]
  const double *gradphis;             // [ie][2][16]
  const int *dss_src;                 // [ie][16][3]: neighbour source (ie2*16+p2), -1 none; Fortran unpack order
};

struct Params {
  int nelem, nlev;
  int rsplit, qsplit, tstep_type, theta_advect_form, pgrad_correction;
  int hypervis_subcycle, hypervis_subcycle_tom, nlev_tom, hypervis_scaling_on;
  double nu, nu_div, nu_p, nu_s, nu_top, dp3d_thresh, vtheta_thresh;
  double rearth_inv;        // scale_factor_inv == laplacian_rigid_factor
  double ps0, ptop;         // ptop = hyai(1)*ps0
};

struct Derived {            // element_state.F90:51-84 (subset accumulated by the step)
  double *vn0, *eta_dot_dpdn, *omega_p;    // [ie][k][2][16], [ie][nlevp][16], [ie][k][16]
};

struct RefStates {
  const double *theta_ref, *dp_ref, *phi_ref;   // [ie][k][16], phi_ref nlevp
};

// error flag bits written by kernels
enum : int { ERR_EOS_BAD_STATE = 1, ERR_DIRK_NOT_CONVERGED = 2, ERR_REMAP_NEGATIVE_LAYER = 4 };

}  // namespace thetal
