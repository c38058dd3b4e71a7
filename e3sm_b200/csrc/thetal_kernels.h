// Kernel argument blocks and launchers of libthetal_b200.so (kernels in thetal_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include "thetal_ctx.h"

namespace thetal {

struct CaarArgs {
  const int *elist = nullptr;   // optional element list (partition-boundary / interior pass); null = all
  int ecount = 0;
  StatePtrs n0, nm1, tens;
  Geo geo;
  Derived der;
  Params prm;
  double dt2, eta_ave_w, scale1, scale2, scale3;
  int *err;
  // optional: the explicit stiff terms of state(n0) that a later DIRK stage adds (IM:176-205), saved here so that
  // stage does not have to redo the EOS of this state: ew = dpnh_dp_i - 1, ephi = g w_i - gwphis, [ie][k<nlev][16]
  double *ew = nullptr, *ephi = nullptr;
};

struct DssArgs {
  StatePtrs tens, out;
  Geo geo;
  Params prm;
  double dt2, scale1, scale3;
  int do_limiter;
  const int *elist;         // optional element list; null = all
  int ecount;
  StatePtrs combo_b{}, combo_c{};   // tstep_type 9 closing stage: out += (combo_b - combo_c)/4 fused behind the DSS (v == null: off)
};

struct DirkArgs {
  StatePtrs np1, n0, nm1;
  Geo geo;
  Params prm;
  double dt2, alphadt_n0, alphadt_nm1;
  // optional: explicit terms of state(n0) / state(nm1) saved by the CAAR launches that evaluated those states
  const double *ew_n0 = nullptr, *ephi_n0 = nullptr, *ew_nm1 = nullptr, *ephi_nm1 = nullptr;
  int *work = nullptr, *work_other = nullptr;   // element counter of this launch / of the previous one (reset here)
  // optional: the DSS of the CAAR stage that precedes this solve, done by the DIRK kernel while it stages an element
  // (tens.v != null): np1 = rspheremp * (node sum of tens), surface fix-up, limiter_dp3d_k, as caar_dss_post_kernel
  StatePtrs tens{};
  double dss_dt2 = 0.0, dss_scale1 = 0.0;
  int dss_limiter = 0;
  int *err;
  int *max_iter;            // running max of itercount over elements (atomicMax)
  unsigned long long *max_deltaerr_bits;   // running max of deltaerr (as ordered bits)
};

struct HvArgs {
  StatePtrs st;             // state at time level nt (in/out)
  StatePtrs tin, tout;      // tens buffers (gather source / destination)
  Geo geo;
  RefStates ref;
  Params prm;
  double dt, nu_ratio;
  int th_is_theta;          // st.th currently holds theta (not vtheta_dp)
  int last_subcycle;        // convert back to vtheta_dp and reset w b.c. (PA:767-774)
  const int *elist;         // optional element list (partition-boundary / interior pass); null = all
  int ecount;
};

struct ComboArgs {
  StatePtrs a, b, c;     // tt9: a(np1) += (b(nm1) - c(n0))/4 ; tt5: a(nm1) = (5*a - b(n0))/4
  Geo geo;
  Params prm;
  int kind;              // 9 or 5
};

struct RemapArgs {
  StatePtrs st;             // time level np1 (in/out)
  Geo geo;
  Params prm;
  int remap_alg;            // vert_remap_u_alg: 1 (mirrored PPM), 10, 11 (linear extrapolation, 10 = limited)
  int *err;
};

constexpr int NDIAG = 4;    // ps, KE, PE, IE
constexpr int NEXTREMA = 7; // u, v, w_i, theta, T, dz, dp3d
struct DiagArgs {
  StatePtrs st;
  Geo geo;
  Params prm;
  double *part;                    // [nelem][NDIAG] element integrals J(ie)
  unsigned long long *maxbits;     // [NDIAG] running max |J| as bits
  unsigned long long *ps_minmax;   // [2] min / max surface pressure as bits
};

void upload_constants(const double *dvv, const double *mp, const double *hybi, const double *dp0,
                      const double *nu_scale_top, int nlev);
void launch_caar_pre(const CaarArgs &A, cudaStream_t st);
void launch_caar_dss_post(const DssArgs &A, cudaStream_t st);
void launch_dirk(const DirkArgs &A, cudaStream_t st);
void launch_hv(int which, const HvArgs &A, cudaStream_t st);   // 0 laplace1, 1 laplace2, 2 update, 3 sponge laplace, 4 sponge update
void launch_combo(const ComboArgs &A, cudaStream_t st);
void launch_halo_pack(const StatePtrs &t, const int *idx, int nent, int nk, int wbot, int nlev, double *buf,
                      cudaStream_t st);
void launch_halo_unpack(const StatePtrs &t, const int *idx, int nent, int nk, int wbot, int nlev,
                        const double *buf, cudaStream_t st);
void launch_diag_elements(const DiagArgs &A, cudaStream_t st);
void launch_extrema(const DiagArgs &A, unsigned long long *keys, cudaStream_t st);   // keys[2*f] = min, [2*f+1] = max
void launch_diag_accumulate(const double *part, int nelem, const unsigned long long *maxbits,
                            unsigned long long *limbs, cudaStream_t st);
void upload_hyai(const double *hyai, int nlev);
void launch_vertical_remap(const RemapArgs &A, cudaStream_t st);
void launch_forcing(const StatePtrs &s, const double *fm, const double *fvtheta, const double *fphi, const Geo &geo,
                    double dt, int nelem, int nlev, cudaStream_t st);
void launch_ps_from_dp(const double *dp, double *ps, int nelem, int nlev, double ptop, cudaStream_t st);
void launch_dp_from_ps(double *dp, const double *ps, int nelem, int nlev, double ps0, cudaStream_t st);
void launch_dss_generic(const double *f, double *out, const int *dss_src, int nelem, int nl, cudaStream_t st);
void launch_halo_generic(double *f, const int *idx, int nent, int nl, double *buf, int unpack, cudaStream_t st);
double launch_fp64_peak(cudaStream_t st);   // returns the flops the launch performs
void launch_wbc(const StatePtrs &s, const Geo &geo, int nelem, int nlev, cudaStream_t st);

}  // namespace thetal
