// Hand-written sm_100a FP64 kernels for HOMME's theta-l nonhydrostatic dycore step.
// Reference arithmetic (H/ = components/homme/ in the reference tree):
//   PA  = H/src/theta-l/share/prim_advance_mod.F90   (compute_andor_apply_rhs :1023-1819,
//                                                      advance_hypervis :546-862, limiter :1822-1931)
//   DM  = H/src/share/derivative_mod.F90              (sphere operators)
//   IM  = H/src/theta-l/share/imex_mod.F90            (compute_stage_value_dirk :60-364)
//   EOS = H/src/theta-l/share/eos.F90
//   VT  = H/src/theta-l/share/viscosity_theta.F90
// Design: one CTA per element; a half-warp owns the 16 GLL points of one level, so every 4x4
// derivative contraction is 8 shared-memory reads + 8 FMAs held in registers (far too small for
// tensor cores).  Element columns are staged through shared memory once; DSS is a pull-style
// gather from the neighbours' pre-DSS ("tens") values, so there are no edge buffers.
#include "thetal_kernels.h"
#include "thetal_pow.h"
#include "thetal_reprosum.h"
#include <cstdio>
#include <cstdlib>

namespace thetal {

constexpr int MAXLEV = 128;
__constant__ double c_dvv[16];            // c_dvv[l*4+i] = Dvv(i,l)
__constant__ double c_mp[16];
__constant__ double c_hybi[MAXLEV + 1];
__constant__ double c_dp0[MAXLEV];
__constant__ double c_nu_scale_top[MAXLEV];
__constant__ double c_hyai[MAXLEV + 1];

void upload_constants(const double *dvv, const double *mp, const double *hybi, const double *dp0,
                      const double *nu_scale_top, int nlev)
{
  if (dvv) cudaMemcpyToSymbol(c_dvv, dvv, 16 * sizeof(double));
  if (mp) cudaMemcpyToSymbol(c_mp, mp, 16 * sizeof(double));
  if (hybi) cudaMemcpyToSymbol(c_hybi, hybi, (nlev + 1) * sizeof(double));
  if (dp0) cudaMemcpyToSymbol(c_dp0, dp0, nlev * sizeof(double));
  if (nu_scale_top) cudaMemcpyToSymbol(c_nu_scale_top, nu_scale_top, nlev * sizeof(double));
}
void upload_hyai(const double *hyai, int nlev) { cudaMemcpyToSymbol(c_hyai, hyai, (nlev + 1) * sizeof(double)); }

// ------------------------------------------------------------------------------------------
// Per-thread view of one GLL point (a,b) = (i,j) of an element: derivative rows + metric terms.
// ------------------------------------------------------------------------------------------
struct Pt {
  int p, a, b;
  double da[4], db[4];          // Dvv(i,a), Dvv(i,b), times sfi = 1/rearth: ddx, ddy return derivatives per metre
  double Dinv[4], D[4];         // index b2*2+a2 of X(i,j,a2,b2)
  double metdet, rmetdet, sfi;
};

__device__ __forceinline__ void pt_init(Pt &t, const Geo &g, int ie, int p, double sfi)
{
  t.p = p; t.a = p & 3; t.b = p >> 2;
#pragma unroll
  for (int i = 0; i < 4; i++) { t.da[i] = c_dvv[t.a * 4 + i] * sfi; t.db[i] = c_dvv[t.b * 4 + i] * sfi; }
#pragma unroll
  for (int q = 0; q < 4; q++) {
    t.Dinv[q] = g.Dinv[(size_t)ie * 64 + q * 16 + p];
    t.D[q] = g.D[(size_t)ie * 64 + q * 16 + p];
  }
  t.metdet = g.metdet[(size_t)ie * 16 + p];
  t.rmetdet = g.rmetdet[(size_t)ie * 16 + p];
  t.sfi = sfi;
}

// x-derivative  v1(a,b) = sum_i Dvv(i,a) s(i,b)  and y-derivative v2(a,b) = sum_i Dvv(i,b) s(a,i)
// of a 16-point tile s[] (i fastest) in shared memory.
__device__ __forceinline__ double ddx(const Pt &t, const double *s)
{
  const double *r = s + t.b * 4;
  double acc = t.da[0] * r[0];
  acc = fma(t.da[1], r[1], acc); acc = fma(t.da[2], r[2], acc); acc = fma(t.da[3], r[3], acc);
  return acc;
}
__device__ __forceinline__ double ddy(const Pt &t, const double *s)
{
  const double *c = s + t.a;
  double acc = t.db[0] * c[0];
  acc = fma(t.db[1], c[4], acc); acc = fma(t.db[2], c[8], acc); acc = fma(t.db[3], c[12], acc);
  return acc;
}
// DM:410-450
__device__ __forceinline__ void gradient_sphere(const Pt &t, const double *s, double &g1, double &g2)
{
  const double v1 = ddx(t, s), v2 = ddy(t, s);
  g1 = t.Dinv[0] * v1 + t.Dinv[1] * v2;
  g2 = t.Dinv[2] * v1 + t.Dinv[3] * v2;
}
// DM:1021-1068, first half: contravariant, metdet-weighted components to be written to a tile
__device__ __forceinline__ void div_prep(const Pt &t, double v1, double v2, double &gv1, double &gv2)
{
  gv1 = t.metdet * (t.Dinv[0] * v1 + t.Dinv[2] * v2);
  gv2 = t.metdet * (t.Dinv[1] * v1 + t.Dinv[3] * v2);
}
__device__ __forceinline__ double div_finish(const Pt &t, const double *sgv1, const double *sgv2)
{
  return (ddx(t, sgv1) + ddy(t, sgv2)) * t.rmetdet;
}
// DM:915-966
__device__ __forceinline__ void vort_prep(const Pt &t, double v1, double v2, double &vco1, double &vco2)
{
  vco1 = t.D[0] * v1 + t.D[1] * v2;
  vco2 = t.D[2] * v1 + t.D[3] * v2;
}
__device__ __forceinline__ double vort_finish(const Pt &t, const double *svco1, const double *svco2)
{
  return (ddx(t, svco2) - ddy(t, svco1)) * t.rmetdet;
}

// Ask the memory system to pull `bytes` (multiple of 16, 16-byte aligned) into L2 ahead of use; costs
// no registers, unlike an early load.
__device__ __forceinline__ void prefetch_l2(const void *ptr, unsigned bytes)
{
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr), "r"(bytes) : "memory");
}

// TMA bulk copy global -> shared (cp.async.bulk, 1-D) completing on an mbarrier: one thread moves a
// whole contiguous element column block; bytes multiple of 16, both addresses 16-byte aligned.
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned phase)
{
  unsigned ok;
  do {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(phase) : "memory");
  } while (!ok);
}

// EOS:175-183  pnh = p0*(p_over_exner/p0)^(1/(1-kappa))
// 1/(1-kappa) and kappa-1 with g*log2(e) as double-doubles (thetal_pow.h)
__device__ constexpr thetal_pow::PowConst C_GAMMA = {1.0 / (1.0 - KAPPA), 0.0, 2.0194836983862725, 1.3732662420330506e-16};
__device__ constexpr thetal_pow::PowConst C_KM1 = {KAPPA - 1.0, 0.0, -1.0306441110016318, -8.6951363107339526e-17};
static_assert(1.0 / (1.0 - KAPPA) == 1.3997994317232159 && KAPPA - 1.0 == -0.71438805970149244, "pow constants");
using thetal_pow::div_fast;
using thetal_pow::rcp_fast;

// a/d given r = rcp_fast(d): the same two operations div_fast ends with, so a/d == div_fast(a, d) bit for bit
__device__ __forceinline__ double div_by(double a, double d, double r)
{
  const double q = a * r;
  return fma(fma(-q, d, a), r, q);
}
// x / 1e5 to <= 1 ulp without a reciprocal instruction (P0 is a constant)
__device__ __forceinline__ double div_p0(double x)
{
  constexpr double c = 1.0 / P0;
  const double q = x * c;
  return fma(fma(-q, P0, x), c, q);
}
__device__ __forceinline__ double eos_pnh(double p_over_exner)
{
  return P0 * thetal_pow::pow_pos(div_p0(p_over_exner), C_GAMMA);
}

// ==========================================================================================
// compute_andor_apply_rhs, element-local part before the exchange (PA:1113-1702), NH, rsplit>0.
// Writes the mass-weighted, not yet DSS'd new state into `tens`.
// ==========================================================================================

constexpr int CAAR_THREADS = 256;
constexpr int CAAR_SLOTS = 10;    // scratch tiles per half-warp

__host__ __device__ inline size_t caar_smem_bytes(int nlev, bool vadv)
{
  const size_t nlevp = nlev + 1;
  // v(2) dp th pnh exn div : 7*nlev tiles ; w phi : 2*nlevp tiles ; scratch ; rsplit = 0: eta_dot_dpdn at interfaces
  return sizeof(double) * 16 * (7 * (size_t)nlev + 2 * nlevp + (CAAR_THREADS / 16) * CAAR_SLOTS + (vadv ? nlevp : 0));
}

// FORM = theta_advect_form, PGRAD = pgrad_correction, ACCUM = (eta_ave_w != 0), VADV = (rsplit == 0, the
// vertically Eulerian terms of PA:1227-1303 and prim_si_mod.F90:94-148): compile-time so that the level
// loop is one straight-line block the scheduler can interleave freely
// SAVE = also store the explicit stiff terms of state(n0) for the closing DIRK stage (IM:176-205)
template <int FORM, bool PGRAD, bool ACCUM, bool VADV, bool SAVE>
__global__ void __launch_bounds__(CAAR_THREADS, 2) caar_pre_kernel(const CaarArgs A)
{
  extern __shared__ double smem[];
  const int nlev = A.prm.nlev, nlevp = nlev + 1;
  const int ie = A.elist ? A.elist[blockIdx.x] : (int)blockIdx.x, tid = threadIdx.x;
  double *sv = smem;                       // [k][2][16]
  double *sdp = sv + 32 * nlev;            // [k][16]
  double *sth = sdp + 16 * nlev;
  double *sw = sth + 16 * nlev;            // [nlevp][16]
  double *sphi = sw + 16 * nlevp;
  double *spnh = sphi + 16 * nlevp;        // [k][16]
  double *sexn = spnh + 16 * nlev;
  double *sdiv = sexn + 16 * nlev;
  double *scr = sdiv + 16 * nlev + (tid >> 4) * (CAAR_SLOTS * 16);   // this half-warp's tiles
  double *seta = sdiv + 16 * nlev + (CAAR_THREADS / 16) * (CAAR_SLOTS * 16);   // [nlevp][16], VADV only

  // the nm1 columns are only needed by the final update of each level (PA:1674-1690), and the
  // derived accumulators by their read-modify-write: start them towards L2 now
  if (tid < 7) {
    const unsigned bv = 32u * nlev * 8u, bm = 16u * nlev * 8u, bi = 16u * nlevp * 8u;
    const bool other = A.nm1.v != A.n0.v;
    if (tid == 0 && other) prefetch_l2(A.nm1.v + (size_t)ie * 32 * nlev, bv);
    if (tid == 1 && other) prefetch_l2(A.nm1.dp + (size_t)ie * 16 * nlev, bm);
    if (tid == 2 && other) prefetch_l2(A.nm1.th + (size_t)ie * 16 * nlev, bm);
    if (tid == 3 && other) prefetch_l2(A.nm1.w + (size_t)ie * 16 * nlevp, bi);
    if (tid == 4 && other) prefetch_l2(A.nm1.phi + (size_t)ie * 16 * nlevp, bi);
    if (tid == 5 && A.eta_ave_w != 0.0) prefetch_l2(A.der.vn0 + (size_t)ie * 32 * nlev, bv);
    if (tid == 6 && A.eta_ave_w != 0.0) prefetch_l2(A.der.omega_p + (size_t)ie * 16 * nlev, bm);
  }
  // ---- phase 0: stage the n0 columns of this element: five TMA bulk copies (each field of an
  // element is one contiguous block), completion counted in bytes on an mbarrier ----
  // two barriers: the EOS phase needs vtheta_dp, dp3d and phinh_i only and starts while v and w_i still travel
  __shared__ unsigned long long stage_bar[2];
  if (tid == 0) { mbar_init(&stage_bar[0], 1); mbar_init(&stage_bar[1], 1); }
  __syncthreads();
  if (tid == 0) {
    const unsigned bv = 32u * nlev * 8u, bm = 16u * nlev * 8u, bi = 16u * nlevp * 8u;
    mbar_expect_tx(&stage_bar[0], 2 * bm + bi);
    bulk_g2s(sth, A.n0.th + (size_t)ie * 16 * nlev, bm, &stage_bar[0]);
    bulk_g2s(sphi, A.n0.phi + (size_t)ie * 16 * nlevp, bi, &stage_bar[0]);
    bulk_g2s(sdp, A.n0.dp + (size_t)ie * 16 * nlev, bm, &stage_bar[0]);
    mbar_expect_tx(&stage_bar[1], bv + bi);
    bulk_g2s(sv, A.n0.v + (size_t)ie * 32 * nlev, bv, &stage_bar[1]);
    bulk_g2s(sw, A.n0.w + (size_t)ie * 16 * nlevp, bi, &stage_bar[1]);
  }
  // the per-point geometry of phase 2 is requested while the bulk copies are in flight (its latency would otherwise
  // be paid after the EOS phase, with nothing to overlap it)
  const int p = tid & 15, kslot = tid >> 4;
  Pt t;
  pt_init(t, A.geo, ie, p, A.prm.rearth_inv);
  const double fcor = A.geo.fcor[(size_t)ie * 16 + p], spheremp = A.geo.spheremp[(size_t)ie * 16 + p];
  const double gps1 = A.geo.gradphis[(size_t)ie * 32 + p], gps2 = A.geo.gradphis[(size_t)ie * 32 + 16 + p];
  mbar_wait(&stage_bar[0], 0);
  // ---- phase 1: EOS (EOS:175-183), one pow per point-level ----
  {
    int bad = 0;
    for (int q = tid; q < 16 * nlev; q += CAAR_THREADS) {
      const double dphi = sphi[q + 16] - sphi[q];
      const double th = sth[q], dp = sdp[q];
      if (th < 0 || dp < 0 || dphi > 0) bad = 1;
      const double poe = div_fast(RGAS * th, -dphi);
      const double pnh = eos_pnh(poe);
      spnh[q] = pnh;
      sexn[q] = div_fast(pnh, poe);
    }
    if (bad) atomicOr(A.err, ERR_EOS_BAD_STATE);
  }
  mbar_wait(&stage_bar[1], 0);
  __syncthreads();

  // ---- phase 2: tendencies, a half-warp per level ----
  constexpr int KS = CAAR_THREADS / 16;
  if (VADV) {
    // eta_dot_dpdn at the interfaces (PA:1227-1251) needs the divergence of every level first
    for (int k = kslot; k < ((nlev + KS - 1) / KS) * KS; k += KS) {
      const int kk = k < nlev ? k : nlev - 1, q = kk * 16 + p;
      double gv1, gv2;
      div_prep(t, sv[kk * 32 + p] * sdp[q], sv[kk * 32 + 16 + p] * sdp[q], gv1, gv2);
      scr[p] = gv1; scr[16 + p] = gv2;
      __syncwarp();
      const double dv = div_finish(t, scr, scr + 16);
      if (k < nlev) sdiv[q] = dv;
      __syncwarp();
    }
    __syncthreads();
    if (tid < 16) {
      double sum = 0.0;
      for (int k = 0; k < nlev; k++) { sum = sum + sdiv[k * 16 + tid]; seta[(k + 1) * 16 + tid] = sum; }
      for (int k = 0; k < nlev - 1; k++) seta[(k + 1) * 16 + tid] = c_hybi[k + 1] * sum - seta[(k + 1) * 16 + tid];
      seta[tid] = 0.0;
      seta[nlev * 16 + tid] = 0.0;
    }
    __syncthreads();
  }
  const double s1 = A.scale1, s2 = A.scale2, s3 = A.scale3, dt2 = A.dt2;
  constexpr bool accum = ACCUM;
  const int niter = (nlev + KS - 1) / KS;

  for (int it = 0; it < niter; it++) {
    const int kk = it * KS + kslot;
    const bool act = kk < nlev;
    const int k = act ? kk : nlev - 1;     // inactive half-warps redo the last level, never store
    const int q = k * 16 + p;
    // nm1 values of the final update (PA:1674-1690): issue the loads now, use them ~1000 instructions later
    const size_t gq = (size_t)ie * 16 * nlev + q, gqv = (size_t)ie * 32 * nlev + k * 32 + p;
    const size_t gqi = (size_t)ie * 16 * nlevp + q;
    const double m_v1 = A.nm1.v[gqv], m_v2 = A.nm1.v[gqv + 16], m_th = A.nm1.th[gq];
    const double m_w = A.nm1.w[gqi], m_phi = A.nm1.phi[gqi], m_dp = A.nm1.dp[gq];
    double o_vn1 = 0.0, o_vn2 = 0.0;         // derived%vn0 before this stage: requested with the nm1 values
    if (ACCUM) { o_vn1 = A.der.vn0[gqv]; o_vn2 = A.der.vn0[gqv + 16]; }
    const double dp = sdp[q], th = sth[q], v1 = sv[k * 32 + p], v2 = sv[k * 32 + 16 + p];
    const double wk = sw[q], wk1 = sw[q + 16];
    const double pnh = spnh[q], exner = sexn[q];
    const double theta = div_fast(th, dp);                                         // PA:1116
    // interface quantities at k (top of the layer) and k+1 (bottom)
    double dpi_k, v1i_k, v2i_k, mu_k;                                               // PA:1156-1170, EOS:198-208
    if (k == 0) {
      dpi_k = dp; v1i_k = v1; v2i_k = v2;
      mu_k = div_fast(2 * (pnh - A.prm.ptop), dpi_k);
    } else {
      const double dpm = sdp[q - 16];
      dpi_k = (dp + dpm) / 2;
      const double r2 = thetal_pow::rcp_fast(2 * dpi_k);          // one reciprocal for the three quotients:
      v1i_k = div_by(dp * v1 + dpm * sv[(k - 1) * 32 + p], 2 * dpi_k, r2);
      v2i_k = div_by(dp * v2 + dpm * sv[(k - 1) * 32 + 16 + p], 2 * dpi_k, r2);
      mu_k = div_by(pnh - spnh[q - 16], dpi_k, 2 * r2);           // 1/dpi = 2/(2 dpi), exact scaling
    }
    double mu_k1;
    if (k == nlev - 1) mu_k1 = div_fast(2 * ((pnh + dp / 2) - pnh), dp);             // EOS:193,205
    else mu_k1 = div_fast(spnh[q + 16] - pnh, (sdp[q + 16] + dp) / 2);

    // tiles for the operators that act on computed quantities
    double gv1, gv2, vc1, vc2, gt1, gt2;
    div_prep(t, v1 * dp, v2 * dp, gv1, gv2);                                        // PA:1188-1192
    vort_prep(t, v1, v2, vc1, vc2);
    scr[0 * 16 + p] = gv1; scr[1 * 16 + p] = gv2; scr[2 * 16 + p] = vc1; scr[3 * 16 + p] = vc2;
    if (FORM != 1) {
      div_prep(t, v1 * th, v2 * th, gt1, gt2);                                      // PA:1373-1375
      scr[4 * 16 + p] = gt1; scr[5 * 16 + p] = gt2;
    }
    scr[6 * 16 + p] = (wk * wk + wk1 * wk1) / 4;                                    // PA:1401
    scr[7 * 16 + p] = (v1 * v1 + v2 * v2) / 2;                                      // PA:1409
    if (FORM != 0) { scr[8 * 16 + p] = theta; scr[9 * 16 + p] = theta * exner; }
    __syncwarp();

    const double divdp = div_finish(t, scr, scr + 16);
    const double vort = vort_finish(t, scr + 32, scr + 48);
    double gphi1_k, gphi2_k, gphi1_k1, gphi2_k1, gw1_k, gw2_k, gw1_k1, gw2_k1;
    gradient_sphere(t, sphi + k * 16, gphi1_k, gphi2_k);                            // PA:1323
    gradient_sphere(t, sphi + (k + 1) * 16, gphi1_k1, gphi2_k1);
    gradient_sphere(t, sw + k * 16, gw1_k, gw2_k);                                  // PA:1325
    gradient_sphere(t, sw + (k + 1) * 16, gw1_k1, gw2_k1);
    double wv1, wv2, ke1, ke2, ge1, ge2;
    gradient_sphere(t, scr + 6 * 16, wv1, wv2);                                     // PA:1403
    gradient_sphere(t, scr + 7 * 16, ke1, ke2);                                     // PA:1410
    gradient_sphere(t, sexn + k * 16, ge1, ge2);                                    // PA:1411

    // theta tendency PA:1372-1398
    double div_v_theta;
    if (FORM == 0) {
      div_v_theta = div_finish(t, scr + 64, scr + 80);
    } else {
      double gth1, gth2;
      gradient_sphere(t, scr + 8 * 16, gth1, gth2);
      const double nc = theta * divdp + dp * v1 * gth1 + dp * v2 * gth2;
      if (FORM == 1) div_v_theta = nc;
      else div_v_theta = div_finish(t, scr + 64, scr + 80) / 2 + nc / 2;
    }
    // vertically Eulerian terms (PA:1252-1302); all zero for rsplit > 0 (PA:1219-1225)
    double theta_vadv = 0.0, w_vadv_k = 0.0, phi_vadv_k = 0.0, w_vadv_b = 0.0, vv1 = 0.0, vv2 = 0.0, deta = 0.0;
    if (VADV) {
      const double eta_k = seta[q], eta_k1 = seta[q + 16];
      const int km = k > 0 ? k - 1 : 0, kp = k < nlev - 1 ? k + 1 : nlev - 1;
      const double phm = (sphi[q] + sphi[q + 16]) / 2;                               // PA:1183
      // preq_vertadv_v1 (prim_si_mod.F90:119-147): the one-sided end formulas are the general one with
      // eta_dot = 0 at the model top and bottom
      const double facp = div_fast(0.5 * eta_k1, dp), facm = div_fast(0.5 * eta_k, dp);
      vv1 = facp * (sv[kp * 32 + p] - v1) + facm * (v1 - sv[km * 32 + p]);
      vv2 = facp * (sv[kp * 32 + 16 + p] - v2) + facm * (v2 - sv[km * 32 + 16 + p]);
      // theta at the interfaces k and k+1, energy-conserving average (PA:1268-1272); zero at top and bottom
      double thi_k = 0.0, thi_k1 = 0.0, tmp_m = 0.0;
      if (k > 0) {
        const double phm_m = (sphi[q - 16] + sphi[q]) / 2;
        thi_k = div_fast(div_fast(-mu_k * (phm - phm_m), exner - sexn[q - 16]), CP);
        tmp_m = ((seta[q - 16] + eta_k) / 2) * (wk - sw[q - 16]);                     // temp(k-1), PA:1279
        phi_vadv_k = div_fast(eta_k * (phm - phm_m), dpi_k);                           // PA:1289,1298
      }
      if (k < nlev - 1) {
        const double phm_p = (sphi[q + 16] + sphi[q + 32]) / 2;
        thi_k1 = div_fast(div_fast(-mu_k1 * (phm_p - phm), sexn[q + 16] - exner), CP);
      }
      const double tmp_k = ((eta_k + eta_k1) / 2) * (wk1 - wk);
      theta_vadv = eta_k1 * thi_k1 - eta_k * thi_k;                                   // PA:1283-1284
      w_vadv_k = div_fast(k > 0 ? (tmp_m + tmp_k) / 2 : tmp_k, dpi_k);                // PA:1288,1291,1297
      w_vadv_b = div_fast(tmp_k, dp);                                                 // interface nlevp: dp3d_i = dp3d(nlev)
      deta = eta_k1 - eta_k;
      if (ACCUM && act) A.der.eta_dot_dpdn[(size_t)ie * 16 * nlevp + q] += A.eta_ave_w * eta_k;   // PA:1309-1316
    }
    const double theta_tens = (-theta_vadv - div_v_theta) * s1;

    // w, phi tendencies at interface k  PA:1326-1344
    const double v_gradw = v1i_k * gw1_k + v2i_k * gw2_k;
    const double w_tens = (-w_vadv_k - v_gradw) * s1 - s2 * G_GRAV * (1 - mu_k);
    const double v_gradphi = v1i_k * gphi1_k + v2i_k * gphi2_k;
    double phi_tens = (-phi_vadv_k - v_gradphi) * s1 + s2 * G_GRAV * wk;
    if (s1 != s2) phi_tens = phi_tens + (s1 - s2) * (v1i_k * gps1 + v2i_k * gps2) * c_hybi[k];
    if (SAVE && act) {          // IM:176-205 for a later DIRK stage: dpnh_dp_i - 1 and g w - gwphis (IM:42-53) of this state
      A.ew[gq] = mu_k - 1;
      A.ephi[gq] = G_GRAV * wk - (k == 0 ? 0.0 : (v1i_k * gps1 + v2i_k * gps2) * c_hybi[k]);
    }

    // momentum PA:1404-1509
    wv1 = wv1 - (wk * gw1_k + wk1 * gw1_k1) / 2;
    wv2 = wv2 - (wk * gw2_k + wk1 * gw2_k1) / 2;
    double gp1 = CP * theta * ge1, gp2 = CP * theta * ge2;
    if (FORM == 2) {                                             // PA:1446-1452
      double gte1, gte2, gth1, gth2;
      gradient_sphere(t, scr + 9 * 16, gte1, gte2);
      gradient_sphere(t, scr + 8 * 16, gth1, gth2);
      gp1 = (gp1 + CP * (gte1 - exner * gth1)) / 2;
      gp2 = (gp2 + CP * (gte2 - exner * gth2)) / 2;
    }
    double mg1 = (mu_k * gphi1_k + mu_k1 * gphi1_k1) / 2;                           // PA:1455-1458
    double mg2 = (mu_k * gphi2_k + mu_k1 * gphi2_k1) / 2;
    if (PGRAD) {                                              // PA:1460-1477
      __syncwarp();
      scr[0 * 16 + p] = log(exner);
      __syncwarp();
      double gl1, gl2;
      gradient_sphere(t, scr, gl1, gl2);
      const double T0 = TREF - TREF_LAPSE * TREF * CP / G_GRAV;
      const double rex = thetal_pow::rcp_fast(exner);
      mg1 = mg1 + CP * T0 * (gl1 - div_by(ge1, exner, rex));
      mg2 = mg2 + CP * T0 * (gl2 - div_by(ge2, exner, rex));
    }
    const double vt1 = (-vv1 + v2 * (fcor + vort) - ke1 - mg1 - gp1 - wv1) * s1;    // PA:1495-1506
    const double vt2 = (-vv2 - v1 * (fcor + vort) - ke2 - mg2 - gp2 - wv2) * s1;

    if (act) {
      // PA:1674-1690  X(np1) = spheremp*(scale3*X(nm1) + dt2*tens); reads of nm1 straight from HBM
      A.tens.v[gqv] = spheremp * (s3 * m_v1 + dt2 * vt1);
      A.tens.v[gqv + 16] = spheremp * (s3 * m_v2 + dt2 * vt2);
      A.tens.th[gq] = spheremp * (s3 * m_th + dt2 * theta_tens);
      A.tens.w[gqi] = spheremp * (s3 * m_w + dt2 * w_tens);
      A.tens.phi[gqi] = spheremp * (s3 * m_phi + dt2 * phi_tens);
      A.tens.dp[gq] = spheremp * (s3 * m_dp - s1 * dt2 * (VADV ? divdp + deta : divdp));   // PA:1693-1694
      if (accum) {
        sdiv[q] = divdp;
        A.der.vn0[gqv] = o_vn1 + A.eta_ave_w * (v1 * dp);                           // PA:1190
        A.der.vn0[gqv + 16] = o_vn2 + A.eta_ave_w * (v2 * dp);
      }
      if (k == nlev - 1) {                                                          // PA:1349-1361, :1699
        const double v_gradw_b = v1 * gw1_k1 + v2 * gw2_k1;                         // v_i(nlevp) = v(nlev)
        const double w_tens_b = (-w_vadv_b - v_gradw_b) * s1 - s1 * G_GRAV * (1 - mu_k1);
        A.tens.w[gqi + 16] = spheremp * (s3 * A.nm1.w[gqi + 16] + dt2 * w_tens_b);
      }
    }
    __syncwarp();   // tiles are rewritten by the next iteration
  }

  // ---- phase 3/4: diagnostic omega (PA:1197-1213, 1312-1313), only when eta_ave_w != 0 ----
  if (accum) {
    // (moving the vn0 read-modify-writes behind the level loop was measured slower on C2; their old values are
    //  requested at the top of each level instead, and the omega_p ones one iteration ahead)
    __syncthreads();
    {
      // pi_i(k+1) = pi_i(k) + dp3d(k), omega_i(k+1) = omega_i(k) + divdp(k) (PA:1197-1202) as a two-level scan over
      // all 256 threads: half-warp c sums its chunk of levels, the chunk totals are combined through shared
      // memory.  (The association differs from the serial sum by roundings of 1e-16; omega_p is a diagnostic.)
      const int CH = (nlev + KS - 1) / KS, kb = kslot * CH;
      double *stot = sdiv + 16 * nlev;                       // scratch tiles are free after the level loop
      double sd = 0.0, sv_ = 0.0;
      for (int i = 0; i < CH; i++) {
        const int k = kb + i;
        if (k < nlev) { sd = sd + sdp[k * 16 + p]; sv_ = sv_ + sdiv[k * 16 + p]; }
      }
      stot[kslot * 16 + p] = sd; stot[256 + kslot * 16 + p] = sv_;
      __syncthreads();
      double pi_i = A.prm.ptop, om_i = 0.0;
      for (int c = 0; c < kslot; c++) { pi_i = pi_i + stot[c * 16 + p]; om_i = om_i + stot[256 + c * 16 + p]; }
      for (int i = 0; i < CH; i++) {
        const int k = kb + i;
        if (k < nlev) {
          const double dp = sdp[k * 16 + p], dv = sdiv[k * 16 + p];
          spnh[k * 16 + p] = pi_i + dp / 2;                 // pi(k), PA:1204
          const double om_next = om_i + dv;
          sdiv[k * 16 + p] = (om_i + om_next) / 2;
          pi_i = pi_i + dp;
          om_i = om_next;
        }
      }
    }
    __syncthreads();
    double *gom = A.der.omega_p + (size_t)ie * 16 * nlev;
    double o_next = kslot < nlev ? gom[kslot * 16 + p] : 0.0;      // old value one iteration ahead of its use
    for (int it = 0; it < niter; it++) {
      const int k = it * KS + kslot;
      const double o_om = o_next;
      if (k + KS < nlev) o_next = gom[(k + KS) * 16 + p];
      if (k < nlev) {
        double g1, g2;
        gradient_sphere(t, spnh + k * 16, g1, g2);
        const double vgrad_p = sv[k * 32 + p] * g1 + sv[k * 32 + 16 + p] * g2;
        const double omega = vgrad_p - sdiv[k * 16 + p];
        gom[k * 16 + p] = o_om + A.eta_ave_w * omega;                                   // PA:1312-1313
      }
    }
  }
}

template <int FORM, bool PGRAD, bool ACCUM, bool VADV, bool SAVE>
static void launch_caar_pre_t(const CaarArgs &A, cudaStream_t st)
{
  const size_t sm = caar_smem_bytes(A.prm.nlev, VADV);
  static size_t configured = 0;
  if (sm > configured) {
    cudaFuncSetAttribute(caar_pre_kernel<FORM, PGRAD, ACCUM, VADV, SAVE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    configured = sm;
  }
  const int nb = A.elist ? A.ecount : A.prm.nelem;
  if (nb > 0) caar_pre_kernel<FORM, PGRAD, ACCUM, VADV, SAVE><<<nb, CAAR_THREADS, sm, st>>>(A);
}

template <int FORM, bool PGRAD>
static void launch_caar_pre_a(const CaarArgs &A, cudaStream_t st)
{
  const bool acc = A.eta_ave_w != 0.0, vadv = A.prm.rsplit == 0, save = A.ew != nullptr && !vadv;
  if (vadv) { acc ? launch_caar_pre_t<FORM, PGRAD, true, true, false>(A, st) : launch_caar_pre_t<FORM, PGRAD, false, true, false>(A, st); }
  else if (save) { acc ? launch_caar_pre_t<FORM, PGRAD, true, false, true>(A, st) : launch_caar_pre_t<FORM, PGRAD, false, false, true>(A, st); }
  else { acc ? launch_caar_pre_t<FORM, PGRAD, true, false, false>(A, st) : launch_caar_pre_t<FORM, PGRAD, false, false, false>(A, st); }
}

void launch_caar_pre(const CaarArgs &A, cudaStream_t st)
{
  const bool pg = A.prm.pgrad_correction == 1;
  switch (A.prm.theta_advect_form) {
    case 0: pg ? launch_caar_pre_a<0, true>(A, st) : launch_caar_pre_a<0, false>(A, st); break;
    case 1: pg ? launch_caar_pre_a<1, true>(A, st) : launch_caar_pre_a<1, false>(A, st); break;
    default: pg ? launch_caar_pre_a<2, true>(A, st) : launch_caar_pre_a<2, false>(A, st); break;
  }
}

// ==========================================================================================
// DSS (edge_mod.F90:500-629,862-944 + bndry_mod.F90:51-169 as a pull-style node sum), then
// x rspheremp (PA:1741-1753), surface dpnh_dp_i fix-up (PA:1757-1772), limiter_dp3d_k
// (PA:1812-1815, 1822-1931).  Reads `tens`, writes the finished state `out` (= np1).
// ==========================================================================================
__device__ __forceinline__ double dss_sum(const double *f, size_t estride, size_t off, double own,
                                          int s0, int s1, int s2)
{
  double s = own;
#ifndef DSS_NO_LDG
  if (s0 >= 0) s = s + __ldg(f + (size_t)(s0 >> 4) * estride + off + (s0 & 15));
  if (s1 >= 0) s = s + __ldg(f + (size_t)(s1 >> 4) * estride + off + (s1 & 15));
  if (s2 >= 0) s = s + __ldg(f + (size_t)(s2 >> 4) * estride + off + (s2 & 15));
#else
  if (s0 >= 0) s = s + f[(size_t)(s0 >> 4) * estride + off + (s0 & 15)];
  if (s1 >= 0) s = s + f[(size_t)(s1 >> 4) * estride + off + (s1 & 15)];
  if (s2 >= 0) s = s + f[(size_t)(s2 >> 4) * estride + off + (s2 & 15)];
#endif
  return s;
}

// slow path of limiter_dp3d_k for one element already written to `out`; executed by 16 threads
__device__ void limiter_columns(const StatePtrs &out, int ie, int nlev, int p, double spheremp,
                                double dp3d_thresh)
{
  double *dp = out.dp + (size_t)ie * 16 * nlev, *th = out.th + (size_t)ie * 16 * nlev;
  double mn = dp[p] - dp3d_thresh * c_dp0[0];
  for (int k = 1; k < nlev; k++) mn = fmin(mn, dp[k * 16 + p] - dp3d_thresh * c_dp0[k]);
  // vtheta_dp -> theta for the whole element (PA:1859), this thread's column
  for (int k = 0; k < nlev; k++) th[k * 16 + p] = th[k * 16 + p] / dp[k * 16 + p];
  if (mn < 0) {                                                                     // PA:1863-1893
    double mass = 0, mass_new = 0;
    for (int k = 0; k < nlev; k++) mass = mass + (dp[k * 16 + p] - dp3d_thresh * c_dp0[k]) * spheremp;
    const double sgn = (mass < 0) ? -1.0 : 1.0;
    for (int k = 0; k < nlev; k++) {
      double Q = (dp[k * 16 + p] - dp3d_thresh * c_dp0[k]) * spheremp;
      if (mass < 0) Q = -Q;
      if (!(Q < 0)) mass_new = mass_new + Q;
    }
    for (int k = 0; k < nlev; k++) {
      double Q = (dp[k * 16 + p] - dp3d_thresh * c_dp0[k]) * spheremp;
      if (mass < 0) Q = -Q;
      if (Q < 0) Q = 0;
      if (mass_new > 0) Q = Q * (fabs(mass) / mass_new);
      if (sgn < 0) Q = -Q;
      dp[k * 16 + p] = Q / spheremp + dp3d_thresh * c_dp0[k];
    }
  }
  for (int k = 0; k < nlev; k++) th[k * 16 + p] = th[k * 16 + p] * dp[k * 16 + p];    // PA:1897
}


constexpr int DSS_THREADS = 256;
#ifndef DSS_MINB
#define DSS_MINB 4
#endif
#ifndef DSS_UNROLL
#define DSS_UNROLL 1
#endif
constexpr int DSS_UNROLL_N = DSS_UNROLL;
#ifndef DSS_NO_LDG
#define DSS_OWN(ptr) __ldg(&(ptr))
#else
#define DSS_OWN(ptr) (ptr)
#endif

// tstep_type 9 closing combination (PA:313-326): a(np1) += (b(nm1) - c(n0))/4 for one element, then limiter_dp3d_k.
// Called by all DSS_THREADS threads of the CTA that owns element ie.
__device__ __forceinline__ void combo9_element(const StatePtrs &a, const StatePtrs &b, const StatePtrs &c, const Geo &geo,
                                               const Params &prm, int ie, int tid)
{
  const int nlev = prm.nlev, nlevp = nlev + 1, p = tid & 15, kslot = tid >> 4;
  constexpr int KS = DSS_THREADS / 16;
  const size_t em = (size_t)ie * 16 * nlev, ei = (size_t)ie * 16 * nlevp, ev = (size_t)ie * 32 * nlev;
  int lim_dp = 0, lim_th = 0;
  for (int k = kslot; k < nlevp; k += KS) {
    const int q = k * 16 + p;
    a.w[ei + q] = a.w[ei + q] + (b.w[ei + q] - c.w[ei + q]) / 4;
    if (k < nlev) {
      const size_t qv = ev + k * 32 + p;
      a.v[qv] = a.v[qv] + (b.v[qv] - c.v[qv]) / 4;
      a.v[qv + 16] = a.v[qv + 16] + (b.v[qv + 16] - c.v[qv + 16]) / 4;
      const double th = a.th[em + q] + (b.th[em + q] - c.th[em + q]) / 4;
      const double dp = a.dp[em + q] + (b.dp[em + q] - c.dp[em + q]) / 4;
      a.th[em + q] = th; a.dp[em + q] = dp;
      a.phi[ei + q] = a.phi[ei + q] + (b.phi[ei + q] - c.phi[ei + q]) / 4;
      if (dp < prm.dp3d_thresh * c_dp0[k]) lim_dp = 1;
      if (th - prm.vtheta_thresh * dp < 0) lim_th = 1;
    }
  }
  if (!__syncthreads_or(lim_dp | lim_th)) return;          // common case: one barrier (block-uniform)
  lim_dp = __syncthreads_or(lim_dp);                                                // PA:324-325
  if (lim_dp) {
    __threadfence_block();
    if (tid < 16) limiter_columns(a, ie, nlev, tid, geo.spheremp[(size_t)ie * 16 + tid], prm.dp3d_thresh);
    __threadfence_block();
    __syncthreads();
    lim_th = 0;
    for (int k = kslot; k < nlev; k += KS) {
      const size_t q = em + (size_t)k * 16 + p;
      if (a.th[q] - prm.vtheta_thresh * a.dp[q] < 0) lim_th = 1;
    }
  }
  lim_th = __syncthreads_or(lim_th);
  if (lim_th) {
    for (int k = kslot; k < nlev; k += KS) {
      const size_t q = em + (size_t)k * 16 + p;
      const double dp = a.dp[q];
      if (a.th[q] - prm.vtheta_thresh * dp < 0) a.th[q] = prm.vtheta_thresh * dp;
    }
  }
}

__global__ void __launch_bounds__(DSS_THREADS, DSS_MINB) caar_dss_post_kernel(const DssArgs A)
{
  const int nlev = A.prm.nlev, nlevp = nlev + 1;
  const int ie = A.elist ? A.elist[blockIdx.x] : (int)blockIdx.x;
  const int tid = threadIdx.x, p = tid & 15, kslot = tid >> 4;
  constexpr int KS = DSS_THREADS / 16;
  const int *src = A.geo.dss_src + ((size_t)ie * 16 + p) * 3;
  const int s0 = src[0], s1 = src[1], s2 = src[2];
  const double rs = A.geo.rspheremp[(size_t)ie * 16 + p];
  const size_t ev = 32 * (size_t)nlev, em = 16 * (size_t)nlev, ei = 16 * (size_t)nlevp;
  int lim_dp = 0, lim_th = 0;
  if (tid < 5) {      // own pre-DSS columns towards L2 (the neighbours' CTAs do the same for theirs)
    if (tid == 0) prefetch_l2(A.tens.v + ie * ev, (unsigned)(ev * 8));
    if (tid == 1) prefetch_l2(A.tens.dp + ie * em, (unsigned)(em * 8));
    if (tid == 2) prefetch_l2(A.tens.th + ie * em, (unsigned)(em * 8));
    if (tid == 3) prefetch_l2(A.tens.w + ie * ei, (unsigned)(ei * 8));
    if (tid == 4) prefetch_l2(A.tens.phi + ie * ei, (unsigned)(ei * 8));
  }

#pragma unroll DSS_UNROLL_N
  for (int k = kslot; k < nlev; k += KS) {
    const size_t ov = (size_t)k * 32, om = (size_t)k * 16;
    const double v1s = dss_sum(A.tens.v, ev, ov, DSS_OWN(A.tens.v[ie * ev + ov + p]), s0, s1, s2);
    const double v2s = dss_sum(A.tens.v, ev, ov + 16, DSS_OWN(A.tens.v[ie * ev + ov + 16 + p]), s0, s1, s2);
    const double dps = dss_sum(A.tens.dp, em, om, DSS_OWN(A.tens.dp[ie * em + om + p]), s0, s1, s2);
    const double ths = dss_sum(A.tens.th, em, om, DSS_OWN(A.tens.th[ie * em + om + p]), s0, s1, s2);
    const double ws = dss_sum(A.tens.w, ei, om, DSS_OWN(A.tens.w[ie * ei + om + p]), s0, s1, s2);
    const double phs = dss_sum(A.tens.phi, ei, om, DSS_OWN(A.tens.phi[ie * ei + om + p]), s0, s1, s2);
    double v1 = rs * v1s, v2 = rs * v2s;
    const double dp = rs * dps, th = rs * ths;
    A.out.dp[ie * em + om + p] = dp;
    A.out.th[ie * em + om + p] = th;
    A.out.w[ie * ei + om + p] = rs * ws;
    A.out.phi[ie * ei + om + p] = rs * phs;
    if (k == nlev - 1) {                                                            // PA:1751-1772
      const size_t ob = (size_t)nlev * 16;
      double wb = rs * dss_sum(A.tens.w, ei, ob, A.tens.w[ie * ei + ob + p], s0, s1, s2);
      const double g1 = A.geo.gradphis[(size_t)ie * 32 + p], g2 = A.geo.gradphis[(size_t)ie * 32 + 16 + p];
      const double dpnh = 1 + (((v1 * g1 + v2 * g2) / G_GRAV - wb) / (G_GRAV + (g1 * g1 + g2 * g2) / (2 * G_GRAV))) / A.dt2;
      wb = wb + A.scale1 * A.dt2 * G_GRAV * (dpnh - 1);
      v1 = v1 - A.scale1 * A.dt2 * (dpnh - 1) * g1 / 2;
      v2 = v2 - A.scale1 * A.dt2 * (dpnh - 1) * g2 / 2;
      A.out.w[ie * ei + ob + p] = wb;
    }
    A.out.v[ie * ev + ov + p] = v1;
    A.out.v[ie * ev + ov + 16 + p] = v2;
    if (dp < A.prm.dp3d_thresh * c_dp0[k]) lim_dp = 1;
    if (th - A.prm.vtheta_thresh * dp < 0) lim_th = 1;
  }
  // limiter_dp3d_k: rare; block-uniform decisions behind one barrier in the common case
  if (A.do_limiter && __syncthreads_or(lim_dp | lim_th)) {
  lim_dp = __syncthreads_or(lim_dp);
  if (lim_dp) {
    __threadfence_block();
    if (tid < 16) limiter_columns(A.out, ie, nlev, tid, A.geo.spheremp[(size_t)ie * 16 + tid], A.prm.dp3d_thresh);
    __threadfence_block();
    __syncthreads();
    lim_th = 0;
    for (int k = kslot; k < nlev; k += KS) {
      const size_t q = ie * em + (size_t)k * 16 + p;
      if (A.out.th[q] - A.prm.vtheta_thresh * A.out.dp[q] < 0) lim_th = 1;
    }
  }
  lim_th = __syncthreads_or(lim_th);
  if (lim_th) {                                                                     // PA:1903-1925
    for (int k = kslot; k < nlev; k += KS) {
      const size_t q = ie * em + (size_t)k * 16 + p;
      const double dp = A.out.dp[q];
      if (A.out.th[q] - A.prm.vtheta_thresh * dp < 0) A.out.th[q] = A.prm.vtheta_thresh * dp;
    }
  }
  }
  // the tstep_type 9 combination that follows the last stage (PA:313-326) is applied to the element this CTA has
  // just finished, while its columns are still in L2: np1 is not streamed through HBM a second time
  if (A.combo_b.v) {
    __threadfence_block();
    __syncthreads();
    combo9_element(A.out, A.combo_b, A.combo_c, A.geo, A.prm, ie, tid);
  }
}

void launch_caar_dss_post(const DssArgs &A, cudaStream_t st)
{
  const int nb = A.elist ? A.ecount : A.prm.nelem;
  if (nb > 0) caar_dss_post_kernel<<<nb, DSS_THREADS, 0, st>>>(A);
}

// ==========================================================================================
// compute_stage_value_dirk (IM:60-364): per-column Newton solve, one CTA per element.
// (k,p)-parallel phases for EOS / residual / update, the tridiagonal recurrence (IM:576-599)
// runs on one warp; convergence is decided per element as in IM:330-335.
// ==========================================================================================

constexpr int DIRK_THREADS = 256;
constexpr int DIRK_ARRAYS = 6;
#ifndef DIRK_MINB
#define DIRK_MINB 3
#endif

__host__ __device__ inline size_t dirk_smem_bytes(int nlev)
{
  return sizeof(double) * (16 * ((size_t)DIRK_ARRAYS * nlev + 1) + 144);
}

__device__ __forceinline__ double block_max(double v, double *red, int tid)
{
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if ((tid & 31) == 0) red[tid >> 5] = v;
  __syncthreads();
  double r = red[0];
  for (int i = 1; i < DIRK_THREADS / 32; i++) r = fmax(r, red[i]);
  return r;
}

// IM:42-53 at interface k (1 <= k < nlev), state arrays of one element in global memory
__device__ __forceinline__ double gwphis_at(const double *v, const double *dp, int k, int p,
                                            double g1, double g2)
{
  const double dk = dp[k * 16 + p], dm = dp[(k - 1) * 16 + p];
  const double vi1 = div_fast(dk * v[k * 32 + p] + dm * v[(k - 1) * 32 + p], dk + dm);
  const double vi2 = div_fast(dk * v[k * 32 + 16 + p] + dm * v[(k - 1) * 32 + 16 + p], dk + dm);
  return (vi1 * g1 + vi2 * g2) * c_hybi[k];
}

// Layout of the solve: thread tid owns the point-levels q = tid + j*256 (j < ITEMS) of the element
// for the whole kernel, so everything only its owner touches (Rgas*vtheta_dp, the stage right-hand
// sides w0 and dphi0) stays in registers; only what a vertical neighbour or the tridiagonal sweep
// reads lives in shared memory (6 tiles per level: 1/dp_i, w, pnh, pnh/dphi, x, dphi).  That is
// 55 KB per element at L72, so three to four elements are resident per SM and the serial sweeps of
// one element overlap the pow-heavy parallel phases of the others.
// Deviations from the Fortran operation order (all far below the 1e-10 parity budget):
//  * (pnh_k - pnh_{k-1}) is multiplied by a stored 1/dp_i instead of divided by dp_i (<= 1 ulp of mu);
//  * the tridiagonal solve multiplies by reciprocals of the pivots (the solve only steers Newton);
//  * the EOS after the last Newton update is skipped (its result is never used, IM:324-336).
template <int ITEMS>
__global__ void __launch_bounds__(DIRK_THREADS, (ITEMS >= 6 ? 2 : DIRK_MINB)) dirk_kernel(const DirkArgs A)
{
  extern __shared__ double smem[];
  const int nlev = A.prm.nlev, nlevp = nlev + 1;
  const int ie = blockIdx.x, tid = threadIdx.x, n16 = 16 * nlev, p = tid & 15;
  double *srdpi = smem, *sw = srdpi + n16, *spnh = sw + n16, *sc = spnh + n16, *sr = sc + n16;
  double *sx = sr + n16;                   // nlevp tiles (phi0 incl. phis during set-up)
  double *red = sx + n16 + 16;
  const double g = G_GRAV, dt2 = A.dt2;
  const double ptop = A.prm.ptop;
  const int medge = (nlev >> 1) - 1;       // last row of the downward Thomas sweep

  const double *gth = A.np1.th + (size_t)ie * n16, *gdp = A.np1.dp + (size_t)ie * n16;
  const double *gv = A.np1.v + (size_t)ie * 2 * n16;
  double *gw = A.np1.w + (size_t)ie * 16 * nlevp, *gphi = A.np1.phi + (size_t)ie * 16 * nlevp;
  const double gp1 = A.geo.gradphis[(size_t)ie * 32 + p], gp2 = A.geo.gradphis[(size_t)ie * 32 + 16 + p];
  int bad = 0;
  double rth[ITEMS], w0[ITEMS], dphi0[ITEMS], phi0[ITEMS];
  if (tid < 5) {      // start the element's columns towards L2 (v is only read by compute_gwphis)
    const unsigned bv = 32u * nlev * 8u, bm = 16u * nlev * 8u, bi = 16u * nlevp * 8u;
    if (tid == 0) prefetch_l2(gdp, bm);
    if (tid == 1) prefetch_l2(gth, bm);
    if (tid == 2) prefetch_l2(gw, bi);
    if (tid == 3) prefetch_l2(gphi, bi);
    if (tid == 4) prefetch_l2(gv, bv);
  }

  // ---- load, wmax (IM:168-170) ----
  double wmax = 0.0;
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const int q = tid + j * DIRK_THREADS;
    rth[j] = 0; w0[j] = 0; dphi0[j] = 0; phi0[j] = 0;
    if (q < n16) {
      const double dp = gdp[q], th = gth[q];
      const double dpi = (q >= 16) ? (dp + gdp[q - 16]) / 2 : dp;
      srdpi[q] = div_fast(1.0, dpi);
      sc[q] = dp;                           // temporary: pressure scan below
      if (th < 0 || dp < 0) bad = 1;
      rth[j] = RGAS * th;
      const double w = gw[q];
      w0[j] = w; phi0[j] = gphi[q];
      wmax = fmax(wmax, fabs(w));
    }
  }
  if (tid < 16) wmax = fmax(wmax, fabs(gw[n16 + tid]));
  wmax = fmax(1.0, block_max(wmax, red, tid));

  // ---- explicit stiff terms of earlier stages (IM:176-205) ----
  for (int pass = 0; pass < 2; pass++) {
    const double dt3 = pass == 0 ? A.alphadt_n0 : A.alphadt_nm1;
    if (dt3 == 0.0) continue;
    const StatePtrs &S = pass == 0 ? A.n0 : A.nm1;
    const double *tth = S.th + (size_t)ie * n16, *tdp = S.dp + (size_t)ie * n16;
    const double *tphi = S.phi + (size_t)ie * 16 * nlevp, *tw = S.w + (size_t)ie * 16 * nlevp;
    const double *tv = S.v + (size_t)ie * 2 * n16;
    for (int q = tid; q < n16; q += DIRK_THREADS) {
      const double dphi = tphi[q + 16] - tphi[q], th = tth[q], dp = tdp[q];
      if (th < 0 || dp < 0 || dphi > 0) bad = 1;
      spnh[q] = eos_pnh(div_fast(RGAS * th, -dphi));
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
      const int q = tid + j * DIRK_THREADS;
      if (q < n16) {
        const int k = q >> 4;
        double mu;
        if (k == 0) mu = div_fast(2 * (spnh[q] - ptop), tdp[q]);
        else mu = div_fast(spnh[q] - spnh[q - 16], (tdp[q] + tdp[q - 16]) / 2);
        w0[j] = w0[j] + dt3 * g * (mu - 1);
        const double gwh = (k == 0) ? 0.0 : gwphis_at(tv, tdp, k, p, gp1, gp2);
        phi0[j] = phi0[j] + dt3 * g * tw[q] - dt3 * gwh;
      }
    }
    __syncthreads();
  }
  // ---- wphis(np1) term (IM:208-211); phi0 goes to global (it is the base of the final update) ----
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const int q = tid + j * DIRK_THREADS;
    if (q < n16) {
      const int k = q >> 4;
      const double gwh = (k == 0) ? 0.0 : gwphis_at(gv, gdp, k, p, gp1, gp2);
      phi0[j] = phi0[j] - dt2 * gwh;
      sx[q] = phi0[j];
      gphi[q] = phi0[j];
    }
  }
  if (tid < 16) sx[n16 + tid] = gphi[n16 + tid];
  __syncthreads();
  // ---- hydrostatic initial guess phi_from_eos (EOS:286-302, IM:228-230) ----
  if (tid < 16) {
    double p_i = ptop;
    for (int k = 0; k < nlev; k++) { const double pn = p_i + sc[k * 16 + tid]; spnh[k * 16 + tid] = (pn + p_i) / 2; p_i = pn; }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const int q = tid + j * DIRK_THREADS;
    const int qq = min(q, n16 - 1);
    const double term = div_fast(rth[j] * thetal_pow::pow_pos(div_fast(spnh[qq], P0), C_KM1), P0);
    if (q < n16) sr[q] = term;
  }
  __syncthreads();
  if (tid < 16) {   // sr <- phi_np1 (hydrostatic), bottom-up
    double ph = sx[n16 + tid];
    for (int k = nlev - 1; k >= 0; k--) { ph = ph + sr[k * 16 + tid]; sr[k * 16 + tid] = ph; }
  }
  __syncthreads();
  int nsafe = 0;
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {                                                 // IM:234-253
    const int q = tid + j * DIRK_THREADS;
    if (q < n16) {
      const double below = (q + 16 < n16) ? sr[q + 16] : sx[q + 16];
      double dphi = below - sr[q];
      dphi0[j] = sx[q + 16] - sx[q];
      if (dphi > -g) { dphi = -g; nsafe = 1; }
      spnh[q] = dphi;                       // temporary
    }
  }
  nsafe = __syncthreads_or(nsafe);
  if (nsafe) {                                                                      // IM:254-260
    if (tid < 16) {
      double ph = sx[n16 + tid];
      for (int k = nlev - 1; k >= 0; k--) { ph = ph - spnh[k * 16 + tid]; sr[k * 16 + tid] = ph; }
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const int q = tid + j * DIRK_THREADS;
    const int qq = min(q, n16 - 1);           // branch-free body: the point-levels of a thread overlap
    const double w_guess = div_fast(sr[qq] - sx[qq], dt2 * g);
    const double dphi = spnh[qq];
    if (q < n16 && dphi > 0) bad = 1;   // (a clamped duplicate may read a slot already overwritten)
    const double pnh = eos_pnh(div_fast(rth[j], -dphi));                            // IM:261
    const double cq = pnh * rcp_fast(dphi);
    if (q < n16) {
      sw[q] = w_guess;
      spnh[q] = pnh;
      sc[q] = cq;
      if ((q >> 4) == medge) red[64 + p] = cq;
    }
  }
  __syncthreads();

  // ---- Newton iteration (IM:268-336) ----
  int itercount = 0;
  double deltaerr = 0.0;
  const int maxiter = 20;
  const double ajac = (dt2 * g) * (dt2 * g) / (1 - KAPPA);                          // IM:414
  unsigned long long *amin = (unsigned long long *)(red + 32);
  int *cflag = (int *)(red + 48);
  while (itercount < maxiter) {
    // residual Fn (IM:263-264 / 326), x = -Fn
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
      const int q = tid + j * DIRK_THREADS;
      if (q < n16) {
        double mu;
        if (q < 16) mu = 2 * (spnh[q] - ptop) * srdpi[q];
        else mu = (spnh[q] - spnh[q - 16]) * srdpi[q];
        const double Fn = sw[q] - (w0[j] + g * dt2 * (mu - 1));
        sx[q] = -Fn;
      }
    }
    __syncthreads();
    // tridiagonal solve J x = -Fn.  Jacobian rows (IM:414-433) are assembled on the fly from
    // c = pnh/dphi and b = a/dp_i.  The Thomas recurrence (IM:590-598) is run from both ends at once:
    // warp 0 eliminates rows 0..m-1 downwards, warp 1 rows nlev-1..m upwards (same arithmetic with
    // the sub- and super-diagonal swapped); the reduced rows m-1 and m form a 2x2 system, then both
    // back-substitute away from the middle.  Rows are stored normalised (x_k + sc_k x_next = sx_k),
    // so back-substitution is one FMA per row.  One column per lane (16 lanes of each warp).
    if (tid < 64) {
      const int m = nlev >> 1;
      const bool up = tid >= 32;
      double myc = 0.0, myy = 0.0;
      if ((tid & 31) < 16) {
        if (!up) {
          const double *pc = sc + p, *pd = srdpi + p;
          double *wc = sc + p, *wx = sx + p;
          double cprev = pc[0];
          double Uprev = 2 * (ajac * pd[0]) * cprev;
          double r = rcp_fast(1 - Uprev);
          double y = wx[0];
          myc = Uprev * r; myy = y * r;
          wc[0] = myc; wx[0] = myy;
          double cn = pc[16], bn = ajac * pd[16], rn = wx[16];
          for (int k = 1; k < m; k++) {
            pc += 16; pd += 16; wc += 16; wx += 16;
            const double c = cn, b = bn, rhs = rn;
            if (k + 1 < m) { cn = pc[16]; bn = ajac * pd[16]; rn = wx[16]; }
            const double L = b * cprev, U = b * c;
            const double diag = 1 - L - U;
            const double P = fma(-(L * Uprev), r, diag);
            y = fma(-(L * r), y, rhs);
            r = rcp_fast(P);
            myc = U * r; myy = y * r;
            wc[0] = myc; wx[0] = myy;
            Uprev = U; cprev = c;
          }
        } else {
          const int q0 = (nlev - 1) * 16 + p;
          const double c_edge = red[64 + p];            // c_{m-1}, saved where c is computed
          const double *pc = sc + q0, *pd = srdpi + q0;
          double *wc = sc + q0, *wx = sx + q0;
          double c = pc[0];
          double cm1 = (nlev - 2 >= m) ? pc[-16] : c_edge;
          const double b0 = ajac * pd[0];
          double Lprev = b0 * cm1;
          double r = rcp_fast(1 - Lprev - b0 * c);
          double y = wx[0];
          myc = Lprev * r; myy = y * r;
          wc[0] = myc; wx[0] = myy;
          double bn = 0.0, rn = 0.0, cn = 0.0;         // row k-1 operands: b, rhs and c_{k-2}
          if (nlev - 2 >= m) { bn = ajac * pd[-16]; rn = wx[-16]; cn = (nlev - 3 >= m) ? pc[-32] : c_edge; }
          for (int k = nlev - 2; k >= m; k--) {
            pc -= 16; pd -= 16; wc -= 16; wx -= 16;
            c = cm1; cm1 = cn;
            const double b = bn, rhs = rn;
            if (k - 1 >= m) { bn = ajac * pd[-16]; rn = wx[-16]; cn = (k - 2 >= m) ? pc[-32] : c_edge; }
            const double L = b * cm1, U = b * c;
            const double diag = 1 - L - U;
            const double P = fma(-(U * Lprev), r, diag);
            y = fma(-(U * r), y, rhs);
            r = rcp_fast(P);
            myc = L * r; myy = y * r;
            wc[0] = myc; wx[0] = myy;
            Lprev = L;
          }
        }
        red[(up ? 112 : 80) + p] = myc;
        red[(up ? 128 : 96) + p] = myy;
      }
      asm volatile("bar.sync 1, 64;" ::: "memory");
      if ((tid & 31) < 16) {
        const double cp = red[80 + p], yp = red[96 + p], lp = red[112 + p], zp = red[128 + p];
        const double xm = (zp - lp * yp) * rcp_fast(1 - lp * cp);
        if (!up) {
          double x = yp - cp * xm;
          double *wx = sx + (m - 1) * 16 + p;
          const double *pc = sc + (m - 1) * 16 + p;
          wx[0] = x;
          double un = 0.0, yn = 0.0;
          if (m >= 2) { un = pc[-16]; yn = wx[-16]; }
          for (int k = m - 2; k >= 0; k--) {
            pc -= 16; wx -= 16;
            const double u = un, yy = yn;
            if (k > 0) { un = pc[-16]; yn = wx[-16]; }
            x = fma(-u, x, yy);
            wx[0] = x;
          }
        } else {
          double x = xm;
          double *wx = sx + m * 16 + p;
          const double *pc = sc + m * 16 + p;
          wx[0] = x;
          double un = 0.0, yn = 0.0;
          if (m + 1 < nlev) { un = pc[16]; yn = wx[16]; }
          for (int k = m + 1; k < nlev; k++) {
            pc += 16; wx += 16;
            const double u = un, yy = yn;
            if (k + 1 < nlev) { un = pc[16]; yn = wx[16]; }
            x = fma(-u, x, yy);
            wx[0] = x;
          }
        }
      }
    }
    __syncthreads();
    // dphi update (IM:278-283) and positivity check
    int anypos = 0;
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
      const int q = tid + j * DIRK_THREADS;
      if (q < n16) {
        double dphi;
        if (q + 16 < n16) dphi = dphi0[j] + dt2 * g * ((sw[q + 16] - sw[q]) + (sx[q + 16] - sx[q]));
        else dphi = dphi0[j] - dt2 * g * (sw[q] + sx[q]);
        sr[q] = dphi;
        if (dphi >= 0) anypos = 1;
      }
    }
    anypos = __syncthreads_or(anypos);
    double alpha = 1.0;
    if (anypos) {                                                                   // IM:285-319 (rare)
      if (tid < 16) {
        int colpos = 0;
        for (int k = 0; k < nlev; k++) if (sr[k * 16 + tid] >= 0) colpos = 1;
        cflag[tid] = colpos;
        amin[tid] = (unsigned long long)__double_as_longlong(1.0);
      }
      __syncthreads();
      if (cflag[p]) {
#pragma unroll
        for (int j = 0; j < ITEMS; j++) {
          const int q = tid + j * DIRK_THREADS;
          if (q < n16) {
            double dx, dw;
            if (q + 16 < n16) { dx = sx[q + 16] - sx[q]; dw = sw[q + 16] - sw[q]; }
            else { dx = -sx[q]; dw = -sw[q]; }
            if (dx != 0) {
              const double alpha_k = -(dphi0[j] + dt2 * g * dw) / (dt2 * g * dx);
              if (alpha_k >= 0) atomicMin(&amin[p], (unsigned long long)__double_as_longlong(fabs(alpha_k)));
            }
          }
        }
      }
      __syncthreads();
      if (cflag[p]) {
        alpha = __longlong_as_double((long long)amin[p]) / 2;
#pragma unroll
        for (int j = 0; j < ITEMS; j++) {
          const int q = tid + j * DIRK_THREADS;
          if (q < n16) {
            if (q + 16 < n16) sr[q] = dphi0[j] + dt2 * g * ((sw[q + 16] - sw[q]) + alpha * (sx[q + 16] - sx[q]));
            else sr[q] = dphi0[j] - dt2 * g * (sw[q] + alpha * sx[q]);
          }
        }
      }
      __syncthreads();
    }
    double xmax = 0.0;
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
      const int q = tid + j * DIRK_THREADS;
      if (q < n16) {
        const double x = sx[q];
        sw[q] = sw[q] + alpha * x;                                                  // IM:320-322
        xmax = fmax(xmax, fabs(x));
      }
    }
    xmax = block_max(xmax, red, tid);
    deltaerr = div_fast(xmax, wmax);                                                         // IM:330
    itercount++;
    if (deltaerr < 1.0e-11) break;                                                  // IM:335
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
      const int q = tid + j * DIRK_THREADS;
      const int qq = min(q, n16 - 1);
      const double dphi = sr[qq];
      if (q < n16 && dphi > 0) bad = 1;
      const double pnh = eos_pnh(div_fast(rth[j], -dphi));                          // IM:324
      const double cq = pnh * rcp_fast(dphi);
      if (q < n16) {
        spnh[q] = pnh;
        sc[q] = cq;
        if ((q >> 4) == medge) red[64 + p] = cq;
      }
    }
    __syncthreads();
  }
  // ---- phi update (IM:339) and write back ----
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const int q = tid + j * DIRK_THREADS;
    if (q < n16) {
      const double w = sw[q];
      gw[q] = w;
      gphi[q] = gphi[q] + dt2 * g * w;      // gphi holds phi0 since the set-up phase
    }
  }
  if (bad) atomicOr(A.err, ERR_EOS_BAD_STATE);
  if (tid == 0) {
    atomicMax(A.max_iter, itercount);
    atomicMax(A.max_deltaerr_bits, (unsigned long long)__double_as_longlong(deltaerr));
    if (itercount >= maxiter) atomicOr(A.err, ERR_DIRK_NOT_CONVERGED);
  }
}

template <int ITEMS>
static void launch_dirk_t(const DirkArgs &A, cudaStream_t st, size_t sm)
{
  static size_t configured = 0;
  if (sm > configured) {
    cudaFuncSetAttribute(dirk_kernel<ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    configured = sm;
  }
  dirk_kernel<ITEMS><<<A.prm.nelem, DIRK_THREADS, sm, st>>>(A);
}

}  // namespace thetal
namespace thetal {
#include "thetal_dirk_col.cuh"

void launch_dirk(const DirkArgs &A, cudaStream_t st)
{
  static const int variant = [] { const char *e = getenv("THETAL_B200_DIRK"); return e ? atoi(e) : 1; }();
  if (variant == 1 && A.tens.v) {     // the same solve behind the DSS of the preceding CAAR stage (one pass over the state)
    switch ((A.prm.nlev + 15) / 16) {
      case 0: case 1: case 2: launch_dirk_col_t<2, 16, true>(A, st); break;
      case 3: launch_dirk_col_t<3, 16, true>(A, st); break;
      case 4: launch_dirk_col_t<4, 16, true>(A, st); break;
      case 5: launch_dirk_col_t<5, 16, true>(A, st); break;
      case 6: launch_dirk_col_t<6, 16, true>(A, st); break;
      default: launch_dirk_col_t<4, 32, true>(A, st); break;
    }
    return;
  }
  if (variant == 1) {       // column-parallel solve (half-warp per column, partition + cyclic reduction)
    switch ((A.prm.nlev + 15) / 16) {
      case 0: case 1: case 2: launch_dirk_col_t<2, 16>(A, st); break;
      case 3: launch_dirk_col_t<3, 16>(A, st); break;
      case 4: launch_dirk_col_t<4, 16>(A, st); break;
      case 5: launch_dirk_col_t<5, 16>(A, st); break;
      case 6: launch_dirk_col_t<6, 16>(A, st); break;
      default: {      // 97..128 levels: a full warp per column (4 levels per lane, 512 threads) keeps the registers at 128
        static const int wide = [] { const char *e = getenv("THETAL_B200_DIRK_TPC32"); return e ? atoi(e) : 1; }();
        if (wide) launch_dirk_col_t<4, 32>(A, st);
        else if (A.prm.nlev <= 112) launch_dirk_col_t<7, 16>(A, st);
        else launch_dirk_col_t<8, 16>(A, st);
        break;
      }
    }
    return;
  }
  const size_t sm = dirk_smem_bytes(A.prm.nlev);
  const int need = (16 * A.prm.nlev + DIRK_THREADS - 1) / DIRK_THREADS;
  if (need <= 2) launch_dirk_t<2>(A, st, sm);
  else if (need <= 4) launch_dirk_t<4>(A, st, sm);
  else if (need == 5) launch_dirk_t<5>(A, st, sm);
  else if (need == 6) launch_dirk_t<6>(A, st, sm);
  else launch_dirk_t<8>(A, st, sm);
}

// ==========================================================================================
// Hyperviscosity (PA:546-862, VT:39-169).  Weak-form operators (DM:453-580, 784-858, 1072-1251).
// ==========================================================================================
struct PtW {                 // extra per-point data of the weak operators
  double dta[4], dtb[4];     // -sfi Dvv(a,j), -sfi Dvv(b,j): the test-function derivative rows of DM:821-833 with their sign and scale
  double ca[4], cb[4];       // sfi mp(j,b) Dvv(a,j) , sfi mp(a,j) Dvv(b,j): the level-independent factors of DM:493-503, 555-569
  double mm[4];              // metinv (index b2*2+a2) times metdet
  double ms[3];              // spheremp * Dinv^T Dinv (symmetric: 11, 12, 22): gradient_sphere followed by the
                             // first half of divergence_sphere_wk (DM:410-450, 814-819) as one 2x2 product
  double spheremp;
  double lrf2;               // laplacian_rigid_factor^2
};

__device__ __forceinline__ void ptw_init(PtW &w, const Pt &t, const Geo &g, int ie, double lrf)
{
#pragma unroll
  for (int j = 0; j < 4; j++) {
    w.ca[j] = t.sfi * (c_mp[t.b * 4 + j] * c_dvv[j * 4 + t.a]); w.cb[j] = t.sfi * (c_mp[j * 4 + t.a] * c_dvv[j * 4 + t.b]);
    w.dta[j] = -(t.sfi * c_dvv[j * 4 + t.a]); w.dtb[j] = -(t.sfi * c_dvv[j * 4 + t.b]);
  }
#pragma unroll
  for (int q = 0; q < 4; q++) w.mm[q] = g.metinv[(size_t)ie * 64 + q * 16 + t.p] * t.metdet;
  w.spheremp = g.spheremp[(size_t)ie * 16 + t.p];
  w.lrf2 = lrf * lrf;
  const double sc = w.spheremp;            // ddx, ddy carry sfi
  w.ms[0] = sc * (t.Dinv[0] * t.Dinv[0] + t.Dinv[2] * t.Dinv[2]);
  w.ms[1] = sc * (t.Dinv[0] * t.Dinv[1] + t.Dinv[2] * t.Dinv[3]);
  w.ms[2] = sc * (t.Dinv[1] * t.Dinv[1] + t.Dinv[3] * t.Dinv[3]);
}

constexpr int HV_THREADS = 256;
constexpr int HV_SLOTS = 18;
#ifndef HV_MINB
#define HV_MINB 2
#endif

// Laplacians of the six fields of one level (scalar: DM:1072-1119 with var_coef=.false.;
// vector: DM:1212-1251).  All 16 lanes of the half-warp must call this together.
// scr: HV_SLOTS tiles private to the half-warp.
__device__ __forceinline__ void laplace_level(const Pt &t, const PtW &w, double *scr, double nu_ratio,
                                              const double s_in[4], double v1, double v2,
                                              double lap_s[4], double &lap_v1, double &lap_v2)
{
  const int p = t.p;
  double gv1, gv2, vc1, vc2;
  div_prep(t, v1, v2, gv1, gv2);
  vort_prep(t, v1, v2, vc1, vc2);
#pragma unroll
  for (int f = 0; f < 4; f++) scr[f * 16 + p] = s_in[f];
  scr[4 * 16 + p] = gv1; scr[5 * 16 + p] = gv2; scr[6 * 16 + p] = vc1; scr[7 * 16 + p] = vc2;
  __syncwarp();
#pragma unroll
  for (int f = 0; f < 4; f++) {
    // gradient_sphere, then divergence_sphere_wk first half (DM:814-819) times spheremp: one symmetric 2x2 product
    const double dx = ddx(t, scr + f * 16), dy = ddy(t, scr + f * 16);
    scr[(8 + 2 * f) * 16 + p] = w.ms[0] * dx + w.ms[1] * dy;
    scr[(9 + 2 * f) * 16 + p] = w.ms[1] * dx + w.ms[2] * dy;
  }
  const double div = nu_ratio * div_finish(t, scr + 4 * 16, scr + 5 * 16);
  const double vor = vort_finish(t, scr + 6 * 16, scr + 7 * 16);
  scr[16 * 16 + p] = div;
  scr[17 * 16 + p] = vor;
  __syncwarp();
#pragma unroll
  for (int f = 0; f < 4; f++) {                                                     // DM:821-833
    const double *T1 = scr + (8 + 2 * f) * 16 + t.b * 4, *T2 = scr + (9 + 2 * f) * 16 + t.a;
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < 4; j++) acc = fma(T1[j], w.dta[j], fma(T2[4 * j], w.dtb[j], acc));
    lap_s[f] = acc;
  }
  {
    const double *sd_r = scr + 16 * 16 + t.b * 4, *sd_c = scr + 16 * 16 + t.a;      // div(j,b), div(a,j)
    const double *sv_r = scr + 17 * 16 + t.b * 4, *sv_c = scr + 17 * 16 + t.a;
    // gradient_sphere_wk_testcov (DM:555-569) and curl_sphere_wk_testcov (DM:493-503): the products of mp, Dvv,
    // metinv and metdet do not depend on the level, so each of the four sums over j is taken once with the
    // combined factors (the reference multiplies them out inside the sum: same value up to rounding)
    double sa = 0.0, sb = 0.0, va = 0.0, vb = 0.0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      sa = fma(w.ca[j], sd_r[j], sa); sb = fma(w.cb[j], sd_c[4 * j], sb);
      va = fma(w.ca[j], sv_r[j], va); vb = fma(w.cb[j], sv_c[4 * j], vb);
    }
    const double gc1 = -(w.mm[0] * sa + w.mm[1] * sb), gc2 = -(w.mm[2] * sa + w.mm[3] * sb);
    const double cc1 = -vb, cc2 = va;
    // covariant -> lat-lon of (grad - curl) in one product, DM:1241-1247
    const double d1 = gc1 - cc1, d2 = gc2 - cc2;
    lap_v1 = (t.D[0] * d1 + t.D[2] * d2) + 2 * w.spheremp * v1 * w.lrf2;
    lap_v2 = (t.D[1] * d1 + t.D[3] * d2) + 2 * w.spheremp * v2 * w.lrf2;
  }
  __syncwarp();
}


// Tensor hyperviscosity (hypervis_scaling > 0): weak laplacians of NF scalar tiles with the
// gradient multiplied by elem%tensorVisc before the weak divergence (laplace_sphere_wk with var_coef,
// DM:1090-1117).  tv[b*2+a] = tensorVisc(i,j,a,b).  Uses 3*NF scratch tiles; all 16 lanes call it together.
template <int NF, bool TENSOR>
__device__ __forceinline__ void laplace_scalars(const Pt &t, const PtW &w, const double tv[4], double *scr,
                                                const double *s_in, double *lap_s)
{
  const int p = t.p;
#pragma unroll
  for (int f = 0; f < NF; f++) scr[f * 16 + p] = s_in[f];
  __syncwarp();
#pragma unroll
  for (int f = 0; f < NF; f++) {
    if (TENSOR) {
      double g1, g2;
      gradient_sphere(t, scr + f * 16, g1, g2);
      const double o1 = g1, o2 = g2;
      g1 = o1 * tv[0] + o2 * tv[2];
      g2 = o1 * tv[1] + o2 * tv[3];
      scr[(NF + 2 * f) * 16 + p] = w.spheremp * (t.Dinv[0] * g1 + t.Dinv[2] * g2);      // DM:814-819
      scr[(NF + 2 * f + 1) * 16 + p] = w.spheremp * (t.Dinv[1] * g1 + t.Dinv[3] * g2);
    } else {
      const double dx = ddx(t, scr + f * 16), dy = ddy(t, scr + f * 16);
      scr[(NF + 2 * f) * 16 + p] = w.ms[0] * dx + w.ms[1] * dy;
      scr[(NF + 2 * f + 1) * 16 + p] = w.ms[1] * dx + w.ms[2] * dy;
    }
  }
  __syncwarp();
#pragma unroll
  for (int f = 0; f < NF; f++) {                                                      // DM:821-833
    const double *T1 = scr + (NF + 2 * f) * 16 + t.b * 4, *T2 = scr + (NF + 2 * f + 1) * 16 + t.a;
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < 4; j++) acc = fma(T1[j], w.dta[j], fma(T2[4 * j], w.dtb[j], acc));
    lap_s[f] = acc;
  }
  __syncwarp();
}

// vlaplace_sphere_wk_cartesian (DM:1164-1208): lat-lon -> three Cartesian components, tensor laplacian of each,
// back to lat-lon, plus the rigid-rotation correction.  s2c[c2*3+c3] = vec_sphere2cart(i,j,c3,c2).
__device__ __forceinline__ void vlaplace_cartesian(const Pt &t, const PtW &w, const double tv[4], const double s2c[6],
                                                   double *scr, double v1, double v2, double &lap_v1, double &lap_v2)
{
  double dum[3], out[3];
#pragma unroll
  for (int c = 0; c < 3; c++) dum[c] = s2c[c] * v1 + s2c[3 + c] * v2;
  laplace_scalars<3, true>(t, w, tv, scr, dum, out);
  lap_v1 = out[0] * s2c[0] + out[1] * s2c[1] + out[2] * s2c[2];
  lap_v2 = out[0] * s2c[3] + out[1] * s2c[4] + out[2] * s2c[5];
  lap_v1 = lap_v1 + 2 * w.spheremp * v1 * w.lrf2;
  lap_v2 = lap_v2 + 2 * w.spheremp * v2 * w.lrf2;
}

__host__ __device__ inline size_t hv_smem_bytes() { return sizeof(double) * (HV_THREADS / 16) * HV_SLOTS * 16; }

// pass 1 (PA:629-636 + VT:93-107): laplacians of (state - reference state) -> tout
__global__ void __launch_bounds__(HV_THREADS, HV_MINB) hv_laplace1_kernel(const HvArgs A)
{
  extern __shared__ double smem[];
  const int nlev = A.prm.nlev, nlevp = nlev + 1, ie = A.elist ? A.elist[blockIdx.x] : (int)blockIdx.x, tid = threadIdx.x;
  const int p = tid & 15, kslot = tid >> 4;
  constexpr int KS = HV_THREADS / 16;
  double *scr = smem + kslot * (HV_SLOTS * 16);
  Pt t; PtW w;
  pt_init(t, A.geo, ie, p, A.prm.rearth_inv);
  ptw_init(w, t, A.geo, ie, A.prm.rearth_inv);
  const size_t em = (size_t)ie * 16 * nlev, ei = (size_t)ie * 16 * nlevp, ev = (size_t)ie * 32 * nlev;
  if (tid < 8) {
    const unsigned bv = 32u * nlev * 8u, bm = 16u * nlev * 8u, bi = 16u * nlevp * 8u;
    if (tid == 0) prefetch_l2(A.st.v + ev, bv);
    if (tid == 1) prefetch_l2(A.st.dp + em, bm);
    if (tid == 2) prefetch_l2(A.st.th + em, bm);
    if (tid == 3) prefetch_l2(A.st.w + ei, bi);
    if (tid == 4) prefetch_l2(A.st.phi + ei, bi);
    if (tid == 5) prefetch_l2(A.ref.dp_ref + em, bm);
    if (tid == 6) prefetch_l2(A.ref.theta_ref + em, bm);
    if (tid == 7) prefetch_l2(A.ref.phi_ref + ei, bi);
  }
  const int niter = (nlev + KS - 1) / KS;
  // software pipeline: the nine input values of the next level are requested before the laplacians of
  // the current one are computed
  double n_dp, n_th, n_w, n_phi, n_v1, n_v2, n_rdp, n_rth, n_rphi;
  auto fetch = [&](int it_) {
    const int kk_ = it_ * KS + kslot;
    const int k_ = kk_ < nlev ? kk_ : nlev - 1, q_ = k_ * 16 + p;
    n_dp = A.st.dp[em + q_]; n_th = A.st.th[em + q_]; n_w = A.st.w[ei + q_]; n_phi = A.st.phi[ei + q_];
    n_v1 = A.st.v[ev + k_ * 32 + p]; n_v2 = A.st.v[ev + k_ * 32 + 16 + p];
    n_rdp = A.ref.dp_ref[em + q_]; n_rth = A.ref.theta_ref[em + q_]; n_rphi = A.ref.phi_ref[ei + q_];
  };
  fetch(0);
  for (int it = 0; it < niter; it++) {
    const int kk = it * KS + kslot;
    const bool act = kk < nlev;
    const int k = act ? kk : nlev - 1, q = k * 16 + p;
    const double dp = n_dp, thv = n_th, wv = n_w, phiv = n_phi, v1 = n_v1, v2 = n_v2;
    const double rdp = n_rdp, rth = n_rth, rphi = n_rphi;
    if (it + 1 < niter) fetch(it + 1);
    const double theta = A.th_is_theta ? thv : div_fast(thv, dp);                    // PA:615-621
    double s[4], lap[4], l1, l2;
    s[0] = dp - rdp;
    s[1] = theta - rth;
    s[2] = wv;
    s[3] = phiv - rphi;
    laplace_level(t, w, scr, A.nu_ratio, s, v1, v2, lap, l1, l2);
    if (act) {
      A.tout.dp[em + q] = lap[0]; A.tout.th[em + q] = lap[1];
      A.tout.w[ei + q] = lap[2]; A.tout.phi[ei + q] = lap[3];
      A.tout.v[ev + k * 32 + p] = l1; A.tout.v[ev + k * 32 + 16 + p] = l2;
    }
  }
}

// pass 2 (VT:134-164 + PA:657-663): DSS(tin) * rspheremp -> laplacian again -> * (-nu) -> tout
template <bool TENSOR>
__global__ void __launch_bounds__(HV_THREADS, HV_MINB) hv_laplace2_kernel(const HvArgs A)
{
  extern __shared__ double smem[];
  const int nlev = A.prm.nlev, nlevp = nlev + 1, ie = A.elist ? A.elist[blockIdx.x] : (int)blockIdx.x, tid = threadIdx.x;
  const int p = tid & 15, kslot = tid >> 4;
  constexpr int KS = HV_THREADS / 16;
  double *scr = smem + kslot * (HV_SLOTS * 16);
  Pt t; PtW w;
  pt_init(t, A.geo, ie, p, A.prm.rearth_inv);
  ptw_init(w, t, A.geo, ie, A.prm.rearth_inv);
  const int *src = A.geo.dss_src + ((size_t)ie * 16 + p) * 3;
  const int s0 = src[0], s1 = src[1], s2 = src[2];
  const double rs = A.geo.rspheremp[(size_t)ie * 16 + p];
  const size_t sm_ = 16 * (size_t)nlev, si = 16 * (size_t)nlevp, sv = 32 * (size_t)nlev;
  const size_t em = ie * sm_, ei = ie * si, ev = ie * sv;
  if (tid < 5) {
    if (tid == 0) prefetch_l2(A.tin.v + ev, (unsigned)(sv * 8));
    if (tid == 1) prefetch_l2(A.tin.dp + em, (unsigned)(sm_ * 8));
    if (tid == 2) prefetch_l2(A.tin.th + em, (unsigned)(sm_ * 8));
    if (tid == 3) prefetch_l2(A.tin.w + ei, (unsigned)(si * 8));
    if (tid == 4) prefetch_l2(A.tin.phi + ei, (unsigned)(si * 8));
  }
  double tv[4] = {0, 0, 0, 0}, s2c[6] = {0, 0, 0, 0, 0, 0};
  if (TENSOR) {
#pragma unroll
    for (int q = 0; q < 4; q++) tv[q] = A.geo.tensorvisc[(size_t)ie * 64 + q * 16 + p];
#pragma unroll
    for (int q = 0; q < 6; q++) s2c[q] = A.geo.vec_sph2cart[(size_t)ie * 96 + q * 16 + p];
  }
  const int niter = (nlev + KS - 1) / KS;
  double n_s[6];
  auto gather = [&](int it_) {
    const int kk_ = it_ * KS + kslot;
    const int k_ = kk_ < nlev ? kk_ : nlev - 1, q_ = k_ * 16 + p;
    const size_t om = (size_t)k_ * 16, ov_ = (size_t)k_ * 32;
    n_s[0] = rs * dss_sum(A.tin.dp, sm_, om, A.tin.dp[em + q_], s0, s1, s2);
    n_s[1] = rs * dss_sum(A.tin.th, sm_, om, A.tin.th[em + q_], s0, s1, s2);
    n_s[2] = rs * dss_sum(A.tin.w, si, om, A.tin.w[ei + q_], s0, s1, s2);
    n_s[3] = rs * dss_sum(A.tin.phi, si, om, A.tin.phi[ei + q_], s0, s1, s2);
    n_s[4] = rs * dss_sum(A.tin.v, sv, ov_, A.tin.v[ev + ov_ + p], s0, s1, s2);
    n_s[5] = rs * dss_sum(A.tin.v, sv, ov_ + 16, A.tin.v[ev + ov_ + 16 + p], s0, s1, s2);
  };
  gather(0);
  for (int it = 0; it < niter; it++) {
    const int kk = it * KS + kslot;
    const bool act = kk < nlev;
    const int k = act ? kk : nlev - 1, q = k * 16 + p;
    const size_t ov = (size_t)k * 32;
    double s[4], lap[4], l1, l2;
    s[0] = n_s[0]; s[1] = n_s[1]; s[2] = n_s[2]; s[3] = n_s[3];
    const double v1 = n_s[4], v2 = n_s[5];
    if (it + 1 < niter) gather(it + 1);        // next level's DSS sums are requested before this level's laplacians
    if (TENSOR) {                              // VT:145-163 with var_coef: tensor on the second laplacian only
      laplace_scalars<4, true>(t, w, tv, scr, s, lap);
      vlaplace_cartesian(t, w, tv, s2c, scr, v1, v2, l1, l2);
    } else {
      laplace_level(t, w, scr, A.nu_ratio, s, v1, v2, lap, l1, l2);
    }
    if (act) {
      A.tout.dp[em + q] = -A.prm.nu_p * lap[0]; A.tout.th[em + q] = -A.prm.nu * lap[1];
      A.tout.w[ei + q] = -A.prm.nu * lap[2]; A.tout.phi[ei + q] = -A.prm.nu_s * lap[3];
      A.tout.v[ev + ov + p] = -A.prm.nu * l1; A.tout.v[ev + ov + 16 + p] = -A.prm.nu * l2;
    }
  }
}

// pass 3 (PA:674-758, 767-774): DSS(tin) * dt * rspheremp added to the state
__global__ void __launch_bounds__(HV_THREADS, DSS_MINB) hv_update_kernel(const HvArgs A)
{
  const int nlev = A.prm.nlev, nlevp = nlev + 1, ie = A.elist ? A.elist[blockIdx.x] : (int)blockIdx.x, tid = threadIdx.x;
  const int p = tid & 15, kslot = tid >> 4;
  constexpr int KS = HV_THREADS / 16;
  const int *src = A.geo.dss_src + ((size_t)ie * 16 + p) * 3;
  const int s0 = src[0], s1 = src[1], s2 = src[2];
  const double rs = A.geo.rspheremp[(size_t)ie * 16 + p], dt = A.dt;
  const size_t sm_ = 16 * (size_t)nlev, si = 16 * (size_t)nlevp, sv = 32 * (size_t)nlev;
  const size_t em = ie * sm_, ei = ie * si, ev = ie * sv;
#pragma unroll DSS_UNROLL_N
  for (int k = kslot; k < nlev; k += KS) {
    const int q = k * 16 + p;
    const size_t om = (size_t)k * 16, ov = (size_t)k * 32;
    const double tdp = dt * dss_sum(A.tin.dp, sm_, om, DSS_OWN(A.tin.dp[em + q]), s0, s1, s2) * rs;
    const double tth = dt * dss_sum(A.tin.th, sm_, om, DSS_OWN(A.tin.th[em + q]), s0, s1, s2) * rs;
    const double tw = dt * dss_sum(A.tin.w, si, om, DSS_OWN(A.tin.w[ei + q]), s0, s1, s2) * rs;
    const double tph = dt * dss_sum(A.tin.phi, si, om, DSS_OWN(A.tin.phi[ei + q]), s0, s1, s2) * rs;
    const double tv1 = dt * dss_sum(A.tin.v, sv, ov, DSS_OWN(A.tin.v[ev + ov + p]), s0, s1, s2) * rs;
    const double tv2 = dt * dss_sum(A.tin.v, sv, ov + 16, DSS_OWN(A.tin.v[ev + ov + 16 + p]), s0, s1, s2) * rs;
    const double dp_old = A.st.dp[em + q];
    const double theta_old = A.th_is_theta ? A.st.th[em + q] : div_fast(A.st.th[em + q], dp_old);
    // PA:629-647 subtracts the reference state before the laplacians and adds it back afterwards; the state itself
    // differs from (x - r) + r by at most one rounding of x, far below the parity budget, so the three reference
    // fields are not read again here (27.8 KB per element and subcycle at L72)
    const double dp_new = dp_old + tdp;
    const double theta_new = theta_old + tth;
    const double v1 = A.st.v[ev + ov + p] + tv1, v2 = A.st.v[ev + ov + 16 + p] + tv2;
    A.st.v[ev + ov + p] = v1; A.st.v[ev + ov + 16 + p] = v2;
    A.st.w[ei + q] = A.st.w[ei + q] + tw;
    A.st.dp[em + q] = dp_new;
    A.st.phi[ei + q] = A.st.phi[ei + q] + tph;
    A.st.th[em + q] = A.last_subcycle ? theta_new * dp_new : theta_new;
    if (A.last_subcycle && k == nlev - 1) {
      const double g1 = A.geo.gradphis[(size_t)ie * 32 + p], g2 = A.geo.gradphis[(size_t)ie * 32 + 16 + p];
      A.st.w[ei + (size_t)nlev * 16 + p] = (v1 * g1 + v2 * g2) / G_GRAV;
    }
  }
}

// The sponge works on the top nlev_tom levels only (3 to 6 of them): a CTA takes as many elements as its 16
// half-warps can hold at one level per half-warp (nlev_tom = 6: two elements; one element if nlev_tom > 8).
__host__ __device__ inline int sponge_elems_per_cta(int ntom) { return ntom > 0 && ntom <= HV_THREADS / 32 ? (HV_THREADS / 16) / ntom : 1; }
// this half-warp's element: position in the launch's element list, clamped; `mine` = it owns real work
__device__ __forceinline__ int sponge_element(const HvArgs &A, int kslot, int ntom, int &k0, bool &mine)
{
  const int epc = sponge_elems_per_cta(ntom), nb = A.elist ? A.ecount : A.prm.nelem;
  int el = 0;
  k0 = kslot; mine = true;
  if (epc > 1) { el = kslot / ntom; k0 = kslot - el * ntom; mine = el < epc; if (!mine) { el = epc - 1; k0 = ntom - 1; } }
  int idx = blockIdx.x * epc + el;
  if (idx >= nb) { idx = nb - 1; mine = false; }
  return A.elist ? A.elist[idx] : idx;
}

// sponge pass A (PA:784-800): xfac * laplacian of the top nlev_tom levels of the state -> tout
__global__ void __launch_bounds__(HV_THREADS, 2) hv_sponge_laplace_kernel(const HvArgs A)
{
  extern __shared__ double smem[];
  const int nlev = A.prm.nlev, nlevp = nlev + 1, ntom = A.prm.nlev_tom, tid = threadIdx.x;
  const int p = tid & 15, kslot = tid >> 4;
  constexpr int KS = HV_THREADS / 16;
  int kfirst; bool mine;
  const int ie = sponge_element(A, kslot, ntom, kfirst, mine);
  const int kstep = sponge_elems_per_cta(ntom) > 1 ? ntom : KS;      // several elements per CTA: one level per half-warp
  double *scr = smem + kslot * (HV_SLOTS * 16);
  Pt t; PtW w;
  pt_init(t, A.geo, ie, p, A.prm.rearth_inv);
  ptw_init(w, t, A.geo, ie, A.prm.rearth_inv);
  const size_t em = (size_t)ie * 16 * nlev, ei = (size_t)ie * 16 * nlevp, ev = (size_t)ie * 32 * nlev;
  const int niter = (ntom + kstep - 1) / kstep;
  for (int it = 0; it < niter; it++) {
    const int kk = it * kstep + kfirst;
    const bool act = mine && kk < ntom;
    const int k = kk < ntom ? kk : ntom - 1, q = k * 16 + p;
    double s[4], lap[4], l1, l2;
    s[0] = A.st.dp[em + q]; s[1] = A.st.th[em + q]; s[2] = A.st.w[ei + q]; s[3] = A.st.phi[ei + q];
    const double v1 = A.st.v[ev + k * 32 + p], v2 = A.st.v[ev + k * 32 + 16 + p];
    laplace_level(t, w, scr, 1.0, s, v1, v2, lap, l1, l2);
    if (act) {
      const double xfac = A.dt * c_nu_scale_top[k] * A.prm.nu_top;                    // PA:793
      A.tout.dp[em + q] = xfac * lap[0]; A.tout.th[em + q] = xfac * lap[1];
      A.tout.w[ei + q] = xfac * lap[2]; A.tout.phi[ei + q] = xfac * lap[3];
      A.tout.v[ev + k * 32 + p] = xfac * l1; A.tout.v[ev + k * 32 + 16 + p] = xfac * l2;
    }
  }
}

// sponge pass B (PA:820-852): state += DSS(tin) * rspheremp on the top nlev_tom levels
__global__ void __launch_bounds__(HV_THREADS) hv_sponge_update_kernel(const HvArgs A)
{
  const int nlev = A.prm.nlev, nlevp = nlev + 1, ntom = A.prm.nlev_tom, tid = threadIdx.x;
  const int p = tid & 15, kslot = tid >> 4;
  constexpr int KS = HV_THREADS / 16;
  int kfirst; bool mine;
  const int ie = sponge_element(A, kslot, ntom, kfirst, mine);
  if (!mine) return;
  const int kstep = sponge_elems_per_cta(ntom) > 1 ? ntom : KS;
  const int *src = A.geo.dss_src + ((size_t)ie * 16 + p) * 3;
  const int s0 = src[0], s1 = src[1], s2 = src[2];
  const double rs = A.geo.rspheremp[(size_t)ie * 16 + p];
  const size_t sm_ = 16 * (size_t)nlev, si = 16 * (size_t)nlevp, sv = 32 * (size_t)nlev;
  const size_t em = ie * sm_, ei = ie * si, ev = ie * sv;
  for (int k = kfirst; k < ntom; k += kstep) {
    const int q = k * 16 + p;
    const size_t om = (size_t)k * 16, ov = (size_t)k * 32;
    A.st.v[ev + ov + p] += dss_sum(A.tin.v, sv, ov, A.tin.v[ev + ov + p], s0, s1, s2) * rs;
    A.st.v[ev + ov + 16 + p] += dss_sum(A.tin.v, sv, ov + 16, A.tin.v[ev + ov + 16 + p], s0, s1, s2) * rs;
    A.st.w[ei + q] += dss_sum(A.tin.w, si, om, A.tin.w[ei + q], s0, s1, s2) * rs;
    A.st.dp[em + q] += dss_sum(A.tin.dp, sm_, om, A.tin.dp[em + q], s0, s1, s2) * rs;
    A.st.phi[ei + q] += dss_sum(A.tin.phi, si, om, A.tin.phi[ei + q], s0, s1, s2) * rs;
    A.st.th[em + q] += dss_sum(A.tin.th, sm_, om, A.tin.th[em + q], s0, s1, s2) * rs;
  }
}

void launch_hv(int which, const HvArgs &A, cudaStream_t st)
{
  const size_t sm = hv_smem_bytes();
  const int nb = A.elist ? A.ecount : A.prm.nelem;
  if (nb <= 0) return;
  switch (which) {
    case 0: hv_laplace1_kernel<<<nb, HV_THREADS, sm, st>>>(A); break;
    case 1:
      if (A.prm.hypervis_scaling_on) hv_laplace2_kernel<true><<<nb, HV_THREADS, sm, st>>>(A);
      else hv_laplace2_kernel<false><<<nb, HV_THREADS, sm, st>>>(A);
      break;
    case 2: hv_update_kernel<<<nb, HV_THREADS, 0, st>>>(A); break;
    case 3: { const int epc = sponge_elems_per_cta(A.prm.nlev_tom);
              hv_sponge_laplace_kernel<<<(nb + epc - 1) / epc, HV_THREADS, sm, st>>>(A); break; }
    case 4: { const int epc = sponge_elems_per_cta(A.prm.nlev_tom);
              hv_sponge_update_kernel<<<(nb + epc - 1) / epc, HV_THREADS, 0, st>>>(A); break; }
  }
}

// ==========================================================================================
// Small element-wise kernels of prim_advance_exp (PA:140-143, 197-208, 313-326)
// ==========================================================================================

__global__ void __launch_bounds__(DSS_THREADS) combo_kernel(const ComboArgs A)
{
  const int nlev = A.prm.nlev, nlevp = nlev + 1, ie = blockIdx.x, tid = threadIdx.x;
  if (A.kind == 9) { combo9_element(A.a, A.b, A.c, A.geo, A.prm, ie, tid); return; }
  const int p = tid & 15, kslot = tid >> 4;
  constexpr int KS = DSS_THREADS / 16;
  const size_t em = (size_t)ie * 16 * nlev, ei = (size_t)ie * 16 * nlevp, ev = (size_t)ie * 32 * nlev;
  for (int k = kslot; k < nlevp; k += KS) {                                         // tstep_type 5, PA:197-208
    const int q = k * 16 + p;
    A.a.w[ei + q] = (5 * A.a.w[ei + q] - A.b.w[ei + q]) / 4;
    if (k < nlev) {
      const size_t qv = ev + k * 32 + p;
      A.a.v[qv] = (5 * A.a.v[qv] - A.b.v[qv]) / 4;
      A.a.v[qv + 16] = (5 * A.a.v[qv + 16] - A.b.v[qv + 16]) / 4;
      A.a.th[em + q] = (5 * A.a.th[em + q] - A.b.th[em + q]) / 4;
      A.a.dp[em + q] = (5 * A.a.dp[em + q] - A.b.dp[em + q]) / 4;
      A.a.phi[ei + q] = (5 * A.a.phi[ei + q] - A.b.phi[ei + q]) / 4;
    }
  }
}

void launch_combo(const ComboArgs &A, cudaStream_t st)
{
  combo_kernel<<<A.prm.nelem, DSS_THREADS, 0, st>>>(A);
}

// applyCAMforcing_dynamics (PA:517-541), standalone HOMME variant (FM(:,:,3,:) forces w_i): one CTA per
// element.  fm is derived%FM(np,np,3,nlev,ie), fvtheta derived%FVTheta(np,np,nlev,ie), fphi FPHI(np,np,nlevp,ie).
__global__ void __launch_bounds__(256) forcing_kernel(StatePtrs s, const double *fm, const double *fvtheta,
                                                      const double *fphi, Geo geo, double dt, int nlev)
{
  const int ie = blockIdx.x, p = threadIdx.x & 15, nlevp = nlev + 1;
  const size_t em = (size_t)ie * 16 * nlev, ei = (size_t)ie * 16 * nlevp, ev = (size_t)ie * 32 * nlev;
  const double *f = fm + (size_t)ie * 48 * nlev;
  for (int k = threadIdx.x >> 4; k < nlev; k += 16) {
    const int q = k * 16 + p;
    s.th[em + q] = s.th[em + q] + dt * fvtheta[em + q];
    s.phi[ei + q] = s.phi[ei + q] + dt * fphi[ei + q];
    const double v1 = s.v[ev + k * 32 + p] + dt * f[k * 48 + p];
    const double v2 = s.v[ev + k * 32 + 16 + p] + dt * f[k * 48 + 16 + p];
    s.v[ev + k * 32 + p] = v1;
    s.v[ev + k * 32 + 16 + p] = v2;
    s.w[ei + q] = s.w[ei + q] + dt * f[k * 48 + 32 + p];
    if (k == nlev - 1)
      s.w[ei + (size_t)nlev * 16 + p] = (v1 * geo.gradphis[(size_t)ie * 32 + p] + v2 * geo.gradphis[(size_t)ie * 32 + 16 + p]) / G_GRAV;
  }
}
void launch_forcing(const StatePtrs &s, const double *fm, const double *fvtheta, const double *fphi, const Geo &geo,
                    double dt, int nelem, int nlev, cudaStream_t st)
{
  forcing_kernel<<<nelem, 256, 0, st>>>(s, fm, fvtheta, fphi, geo, dt, nlev);
}

// dss_c: direct stiffness summation of an arbitrary field f[ie][layer][16] (edgeVpack / bndry_exchangeV / edgeVunpack
// of edge_mod.F90:500-629,862-944 as a node sum): gather with the same table and order as the step's kernels
__global__ void __launch_bounds__(256) dss_generic_kernel(const double *f, double *out, const int *dss_src, int nl)
{
  const int ie = blockIdx.x, p = threadIdx.x & 15;
  const int *src = dss_src + ((size_t)ie * 16 + p) * 3;
  const int s0 = src[0], s1 = src[1], s2 = src[2];
  for (int l = threadIdx.x >> 4; l < nl; l += 16) {
    double s = f[((size_t)ie * nl + l) * 16 + p];
    if (s0 >= 0) s = s + f[((size_t)(s0 >> 4) * nl + l) * 16 + (s0 & 15)];
    if (s1 >= 0) s = s + f[((size_t)(s1 >> 4) * nl + l) * 16 + (s1 & 15)];
    if (s2 >= 0) s = s + f[((size_t)(s2 >> 4) * nl + l) * 16 + (s2 & 15)];
    out[((size_t)ie * nl + l) * 16 + p] = s;
  }
}
__global__ void __launch_bounds__(256) halo_generic_kernel(double *f, const int *idx, int nent, int nl, double *buf, int unpack)
{
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)nent * nl) return;
  const int ent = (int)(i / nl), l = (int)(i - (long)ent * nl), ep = idx[ent];
  double *a = f + ((size_t)(ep >> 4) * nl + l) * 16 + (ep & 15);
  if (unpack) *a = buf[i]; else buf[i] = *a;
}
void launch_dss_generic(const double *f, double *out, const int *dss_src, int nelem, int nl, cudaStream_t st)
{
  dss_generic_kernel<<<nelem, 256, 0, st>>>(f, out, dss_src, nl);
}
void launch_halo_generic(double *f, const int *idx, int nent, int nl, double *buf, int unpack, cudaStream_t st)
{
  const long n = (long)nent * nl;
  if (n > 0) halo_generic_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(f, idx, nent, nl, buf, unpack);
}

// ps_v = hyai(1) ps0 + sum_k dp3d (vertremap_mod.F90:79-84) and its inverse on the reference levels,
// dp3d(k) = (hyai(k+1)-hyai(k)) ps0 + (hybi(k+1)-hybi(k)) ps_v (prim_driver.cpp:31-57); one thread per column
__global__ void ps_from_dp_kernel(const double *dp, double *ps, int nelem, int nlev, double ptop)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nelem * 16) return;
  const int ie = i >> 4, p = i & 15;
  double s = ptop;
  for (int k = 0; k < nlev; k++) s = s + dp[(size_t)ie * 16 * nlev + k * 16 + p];
  ps[i] = s;
}
__global__ void dp_from_ps_kernel(double *dp, const double *ps, int nelem, int nlev, double ps0)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nelem * 16) return;
  const int ie = i >> 4, p = i & 15;
  const double psv = ps[i];
  for (int k = 0; k < nlev; k++)
    dp[(size_t)ie * 16 * nlev + k * 16 + p] = (c_hyai[k + 1] - c_hyai[k]) * ps0 + (c_hybi[k + 1] - c_hybi[k]) * psv;
}
void launch_ps_from_dp(const double *dp, double *ps, int nelem, int nlev, double ptop, cudaStream_t st)
{
  ps_from_dp_kernel<<<(nelem * 16 + 255) / 256, 256, 0, st>>>(dp, ps, nelem, nlev, ptop);
}
void launch_dp_from_ps(double *dp, const double *ps, int nelem, int nlev, double ps0, cudaStream_t st)
{
  dp_from_ps_kernel<<<(nelem * 16 + 255) / 256, 256, 0, st>>>(dp, ps, nelem, nlev, ps0);
}

// FP64 FMA peak probe: 8 independent DFMA chains per thread, 1024 threads per CTA, 2 CTAs per SM
__global__ void __launch_bounds__(1024) fp64_peak_kernel(double *sink, int n)
{
  double a[8];
#pragma unroll
  for (int i = 0; i < 8; i++) a[i] = 1.0 + 1e-3 * (threadIdx.x + i);
  const double b = 0.99999999, c = 1e-9;
  for (int it = 0; it < n; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = fma(a[i], b, c);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += a[i];
  if (s == 123.456) sink[0] = s;
}
double launch_fp64_peak(cudaStream_t st)
{
  static double *sink = nullptr;
  if (!sink) cudaMalloc(&sink, sizeof(double));
  int dev = 0, nsm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  const int n = 20000, blocks = 2 * nsm;
  fp64_peak_kernel<<<blocks, 1024, 0, st>>>(sink, n);
  return 2.0 * 8.0 * (double)n * 1024.0 * blocks;
}

// PA:140-143  w_i(nlevp,n0) = (v(nlev) . gradphis)/g
__global__ void wbc_kernel(StatePtrs s, Geo geo, int nelem, int nlev)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nelem * 16) return;
  const int ie = i >> 4, p = i & 15;
  const double *v = s.v + (size_t)ie * 32 * nlev + (nlev - 1) * 32;
  s.w[(size_t)ie * 16 * (nlev + 1) + nlev * 16 + p] =
      (v[p] * geo.gradphis[(size_t)ie * 32 + p] + v[16 + p] * geo.gradphis[(size_t)ie * 32 + 16 + p]) / G_GRAV;
}

void launch_wbc(const StatePtrs &s, const Geo &geo, int nelem, int nlev, cudaStream_t st)
{
  wbc_kernel<<<(nelem * 16 + 255) / 256, 256, 0, st>>>(s, geo, nelem, nlev);
}

// ==========================================================================================
// Halo pack / unpack for the multi-GPU DSS (edgeVpack / edgeVunpack, edge_mod.F90:500-629,
// 862-944; bndry_exchangeV, bndry_mod.F90:51-169).  The DSS kernels pull their neighbours'
// pre-DSS values straight out of the `tens` arrays; a neighbour owned by another rank is a ghost
// element slot behind the local ones, and only its shared edge/corner points are ever filled.
// One entry = one (element, point) on the partition boundary; its message is all layers of the
// exchange: levels k < nk of v1, v2, dp3d, vtheta_dp, w_i, phinh_i (+ w_i(nlevp) if wbot).
// ==========================================================================================
__device__ __forceinline__ double *halo_addr(const StatePtrs &t, int ep, int layer, int nk, int nlev)
{
  const size_t e = (size_t)(ep >> 4);
  const int p = ep & 15, nlevp = nlev + 1;
  if (layer == 6 * nk) return t.w + e * 16 * nlevp + (size_t)nlev * 16 + p;
  const int f = layer / nk, k = layer - f * nk;
  switch (f) {
    case 0: return t.v + e * 32 * nlev + (size_t)k * 32 + p;
    case 1: return t.v + e * 32 * nlev + (size_t)k * 32 + 16 + p;
    case 2: return t.dp + e * 16 * nlev + (size_t)k * 16 + p;
    case 3: return t.th + e * 16 * nlev + (size_t)k * 16 + p;
    case 4: return t.w + e * 16 * nlevp + (size_t)k * 16 + p;
    default: return t.phi + e * 16 * nlevp + (size_t)k * 16 + p;
  }
}

__global__ void __launch_bounds__(256) halo_pack_kernel(StatePtrs t, const int *idx, int nent, int nk, int wbot,
                                                        int nlev, double *buf)
{
  const int nlyr = 6 * nk + wbot;
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)nent * nlyr) return;
  const int ent = (int)(i / nlyr), layer = (int)(i - (long)ent * nlyr);
  buf[i] = *halo_addr(t, idx[ent], layer, nk, nlev);
}

__global__ void __launch_bounds__(256) halo_unpack_kernel(StatePtrs t, const int *idx, int nent, int nk, int wbot,
                                                          int nlev, const double *buf)
{
  const int nlyr = 6 * nk + wbot;
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)nent * nlyr) return;
  const int ent = (int)(i / nlyr), layer = (int)(i - (long)ent * nlyr);
  *halo_addr(t, idx[ent], layer, nk, nlev) = buf[i];
}

void launch_halo_unpack(const StatePtrs &t, const int *idx, int nent, int nk, int wbot, int nlev,
                        const double *buf, cudaStream_t st)
{
  const long n = (long)nent * (6 * nk + wbot);
  if (n > 0) halo_unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(t, idx, nent, nk, wbot, nlev, buf);
}

void launch_halo_pack(const StatePtrs &t, const int *idx, int nent, int nk, int wbot, int nlev, double *buf,
                      cudaStream_t st)
{
  const long n = (long)nent * (6 * nk + wbot);
  if (n > 0) halo_pack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(t, idx, nent, nk, wbot, nlev, buf);
}
// ---------------------------------------------------------------- global diagnostics
// prim_energy_halftimes (prim_state_mod.F90:822-939) per element, then global_integral
// (global_norms_mod.F90:47-90): J(ie) = sum_{j,i} spheremp * h in the reference's point order, and
// an order-independent fixed-point sum over elements (thetal_reprosum.h).
//   q = 0: ps = ptop + sum_k dp3d        1: KE = sum_k dp (u^2+v^2)/2 + dp (w_k^2+w_k+1^2)/4
//   q = 2: PE = sum_k dp (phi_k+phi_k+1)/2   3: IE = sum_k cp vtheta_dp exner + dphi pnh, + ptop phi_top
__global__ void __launch_bounds__(256) diag_element_kernel(const DiagArgs A)
{
  __shared__ double red[NDIAG][16][16];
  __shared__ double col[NDIAG][16];
  const int nlev = A.prm.nlev, ie = blockIdx.x, tid = threadIdx.x, p = tid & 15, ks = tid >> 4;
  const size_t em = (size_t)ie * 16 * nlev, ei = (size_t)ie * 16 * (nlev + 1), ev = (size_t)ie * 32 * nlev;
  double s_ps = 0, s_ke = 0, s_pe = 0, s_ie = 0;
  for (int k = ks; k < nlev; k += 16) {
    const int q = k * 16 + p;
    const double dp = A.st.dp[em + q], th = A.st.th[em + q];
    const double phi0 = A.st.phi[ei + q], phi1 = A.st.phi[ei + q + 16];
    const double w0 = A.st.w[ei + q], w1 = A.st.w[ei + q + 16];
    const double v1 = A.st.v[ev + k * 32 + p], v2 = A.st.v[ev + k * 32 + 16 + p];
    const double dphi = phi1 - phi0;
    const double poe = div_fast(RGAS * th, -dphi);
    const double pnh = eos_pnh(poe), exner = div_fast(pnh, poe);
    s_ps += dp;
    s_ke += 0.5 * (v1 * v1 + v2 * v2) * dp + 0.25 * (w0 * w0 + w1 * w1) * dp;
    s_pe += 0.5 * (phi0 + phi1) * dp;
    s_ie += CP * th * exner + dphi * pnh;
  }
  red[0][ks][p] = s_ps; red[1][ks][p] = s_ke; red[2][ks][p] = s_pe; red[3][ks][p] = s_ie;
  __syncthreads();
  if (tid < NDIAG * 16) {
    const int q = tid >> 4;
    double s = 0;
    for (int j = 0; j < 16; j++) s += red[q][j][p];
    if (q == 0) s += A.prm.ptop;
    if (q == 3) s += A.prm.ptop * A.st.phi[ei + p];
    col[q][p] = s;
  }
  __syncthreads();
  if (tid < NDIAG) {
    double J = 0;
    for (int j = 0; j < 16; j++) J += A.geo.spheremp[(size_t)ie * 16 + j] * col[tid][j];
    A.part[(size_t)ie * NDIAG + tid] = J;
    atomicMax(A.maxbits + tid, (unsigned long long)__double_as_longlong(fabs(J)));
  } else if (tid == 32) {
    double lo = col[0][0], hi = col[0][0];
    for (int j = 1; j < 16; j++) { lo = fmin(lo, col[0][j]); hi = fmax(hi, col[0][j]); }
    atomicMin(A.ps_minmax, (unsigned long long)__double_as_longlong(lo));      // ps > 0: bit order = value order
    atomicMax(A.ps_minmax + 1, (unsigned long long)__double_as_longlong(hi));
  }
}

__global__ void __launch_bounds__(256) diag_accumulate_kernel(const double *part, int nelem,
                                                              const unsigned long long *maxbits,
                                                              unsigned long long *limbs)
{
  const int i = blockIdx.x * 256 + threadIdx.x;
  for (int q = 0; q < NDIAG; q++) {
    long long l[thetal_reprosum::NLIMB] = {0, 0, 0};
    if (i < nelem) thetal_reprosum::split(part[(size_t)i * NDIAG + q], thetal_reprosum::exponent_of_bits(maxbits[q]), l);
    for (int j = 0; j < thetal_reprosum::NLIMB; j++) {
      long long x = l[j];
      for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
      if ((threadIdx.x & 31) == 0 && x != 0) atomicAdd(limbs + q * thetal_reprosum::NLIMB + j, (unsigned long long)x);
    }
  }
}

// prim_printstate's field extrema (prim_state_mod.F90:219-330): min and max over the rank of u, v, w_i, theta =
// vtheta_dp/dp3d, T = theta*exner, dz = -dphi/g and dp3d.  Order-preserving integer keys make atomicMin/atomicMax
// (and an NCCL min/max of integers) exact for either sign.
__device__ __forceinline__ unsigned long long ordered_key(double x)
{
  const unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__global__ void __launch_bounds__(256) extrema_kernel(const DiagArgs A, unsigned long long *keys)
{
  __shared__ double smin[NEXTREMA][8], smax[NEXTREMA][8];
  const int nlev = A.prm.nlev, ie = blockIdx.x, tid = threadIdx.x, p = tid & 15, ks = tid >> 4;
  const size_t em = (size_t)ie * 16 * nlev, ei = (size_t)ie * 16 * (nlev + 1), ev = (size_t)ie * 32 * nlev;
  double lo[NEXTREMA], hi[NEXTREMA];
#pragma unroll
  for (int f = 0; f < NEXTREMA; f++) { lo[f] = 1e300; hi[f] = -1e300; }
  for (int k = ks; k < nlev; k += 16) {
    const int q = k * 16 + p;
    const double dp = A.st.dp[em + q], th = A.st.th[em + q];
    const double dphi = A.st.phi[ei + q + 16] - A.st.phi[ei + q];
    const double poe = div_fast(RGAS * th, -dphi);
    const double exner = div_fast(eos_pnh(poe), poe), theta = div_fast(th, dp);
    const double val[NEXTREMA] = {A.st.v[ev + k * 32 + p], A.st.v[ev + k * 32 + 16 + p], A.st.w[ei + q], theta,
                                  theta * exner, -dphi / G_GRAV, dp};
#pragma unroll
    for (int f = 0; f < NEXTREMA; f++) { lo[f] = fmin(lo[f], val[f]); hi[f] = fmax(hi[f], val[f]); }
  }
  if (ks == 0) { const double wb = A.st.w[ei + (size_t)nlev * 16 + p]; lo[2] = fmin(lo[2], wb); hi[2] = fmax(hi[2], wb); }
#pragma unroll
  for (int f = 0; f < NEXTREMA; f++) {
    for (int o = 16; o > 0; o >>= 1) {
      lo[f] = fmin(lo[f], __shfl_xor_sync(0xffffffffu, lo[f], o));
      hi[f] = fmax(hi[f], __shfl_xor_sync(0xffffffffu, hi[f], o));
    }
    if ((tid & 31) == 0) { smin[f][tid >> 5] = lo[f]; smax[f][tid >> 5] = hi[f]; }
  }
  __syncthreads();
  if (tid < NEXTREMA) {
    double a = smin[tid][0], b = smax[tid][0];
    for (int i = 1; i < 8; i++) { a = fmin(a, smin[tid][i]); b = fmax(b, smax[tid][i]); }
    atomicMin(keys + 2 * tid, ordered_key(a));
    atomicMax(keys + 2 * tid + 1, ordered_key(b));
  }
}
void launch_extrema(const DiagArgs &A, unsigned long long *keys, cudaStream_t st)
{
  extrema_kernel<<<A.prm.nelem, 256, 0, st>>>(A, keys);
}

void launch_diag_elements(const DiagArgs &A, cudaStream_t st)
{
  diag_element_kernel<<<A.prm.nelem, 256, 0, st>>>(A);
}
void launch_diag_accumulate(const double *part, int nelem, const unsigned long long *maxbits,
                            unsigned long long *limbs, cudaStream_t st)
{
  diag_accumulate_kernel<<<(nelem + 255) / 256, 256, 0, st>>>(part, nelem, maxbits, limbs);
}

// Running sums in the reference's k order for one column (pointers already offset by the point index;
// rows are 16 doubles apart).  Eight addends are loaded before the dependent additions start, so one
// shared-memory latency is paid per eight levels instead of per level.
__device__ __forceinline__ double scan_up(const double *in, double *out, double s, int nlev)
{ // out(k+1) = out(k) + in(k), k = 1..nlev
  for (int k0 = 1; k0 <= nlev; k0 += 8) {
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = (k0 + i <= nlev) ? in[(k0 + i) * 16] : 0.0;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (k0 + i <= nlev) { s = s + x[i]; out[(k0 + i + 1) * 16] = s; }
  }
  return s;
}
template <int SIGN>
__device__ __forceinline__ double scan_down(const double *in, double *out, double s, int nlev)
{ // out(k) = out(k+1) +/- in(k), k = nlev..1
  for (int k0 = nlev; k0 >= 1; k0 -= 8) {
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = (k0 - i >= 1) ? in[(k0 - i) * 16] : 0.0;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (k0 - i >= 1) { s = SIGN > 0 ? s + x[i] : s - x[i]; out[(k0 - i) * 16] = s; }
  }
  return s;
}

// ---------------------------------------------------------------- vertical remap
// vertical_remap, vertremap_mod.F90:23-171 (rsplit > 0, qsize = 0) with remap_Q_ppm
// (vertremap_base.F90:514-646; compute_ppm_grids :650-673, compute_ppm :680-723, integrate_parabola
// :730-737, binary_search :744-765, linextrap :768-800).  One CTA per element, 16 columns x 16 level
// slots; the running sums (pressures, masses, phi) are kept in the reference's k order by one thread
// per column, everything else is parallel over levels.  All arrays sit in shared memory as [row][16].
__global__ void __launch_bounds__(256, 2) vertical_remap_kernel(const RemapArgs A)
{
  extern __shared__ double rsm[];
  const int nlev = A.prm.nlev, nlevp = nlev + 1, R = nlev + 4;
  const int ie = blockIdx.x, tid = threadIdx.x, p = tid & 15, ks = tid >> 4;
  double *dpo = rsm, *pio = dpo + R * 16, *pin = pio + R * 16, *z2 = pin + R * 16, *ao = z2 + R * 16;
  double *masso = ao + R * 16, *dma = masso + R * 16, *ai = dma + R * 16, *qm = ai + R * 16;
  double *phr = qm + R * 16, *phs = phr + R * 16, *ps = phs + R * 16;
  int *kid = (int *)(ps + 16);
#define DPO(i) dpo[((i) + 1) * 16 + p]
#define AO(i) ao[((i) + 1) * 16 + p]
#define ROW(a, k) a[(k) * 16 + p]
  const size_t em = (size_t)ie * 16 * nlev, ei = (size_t)ie * 16 * nlevp, ev = (size_t)ie * 32 * nlev;
  double *gdp = A.st.dp + em, *gth = A.st.th + em, *gphi = A.st.phi + ei, *gw = A.st.w + ei, *gv = A.st.v + ev;
  const double ptop = A.prm.ptop;
  if (tid < 4) {      // the columns the field loop reads later, towards L2 now
    if (tid == 0) prefetch_l2(gth, (unsigned)(16 * nlev * 8));
    if (tid == 1) prefetch_l2(gphi, (unsigned)(16 * nlevp * 8));
    if (tid == 2) prefetch_l2(gv, (unsigned)(32 * nlev * 8));
    if (tid == 3) prefetch_l2(gw, (unsigned)(16 * nlevp * 8));
  }
  int bad = 0;
  for (int k = 1 + ks; k <= nlev; k += 16) {
    const double d = gdp[(k - 1) * 16 + p];
    DPO(k) = d;
    if (d < 0) bad = 1;
  }
  if (bad) atomicOr(A.err, ERR_REMAP_NEGATIVE_LAYER);
  if (__syncthreads_or(bad)) return;                          // the reference aborts before remapping (:87-103)
  // phi_from_eos pieces: hydrostatic pressure scan, parallel power terms, phi scan from the bottom
  auto phi_ref_scan = [&](const double *dprow1 /* rows 1..nlev of a [row][16] array */) {
    if (tid < 16) {                                           // p_i(k+1) = p_i(k) + dp(k), eos.F90:287-290
      ROW(masso, 1) = ptop;
      scan_up(dprow1 + p, masso + p, ptop, nlev);
    }
    __syncthreads();
    for (int k = 1 + ks; k <= nlev; k += 16) {                // eos.F90:300
      const double pm = (ROW(masso, k + 1) + ROW(masso, k)) / 2;
      ROW(ai, k) = div_fast(RGAS * gth[(k - 1) * 16 + p] * thetal_pow::pow_pos(div_fast(pm, P0), C_KM1), P0);
    }
    __syncthreads();
    if (tid < 16) {
      const double ph = A.geo.phis[(size_t)ie * 16 + p];
      ROW(phr, nlevp) = ph;
      scan_down<1>(ai + p, phr + p, ph, nlev);
    }
    __syncthreads();
  };
  phi_ref_scan(dpo + 16);                                     // DPO(k) lives in row k+1
  // remove the hydrostatic phi, reference layer thickness, interface pressures of both grids
  for (int k = 1 + ks; k <= nlevp; k += 16) ROW(phs, k) = gphi[(k - 1) * 16 + p] - ROW(phr, k);
  if (tid < 16) {                                             // old interface pressures; their last value is sum(dp3d)
    ROW(pio, 1) = 0;
    const double s = scan_up(dpo + 16 + p, pio + p, 0., nlev);
    ps[p] = ptop + s;                                         // vertremap_mod.F90:76-77
  }
  __syncthreads();
  for (int k = 1 + ks; k <= nlev; k += 16)                    // dp on reference levels, :79-80 (kept in qm for now)
    ROW(z2, k) = (c_hyai[k] - c_hyai[k - 1]) * A.prm.ps0 + (c_hybi[k] - c_hybi[k - 1]) * ps[p];
  __syncthreads();
  if (tid < 16) {
    ROW(pin, 1) = 0;
    scan_up(z2 + p, pin + p, 0., nlev);
    const double b = ROW(pio, nlev + 1);
    ROW(pio, nlev + 2) = b + 1.;
    ROW(pin, nlev + 1) = b;
    DPO(0) = DPO(1); DPO(-1) = DPO(2); DPO(nlev + 1) = DPO(nlev); DPO(nlev + 2) = DPO(nlev - 1);
  }
  __syncthreads();
  // new layer thickness to its own array (z2 is about to be reused): dpn lives in `dma` rows until the
  // field loop needs dma, so copy it to global dp3d at the end from registers instead
  double dpn_r[8];                                            // this thread's levels (nlev <= 128)
#pragma unroll
  for (int j = 0; j < 8; j++) { const int k = 1 + ks + 16 * j; dpn_r[j] = k <= nlev ? ROW(z2, k) : 1.0; }
  __syncthreads();
  for (int k = 1 + ks; k <= nlev; k += 16) {                  // old cell holding the new interface k+1
    const double pv = ROW(pin, k + 1);
    int kk = k;
    if (ROW(pio, kk) <= pv) {
      while (ROW(pio, kk) <= pv) kk = kk + 1;
      kk = kk - 1;
    } else {
      int lo = 1, hi = kk;
      while (hi > lo + 1) {
        kk = (lo + hi) / 2;
        if (ROW(pio, kk) > pv) hi = kk; else lo = kk;
      }
      kk = lo;
    }
    if (kk == nlev + 1) kk = nlev;
    kid[k * 16 + p] = kk;
    ROW(z2, k) = div_fast(pv - (ROW(pio, kk) + ROW(pio, kk + 1)) * 0.5, DPO(kk));
  }
  __syncthreads();
  for (int f = 0; f < 5; f++) {
    for (int k = 1 + ks; k <= nlev; k += 16) {                // ttmp(:,:,k,f), vertremap_mod.F90:112-121
      const int q = (k - 1) * 16 + p;
      double m;
      if (f == 0) m = gv[(k - 1) * 32 + p] * DPO(k);
      else if (f == 1) m = gv[(k - 1) * 32 + 16 + p] * DPO(k);
      else if (f == 2) m = gth[q];
      else if (f == 3) m = ROW(phs, k + 1) - ROW(phs, k);
      else m = gw[q + 16] - gw[q];
      ROW(qm, k) = m;
      AO(k) = div_fast(m, DPO(k));
    }
    __syncthreads();
    if (tid < 16) {
      ROW(masso, 1) = 0.;
      scan_up(qm + p, masso + p, 0., nlev);
      if (A.remap_alg == 10 || A.remap_alg == 11) {
        const bool limit = A.remap_alg == 10;
        double lo = AO(1), hi = AO(1);
        if (limit) for (int k = 2; k <= nlev; k++) { lo = fmin(lo, AO(k)); hi = fmax(hi, AO(k)); }
        auto linextrap = [&](double dx1, double dx2, double dx3, double dx4, double y1, double y2, double &y3, double &y4) {
          const double den = (dx1 + dx2) / 2;
          double num = den + (dx2 + dx3) / 2;
          double a = num / den;
          y3 = (1 - a) * y1 + a * y2;
          num = num + (dx3 + dx4) / 2;
          a = num / den;
          y4 = (1 - a) * y1 + a * y2;
          if (limit) { y3 = fmax(lo, fmin(hi, y3)); y4 = fmax(lo, fmin(hi, y4)); }
        };
        double y3, y4;
        linextrap(DPO(2), DPO(1), DPO(0), DPO(-1), AO(2), AO(1), y3, y4);
        AO(0) = y3; AO(-1) = y4;
        linextrap(DPO(nlev - 1), DPO(nlev), DPO(nlev + 1), DPO(nlev + 2), AO(nlev - 1), AO(nlev), y3, y4);
        AO(nlev + 1) = y3; AO(nlev + 2) = y4;
      } else {
        AO(0) = AO(1); AO(-1) = AO(1); AO(nlev + 1) = AO(nlev); AO(nlev + 2) = AO(nlev);
      }
    }
    __syncthreads();
    for (int j = ks; j <= nlev + 1; j += 16) {                // compute_ppm stage 1 (grid terms :660-664)
      const double dx1 = div_fast(DPO(j), DPO(j - 1) + DPO(j) + DPO(j + 1));
      const double dx2 = div_fast(2. * DPO(j - 1) + DPO(j), DPO(j + 1) + DPO(j));
      const double dx3 = div_fast(DPO(j) + 2. * DPO(j + 1), DPO(j - 1) + DPO(j));
      const double dr = AO(j + 1) - AO(j), dl = AO(j) - AO(j - 1);
      const double da = dx1 * (dx2 * dr + dx3 * dl);
      double m = copysign(fmin(fabs(da), fmin(2. * fabs(dl), 2. * fabs(dr))), da);
      if (dr * dl <= 0.) m = 0.;
      ROW(dma, j) = m;
    }
    __syncthreads();
    for (int j = ks; j <= nlev; j += 16) {                    // stage 2 (grid terms :667-675)
      const double d0 = DPO(j), d1 = DPO(j + 1), dm = DPO(j - 1), d2 = DPO(j + 2);
      const double dx4 = div_fast(d0, d0 + d1);
      const double dx5 = div_fast(1., ((dm + d0) + d1) + d2);
      const double dx6 = div_fast(2. * d1 * d0, d0 + d1);
      const double dx7 = div_fast(dm + d0, 2. * d0 + d1);
      const double dx8 = div_fast(d2 + d1, 2. * d1 + d0);
      const double dx9 = div_fast(d0 * (dm + d0), 2. * d0 + d1);
      const double dx10 = div_fast(d1 * (d1 + d2), d0 + 2. * d1);
      const double dr = AO(j + 1) - AO(j);
      ROW(ai, j) = AO(j) + dx4 * dr + dx5 * (dx6 * (dx7 - dx8) * dr - dx9 * ROW(dma, j + 1) + dx10 * ROW(dma, j));
    }
    __syncthreads();
    for (int k = 1 + ks; k <= nlev; k += 16) {                // stage 3 for the cell the interface sits in + integral
      const int kk = kid[k * 16 + p];
      const double a = AO(kk);
      double al = ROW(ai, kk - 1), ar = ROW(ai, kk);
      if ((ar - a) * (a - al) <= 0.) { al = a; ar = a; }
      if ((ar - al) * (a - (al + ar) / 2.) > (ar - al) * (ar - al) / 6.) al = 3. * a - 2. * ar;
      if ((ar - al) * (a - (al + ar) / 2.) < -((ar - al) * (ar - al)) / 6.) ar = 3. * a - 2. * al;
      const double c0 = 1.5 * a - (al + ar) / 4., c1 = ar - al, c2 = 3. * (-2. * a + (al + ar));
      const double x1 = -0.5, x2 = ROW(z2, k);
      const double par = c0 * (x2 - x1) + c1 * (x2 * x2 - x1 * x1) / 0.2e1 + c2 * (x2 * x2 * x2 - x1 * x1 * x1) / 0.3e1;
      ROW(qm, k) = ROW(masso, kk) + par * DPO(kk);            // massn2(k)
    }
    if (tid < 16) ROW(qm, 0) = 0.;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int k = 1 + ks + 16 * j;
      if (k <= nlev) {
        const double out = ROW(qm, k) - ROW(qm, k - 1);
        const int q = (k - 1) * 16 + p;
        if (f == 0) gv[(k - 1) * 32 + p] = div_fast(out, dpn_r[j]);
        else if (f == 1) gv[(k - 1) * 32 + 16 + p] = div_fast(out, dpn_r[j]);
        else if (f == 2) gth[q] = out;
        else if (f == 3) ROW(pin, k) = out;                   // pin/pio are free after the search
        else ROW(pio, k) = out;
      }
    }
    __syncthreads();
  }
  if (tid < 16) scan_down<-1>(pin + p, phs + p, ROW(phs, nlevp), nlev);                    // vertremap_mod.F90:131-136
  else if (tid < 32) scan_down<-1>(pio + p, gw + p - 16, gw[nlev * 16 + p], nlev);         // out(k) -> gw[(k-1)*16+p]
  // reference-level thickness into a [row][16] array for the second phi_from_eos, and into dp3d (:167)
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const int k = 1 + ks + 16 * j;
    if (k <= nlev) { ROW(dma, k) = dpn_r[j]; gdp[(k - 1) * 16 + p] = dpn_r[j]; }
  }
  __syncthreads();
  phi_ref_scan(dma);                                          // uses the remapped vtheta_dp (:139)
  for (int k = 1 + ks; k <= nlevp; k += 16) gphi[(k - 1) * 16 + p] = ROW(phs, k) + ROW(phr, k);
  if (tid < 16) {                                             // w b.c. from the remapped wind, :145-146
    const double *gp = A.geo.gradphis + (size_t)ie * 32;
    gw[nlev * 16 + p] = (gv[(nlev - 1) * 32 + p] * gp[p] + gv[(nlev - 1) * 32 + 16 + p] * gp[16 + p]) / G_GRAV;
  }
#undef DPO
#undef AO
#undef ROW
}

size_t vertical_remap_smem_bytes(int nlev)
{
  const size_t R = nlev + 4;
  return (11 * R * 16 + 16) * sizeof(double) + R * 16 * sizeof(int);
}
void launch_vertical_remap(const RemapArgs &A, cudaStream_t st)
{
  static size_t configured = 0;
  const size_t bytes = vertical_remap_smem_bytes(A.prm.nlev);
  if (bytes > configured) {
    cudaFuncSetAttribute(vertical_remap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    configured = bytes;
  }
  vertical_remap_kernel<<<A.prm.nelem, 256, bytes, st>>>(A);
}

}  // namespace thetal
