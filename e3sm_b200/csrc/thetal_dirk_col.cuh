// compute_stage_value_dirk (IM:60-364), column-parallel form: TPC lanes of one warp own one column
// (TPC = 16: a half-warp per column, 256 threads per element; TPC = 32: a warp per column, 512 threads),
// lane j holds the ITEMS consecutive levels k = j*ITEMS .. j*ITEMS+ITEMS-1 in registers.  Everything a
// level needs from its vertical neighbours is either in the same thread or one warp shuffle away, so the
// Newton iteration has no shared-memory traffic and ONE block barrier (the per-element convergence test
// of IM:330-335).  The tridiagonal system J x = -F (IM:414-433, 576-599) is solved without a serial
// sweep over the column:
//   1. every lane eliminates its own ITEMS rows (partition method: forward elimination of the
//      sub-diagonal, backward elimination of the super-diagonal) down to one equation that couples the
//      last unknown of the lane to the last unknowns of the two neighbouring lanes;
//   2. the TPC reduced equations of a column are solved by parallel cyclic reduction with warp shuffles
//      (log2 TPC rounds, rows kept normalised so a round costs one reciprocal);
//   3. every lane back-substitutes its interior unknowns (two FMAs per row).
// Levels beyond nlev (the column is padded to TPC*ITEMS rows) are identity rows behind the last real
// row, whose super-diagonal is zero, so they never couple to anything.
// The EOS of IM:261 and IM:324 is evaluated without pow wherever that is exact to rounding (eos_level_*_step below:
// binomial series in the relative change of dphi; the hydrostatic guess returns the hydrostatic pressure).
// With DSS = true the kernel also finishes the CAAR stage before it (node sum, rspheremp, surface fix-up,
// limiter_dp3d_k: PA:1741-1772, 1812-1931) on the staged columns and writes all six prognostics of np1.
// Included by thetal_kernels.cu inside namespace thetal (uses its constants and helpers).
#pragma once

constexpr unsigned FULLMASK = 0xffffffffu;

// sum of v over the lanes below / above this one inside the TPC-lane column group
template <int TPC>
__device__ __forceinline__ double excl_sum_below(double v, int j)
{
  double e = __shfl_up_sync(FULLMASK, v, 1, TPC);
  if (j == 0) e = 0.0;
#pragma unroll
  for (int s = 1; s < TPC; s <<= 1) {
    const double t = __shfl_up_sync(FULLMASK, e, s, TPC);
    if (j >= s) e = e + t;
  }
  return e;
}
template <int TPC>
__device__ __forceinline__ double excl_sum_above(double v, int j)
{
  double e = __shfl_down_sync(FULLMASK, v, 1, TPC);
  if (j == TPC - 1) e = 0.0;
#pragma unroll
  for (int s = 1; s < TPC; s <<= 1) {
    const double t = __shfl_down_sync(FULLMASK, e, s, TPC);
    if (j + s < TPC) e = e + t;
  }
  return e;
}
template <int TPC>
__device__ __forceinline__ double group_max(double v)
{
#pragma unroll
  for (int s = 1; s < TPC; s <<= 1) v = fmax(v, __shfl_xor_sync(FULLMASK, v, s, TPC));
  return v;
}
template <int TPC>
__device__ __forceinline__ double group_min(double v)
{
#pragma unroll
  for (int s = 1; s < TPC; s <<= 1) v = fmin(v, __shfl_xor_sync(FULLMASK, v, s, TPC));
  return v;
}

// IM:42-53 at one interface from the two adjacent levels
__device__ __forceinline__ double gwphis_of(double dk, double dm, double v1k, double v1m, double v2k, double v2m,
                                            double g1, double g2, double hyb)
{
  const double r = rcp_fast(dk + dm);
  const double n1 = dk * v1k + dm * v1m, n2 = dk * v2k + dm * v2m;
  double q1 = n1 * r, q2 = n2 * r;
  q1 = fma(fma(-q1, dk + dm, n1), r, q1);
  q2 = fma(fma(-q2, dk + dm, n2), r, q2);
  return (q1 * g1 + q2 * g2) * hyb;
}

// EOS of one level from dphi (EOS:175-183): pnh and c = pnh/dphi (IM:417) with one reciprocal
__device__ __forceinline__ void eos_level(double rth, double dphi, double &pnh, double &c)
{
  const double r = rcp_fast(dphi);                       // 1/dphi (< 0)
  const double q0 = -(rth * r);
  const double poe = fma(fma(q0, dphi, rth), -r, q0); // Rgas*vtheta_dp/(-dphi), <= 1 ulp
  pnh = P0 * thetal_pow::pow_pos(div_p0(poe), C_GAMMA);
  c = pnh * r;
}

// The same EOS after a Newton update that moved dphi by a small relative amount: with rth fixed,
// pnh_new = pnh_old (dphi_old/dphi_new)^gamma, and (1 + e)^gamma is its binomial series in e = dphi_old/dphi_new - 1.
// For |e| < 2^-4 the terms beyond degree 11 sum to less than 5e-18: the result differs from the full evaluation by
// rounding only; for |e| < 2^-10 the same holds for degree 4 (what remains is below 1e-17).  Thirteen or six FMAs
// instead of the ~65 of pow_pos.  (C2: every update after the hydrostatic guess is below 2^-10.)
#ifndef DIRK_EOS_SERIES
#define DIRK_EOS_SERIES 1
#endif
#ifndef DIRK_EOS_SERIES_MAX
#define DIRK_EOS_SERIES_MAX 0.0625
#endif
constexpr double EOS_SERIES_MAX = DIRK_EOS_SERIES_MAX, EOS_SERIES_TINY = 0.0009765625;
__device__ __forceinline__ void eos_level_tiny_step(double e, double r, double &pnh, double &c)
{
  constexpr double G1 = 1.0 / (1.0 - KAPPA), G2 = G1 * (G1 - 1) / 2, G3 = G2 * (G1 - 2) / 3, G4 = G3 * (G1 - 3) / 4;
  double s = G4;
  s = fma(s, e, G3); s = fma(s, e, G2); s = fma(s, e, G1);
  pnh = fma(pnh * e, s, pnh);
  c = pnh * r;
}
__device__ __forceinline__ void eos_level_small_step(double e, double r, double &pnh, double &c)
{
  constexpr double G1 = 1.0 / (1.0 - KAPPA), G2 = G1 * (G1 - 1) / 2, G3 = G2 * (G1 - 2) / 3, G4 = G3 * (G1 - 3) / 4,
                   G5 = G4 * (G1 - 4) / 5, G6 = G5 * (G1 - 5) / 6, G7 = G6 * (G1 - 6) / 7, G8 = G7 * (G1 - 7) / 8,
                   G9 = G8 * (G1 - 8) / 9, G10 = G9 * (G1 - 9) / 10, G11 = G10 * (G1 - 10) / 11;
  double s = G11;
  s = fma(s, e, G10); s = fma(s, e, G9); s = fma(s, e, G8); s = fma(s, e, G7); s = fma(s, e, G6);
  s = fma(s, e, G5); s = fma(s, e, G4); s = fma(s, e, G3); s = fma(s, e, G2); s = fma(s, e, G1);
  pnh = fma(pnh * e, s, pnh);
  c = pnh * r;
}

template <int ITEMS, int TPC>
struct DirkColCfg {
  static constexpr int NT = 16 * TPC;
  // Shared memory: (a) per-thread private arrays w0, dphi0, phi0 ([slot][thread], conflict-free);
  // (b) the staging area of ONE element's columns, transposed to [field][point][level] with an odd row
  // pitch SS: filled by coalesced 8-byte cp.async (a warp copies 256 contiguous bytes of the [level][point]
  // global layout per instruction), read by each thread along its own column without bank conflicts
  // (ITEMS = 5: lane stride 5 doubles).  Six fields: th, dp, w, phi, v1, v2.
  static constexpr int SS = (ITEMS * TPC + 1) | 1;              // >= nlevp, odd
  static constexpr size_t priv_doubles = 3 * ITEMS * NT, stage_doubles = 6 * 16 * (size_t)SS;
  static constexpr size_t smem_bytes = sizeof(double) * (priv_doubles + stage_doubles);
  static constexpr int minb = (2 * (smem_bytes + 2304) <= 232448) ? 2 : 1;   // CTAs per SM that fit in 227 KB
};

__device__ __forceinline__ void cp_async8(double *dst, const double *src)
{
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(int *dst, const int *src)
{
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// slow path of limiter_dp3d_k (PA:1859-1897) for one column held in the staging area (unit stride over the levels);
// the same arithmetic as limiter_columns, which works on the finished state in global memory
__device__ void limiter_column_staged(double *dp, double *th, int nlev, double spheremp, double dp3d_thresh)
{
  double mn = dp[0] - dp3d_thresh * c_dp0[0];
  for (int k = 1; k < nlev; k++) mn = fmin(mn, dp[k] - dp3d_thresh * c_dp0[k]);
  for (int k = 0; k < nlev; k++) th[k] = th[k] / dp[k];
  if (mn < 0) {
    double mass = 0, mass_new = 0;
    for (int k = 0; k < nlev; k++) mass = mass + (dp[k] - dp3d_thresh * c_dp0[k]) * spheremp;
    const double sgn = (mass < 0) ? -1.0 : 1.0;
    for (int k = 0; k < nlev; k++) {
      double Q = (dp[k] - dp3d_thresh * c_dp0[k]) * spheremp;
      if (mass < 0) Q = -Q;
      if (!(Q < 0)) mass_new = mass_new + Q;
    }
    for (int k = 0; k < nlev; k++) {
      double Q = (dp[k] - dp3d_thresh * c_dp0[k]) * spheremp;
      if (mass < 0) Q = -Q;
      if (Q < 0) Q = 0;
      if (mass_new > 0) Q = Q * (fabs(mass) / mass_new);
      if (sgn < 0) Q = -Q;
      dp[k] = Q / spheremp + dp3d_thresh * c_dp0[k];
    }
  }
  for (int k = 0; k < nlev; k++) th[k] = th[k] * dp[k];
}

// DSS = true: the element's columns are staged from the pre-DSS buffer A.tens of the CAAR stage before this solve, and
// the kernel finishes that stage itself (node sum over the neighbours, x rspheremp, surface fix-up, limiter_dp3d_k:
// what caar_dss_post_kernel does) in the staging area before the Newton iteration starts; it then writes all six
// prognostics of np1.  One pass over the state per stage instead of two (A.tens is read-only here, np1 is only
// written by the CTA that owns the element, so there is no ordering between elements).
template <int ITEMS, int TPC, bool DSS = false>
__global__ void __launch_bounds__(16 * TPC, (DirkColCfg<ITEMS, TPC>::minb)) dirk_col_kernel(const DirkArgs A)
{
  constexpr int NT = 16 * TPC, NW = NT / 32, M = ITEMS;
  static_assert(ITEMS >= 2, "the partition solve needs two rows per lane");
  extern __shared__ double smem[];
  __shared__ double red[3][NW];
  __shared__ int sflag;
  // DSS: the gather table entries and 1/mass of this thread's point, fetched with the element's columns (one element ahead)
  __shared__ int s_src[DSS ? 3 : 1][DSS ? NT : 1];
  __shared__ double s_rs[DSS ? NT : 1];
  const int nlev = A.prm.nlev, nlevp = nlev + 1, n16 = 16 * nlev, nelem = A.prm.nelem;
  const int tid = threadIdx.x, j = tid % TPC, p = tid / TPC, k0 = j * ITEMS;
  double *s_w0 = smem + tid, *s_dphi0 = smem + ITEMS * NT + tid, *s_phi0 = smem + 2 * ITEMS * NT + tid;
  const double g = G_GRAV, dt2 = A.dt2, dtg = dt2 * g, ptop = A.prm.ptop;
  const int km = min(max(k0 - 1, 0), nlev - 1);
  using Cfg = DirkColCfg<ITEMS, TPC>;
  constexpr int SS = Cfg::SS;
  double *st_th = smem + Cfg::priv_doubles, *st_dp = st_th + 16 * SS, *st_w = st_dp + 16 * SS;
  double *st_phi = st_w + 16 * SS, *st_v1 = st_phi + 16 * SS, *st_v2 = st_v1 + 16 * SS;
  // The CTAs are persistent (grid = CTAs resident on the chip, elements strided over them).  While one
  // element iterates, the np1 columns of the CTA's next element travel into the staging area with
  // cp.async, and the n0/nm1 columns the explicit terms need go to L2.
  auto stage_state = [&](const StatePtrs &S, int e) {
    const double *eth = S.th + (size_t)e * n16, *edp = S.dp + (size_t)e * n16, *ev = S.v + (size_t)e * 2 * n16;
    const double *ew = S.w + (size_t)e * 16 * nlevp, *ephi = S.phi + (size_t)e * 16 * nlevp;
    for (int q = tid; q < n16; q += NT) {
      const int o = (q & 15) * SS + (q >> 4);
      cp_async8(st_th + o, eth + q);
      cp_async8(st_dp + o, edp + q);
    }
    for (int q = tid; q < 16 * nlevp; q += NT) {
      const int o = (q & 15) * SS + (q >> 4);
      cp_async8(st_w + o, ew + q);
      cp_async8(st_phi + o, ephi + q);
    }
    for (int q = tid; q < 2 * n16; q += NT) {
      const int o = (q & 15) * SS + (q >> 5);
      cp_async8(((q & 16) ? st_v2 : st_v1) + o, ev + q);
    }
    if constexpr (DSS) {
      const int *src = A.geo.dss_src + ((size_t)e * 16 + (tid & 15)) * 3;
      cp_async4(&s_src[0][tid], src); cp_async4(&s_src[1][tid], src + 1); cp_async4(&s_src[2][tid], src + 2);
      cp_async8(&s_rs[tid], A.geo.rspheremp + (size_t)e * 16 + (tid & 15));
    }
    cp_async_commit();
  };
  auto prefetch_explicit = [&](int e) {
    if (A.ew_n0 && A.ew_nm1) {
      if (tid < 4 && (A.alphadt_n0 != 0.0 || A.alphadt_nm1 != 0.0)) {
        const double *f = tid == 0 ? A.ew_n0 : tid == 1 ? A.ephi_n0 : tid == 2 ? A.ew_nm1 : A.ephi_nm1;
        prefetch_l2(f + (size_t)e * n16, 16u * nlev * 8u);
      }
    } else if (tid < 10) {
      const StatePtrs &S = tid < 5 ? A.n0 : A.nm1;
      const bool on = tid < 5 ? A.alphadt_n0 != 0.0 : A.alphadt_nm1 != 0.0;
      const int f = tid % 5;
      const unsigned bv = 32u * nlev * 8u, bm = 16u * nlev * 8u, bi = 16u * nlevp * 8u;
      if (on) {
        if (f == 0) prefetch_l2(S.dp + (size_t)e * n16, bm);
        if (f == 1) prefetch_l2(S.th + (size_t)e * n16, bm);
        if (f == 2) prefetch_l2(S.w + (size_t)e * 16 * nlevp, bi);
        if (f == 3) prefetch_l2(S.phi + (size_t)e * 16 * nlevp, bi);
        if (f == 4) prefetch_l2(S.v + (size_t)e * 2 * n16, bv);
      }
    }
  };
  const double *cth = st_th + p * SS, *cdp = st_dp + p * SS, *cw = st_w + p * SS, *cphi = st_phi + p * SS;
  const double *cv1 = st_v1 + p * SS, *cv2 = st_v2 + p * SS;       // this thread's column in the staging area

  // Elements beyond the first gridDim.x are handed out dynamically (an atomic counter per launch): elements differ in
  // their Newton work, and a static stride leaves some CTAs with one element more than others.
  __shared__ int s_next;
  int ie = blockIdx.x, ie_next = nelem;
  if (blockIdx.x == 0 && tid == 0) *A.work_other = 0;         // the counter of the previous launch, for the next one
  const StatePtrs &SRC = DSS ? A.tens : A.np1;
  if (ie < nelem) { stage_state(SRC, ie); prefetch_explicit(ie); }
  int bad = 0;
  for (; ie < nelem; ie = ie_next) {
  double *gw = A.np1.w + (size_t)ie * 16 * nlevp, *gphi = A.np1.phi + (size_t)ie * 16 * nlevp;
  const double gp1 = A.geo.gradphis[(size_t)ie * 32 + p], gp2 = A.geo.gradphis[(size_t)ie * 32 + 16 + p];
  if (tid == 0) { sflag = 0; s_next = (int)gridDim.x + atomicAdd(A.work, 1); }

  double w[ITEMS], dphi[ITEMS], rdpi[ITEMS], rth[ITEMS];
  double dp[ITEMS], w0[ITEMS], phi0[ITEMS], v1[ITEMS], v2[ITEMS];
  cp_async_wait_all();
  __syncthreads();
  ie_next = s_next;
  if (ie_next < nelem) prefetch_explicit(ie_next);
  if constexpr (DSS) {
    // ---- DSS and post-processing of the CAAR stage (PA:1741-1772, 1812-1931), [level][point] over the threads ----
    const int pt = tid & 15;
    const int s0 = s_src[0][tid], s1 = s_src[1][tid], s2 = s_src[2][tid];
    const double rs = s_rs[tid];
    const size_t ev = 2 * (size_t)n16, em = (size_t)n16, ei = 16 * (size_t)nlevp;
    // surface data of the fix-up below: requested now so that the warp that owns level nlev does not wait for them
    const double g1 = A.geo.gradphis[(size_t)ie * 32 + pt], g2 = A.geo.gradphis[(size_t)ie * 32 + 16 + pt];
    const double phis_pt = A.np1.phi[(size_t)ie * ei + (size_t)nlev * 16 + pt];   // phinh_i(nlevp) = phis: every time level holds it
    double *ov = A.np1.v + (size_t)ie * ev, *odp = A.np1.dp + (size_t)ie * em, *oth = A.np1.th + (size_t)ie * em;
    int lim_dp = 0, lim_th = 0;
    for (int k = tid >> 4; k < nlev; k += NT / 16) {
      const int o = pt * SS + k;
      const size_t offv = (size_t)k * 32, offm = (size_t)k * 16;
      const double v1s = dss_sum(A.tens.v, ev, offv, st_v1[o], s0, s1, s2);
      const double v2s = dss_sum(A.tens.v, ev, offv + 16, st_v2[o], s0, s1, s2);
      const double dps = dss_sum(A.tens.dp, em, offm, st_dp[o], s0, s1, s2);
      const double ths = dss_sum(A.tens.th, em, offm, st_th[o], s0, s1, s2);
      const double ws = dss_sum(A.tens.w, ei, offm, st_w[o], s0, s1, s2);
      const double phs = dss_sum(A.tens.phi, ei, offm, st_phi[o], s0, s1, s2);
      double v1 = rs * v1s, v2 = rs * v2s;
      const double dpk = rs * dps, thk = rs * ths;
      st_dp[o] = dpk; st_th[o] = thk; st_w[o] = rs * ws; st_phi[o] = rs * phs;
      if (k == nlev - 1) {                                                            // PA:1751-1772
        const size_t ob = (size_t)nlev * 16;
        double wb = rs * dss_sum(A.tens.w, ei, ob, st_w[o + 1], s0, s1, s2);
        const double dpnh = 1 + (((v1 * g1 + v2 * g2) / G_GRAV - wb) / (G_GRAV + (g1 * g1 + g2 * g2) / (2 * G_GRAV))) / A.dss_dt2;
        wb = wb + A.dss_scale1 * A.dss_dt2 * G_GRAV * (dpnh - 1);
        v1 = v1 - A.dss_scale1 * A.dss_dt2 * (dpnh - 1) * g1 / 2;
        v2 = v2 - A.dss_scale1 * A.dss_dt2 * (dpnh - 1) * g2 / 2;
        st_w[o + 1] = wb;
        A.np1.w[(size_t)ie * ei + ob + pt] = wb;
        st_phi[o + 1] = phis_pt;
      }
      st_v1[o] = v1; st_v2[o] = v2;
      ov[offv + pt] = v1; ov[offv + 16 + pt] = v2;
      if (dpk < A.prm.dp3d_thresh * c_dp0[k]) lim_dp = 1;
      if (thk - A.prm.vtheta_thresh * dpk < 0) lim_th = 1;
    }
    // one barrier in the common case: it orders the staging-area writes before the reads below and tells whether
    // any point of the element needs limiter_dp3d_k (rare; block-uniform decisions)
    if (__syncthreads_or(A.dss_limiter ? (lim_dp | lim_th) : 0)) {
      lim_dp = __syncthreads_or(lim_dp);
      if (lim_dp) {
        if (tid < 16) limiter_column_staged(st_dp + tid * SS, st_th + tid * SS, nlev, A.geo.spheremp[(size_t)ie * 16 + tid],
                                            A.prm.dp3d_thresh);
        __syncthreads();
        lim_th = 0;
        for (int k = tid >> 4; k < nlev; k += NT / 16)
          if (st_th[pt * SS + k] - A.prm.vtheta_thresh * st_dp[pt * SS + k] < 0) lim_th = 1;
      }
      lim_th = __syncthreads_or(lim_th);
      if (lim_th) {                                                                   // PA:1903-1925
        for (int k = tid >> 4; k < nlev; k += NT / 16) {
          const int o = pt * SS + k;
          if (st_th[o] - A.prm.vtheta_thresh * st_dp[o] < 0) st_th[o] = A.prm.vtheta_thresh * st_dp[o];
        }
      }
      __syncthreads();
    }
    for (int k = tid >> 4; k < nlev; k += NT / 16) {
      const int o = pt * SS + k;
      odp[(size_t)k * 16 + pt] = st_dp[o]; oth[(size_t)k * 16 + pt] = st_th[o];
    }
  }
  // ---- the column slice, wmax (IM:168-170) ----
  double wmax = 0.0;
  const double dpm0 = cdp[km], v1m0 = cv1[km], v2m0 = cv2[km];   // level above the slice
  const double wbot = cw[nlev];
  const double phis = cphi[nlev];                             // phinh_i(nlevp) = phis, never changed
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const int k = k0 + i, kk = min(k, nlev - 1);
    const bool valid = k < nlev;
    const double th = cth[kk];
    dp[i] = cdp[kk]; phi0[i] = cphi[kk]; v1[i] = cv1[kk]; v2[i] = cv2[kk];
    if (valid && (th < 0 || dp[i] < 0)) bad = 1;
    const double dpi = (k > 0) ? (dp[i] + (i > 0 ? dp[i > 0 ? i - 1 : 0] : dpm0)) / 2 : dp[i];
    // padded rows (k >= nlev) get constants that make them identity rows with a zero right-hand side
    // without any select inside the Newton loop: 1/dp_i = 0, w = 0, w0 = dt2*g, dphi0 = dphi = -g
    rdpi[i] = valid ? div_fast(1.0, dpi) : 0.0;
    rth[i] = RGAS * th;
    w0[i] = valid ? cw[kk] : 0.0;
    wmax = fmax(wmax, fabs(w0[i]));
  }
  if (j == TPC - 1) wmax = fmax(wmax, fabs(wbot));
  wmax = group_max<32>(wmax);
  if ((tid & 31) == 0) red[2][tid >> 5] = wmax;
  __syncthreads();
  wmax = red[2][0];
#pragma unroll
  for (int i = 1; i < NW; i++) wmax = fmax(wmax, red[2][i]);
  wmax = fmax(1.0, wmax);

  // ---- explicit stiff terms of earlier stages (IM:176-205) ----
  const bool expl = A.alphadt_n0 != 0.0 || A.alphadt_nm1 != 0.0;
  if (expl && A.ew_n0 && A.ew_nm1) {
    // saved by the CAAR launches that evaluated state(n0) and state(nm1) (dpnh_dp_i - 1 and g w - gwphis): four
    // interface fields pass through the staging area (every thread has read its np1 slice before the barrier above)
    const double *e0 = A.ew_n0 + (size_t)ie * n16, *p0 = A.ephi_n0 + (size_t)ie * n16;
    const double *e1 = A.ew_nm1 + (size_t)ie * n16, *p1 = A.ephi_nm1 + (size_t)ie * n16;
    for (int q = tid; q < n16; q += NT) {
      const int o = (q & 15) * SS + (q >> 4);
      cp_async8(st_th + o, e0 + q); cp_async8(st_dp + o, p0 + q);
      cp_async8(st_w + o, e1 + q);  cp_async8(st_phi + o, p1 + q);
    }
    cp_async_commit();
    cp_async_wait_all();
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const int kk = min(k0 + i, nlev - 1);
      if (A.alphadt_n0 != 0.0) { w0[i] = w0[i] + A.alphadt_n0 * g * cth[kk]; phi0[i] = phi0[i] + A.alphadt_n0 * cdp[kk]; }
      if (A.alphadt_nm1 != 0.0) { w0[i] = w0[i] + A.alphadt_nm1 * g * cw[kk]; phi0[i] = phi0[i] + A.alphadt_nm1 * cphi[kk]; }
    }
  } else
  // from the states themselves: the n0 / nm1 columns of this element pass through the staging area
  for (int pass = 0; pass < 2; pass++) {
    const double dt3 = pass == 0 ? A.alphadt_n0 : A.alphadt_nm1;
    if (dt3 == 0.0) continue;
    if (pass == 1) __syncthreads();                    // pass 0 has been read by everybody
    stage_state(pass == 0 ? A.n0 : A.nm1, ie);
    cp_async_wait_all();
    __syncthreads();
    double tp[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const int k = k0 + i, kk = min(k, nlev - 1);
      const double dph = cphi[kk + 1] - cphi[kk], th = cth[kk], tdpk = cdp[kk];
      if (k < nlev && (th < 0 || tdpk < 0 || dph > 0)) bad = 1;
      tp[i] = eos_pnh(div_fast(RGAS * th, -dph));
    }
    const double tpm = __shfl_up_sync(FULLMASK, tp[ITEMS - 1], 1, TPC);
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const int k = k0 + i, kk = min(k, nlev - 1), kq = max(kk - 1, 0);
      const double dk = cdp[kk], dm = cdp[kq];
      double mu;
      if (k == 0) mu = div_fast(2 * (tp[i] - ptop), dk);
      else mu = div_fast(tp[i] - (i > 0 ? tp[i > 0 ? i - 1 : 0] : tpm), (dk + dm) / 2);
      w0[i] = w0[i] + dt3 * g * (mu - 1);
      const double gwh = (k == 0) ? 0.0 : gwphis_of(dk, dm, cv1[kk], cv1[kq], cv2[kk], cv2[kq], gp1, gp2, c_hybi[kk]);
      phi0[i] = phi0[i] + dt3 * g * cw[kk] - dt3 * gwh;
    }
  }
  // the staging area is consumed: fetch the next element of this CTA behind the Newton iterations
  if (ie_next < nelem) {
    if (expl) __syncthreads();
    stage_state(SRC, ie_next);
  }
  // ---- wphis(np1) term (IM:208-211) ----
  {
    double v1m = v1m0, v2m = v2m0, dm = dpm0;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const int k = k0 + i, kk = min(k, nlev - 1);
      const double v1k = v1[i], v2k = v2[i];
      const double gwh = (k == 0) ? 0.0 : gwphis_of(dp[i], dm, v1k, v1m, v2k, v2m, gp1, gp2, c_hybi[kk]);
      phi0[i] = phi0[i] - dt2 * gwh;
      v1m = v1k; v2m = v2k; dm = dp[i];
    }
  }
  // ---- hydrostatic initial guess phi_from_eos (EOS:286-302, IM:228-230) ----
  // running sums as (sum inside the lane) + (sum over the lanes above/below): the guess only starts Newton
  double pnh[ITEMS], cc[ITEMS];
  bool guess_kept;               // the hydrostatic guess was not clamped (IM:254-260 not taken)
  {
    double acc = 0.0, pin[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) { pin[i] = acc; acc = acc + ((k0 + i < nlev) ? dp[i] : 0.0); }
    const double ptl = ptop + excl_sum_below<TPC>(acc, j);          // pressure at the top interface of the slice
    double term[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const double pi_k = ptl + pin[i], pn = pi_k + dp[i];
      const double pmid = (pn + pi_k) / 2;
      const double t = div_p0(rth[i] * thetal_pow::pow_pos(div_p0(pmid), C_KM1));
      term[i] = (k0 + i < nlev) ? t : 0.0;
      pnh[i] = pmid;
    }
    // phi_k = phis + sum_{k' >= k} term_k'
    double suf[ITEMS];
    acc = 0.0;
#pragma unroll
    for (int i = ITEMS - 1; i >= 0; i--) { acc = acc + term[i]; suf[i] = acc; }
    const double below = phis + excl_sum_above<TPC>(acc, j);        // phi at the bottom interface of the slice
    double ph[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) ph[i] = below + suf[i];
    const double ph_next = __shfl_down_sync(FULLMASK, ph[0], 1, TPC);       // first level of the lane below
    const double p0_next = __shfl_down_sync(FULLMASK, phi0[0], 1, TPC);
    int nsafe = 0;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {                                               // IM:234-253
      const int k = k0 + i;
      const bool lastlev = (k == nlev - 1), inlane = (i + 1 < ITEMS);
      const double phb = lastlev ? phis : (inlane ? ph[inlane ? i + 1 : i] : ph_next);
      const double p0b = lastlev ? phis : (inlane ? phi0[inlane ? i + 1 : i] : p0_next);
      double d = phb - ph[i];
      s_dphi0[i * NT] = (k < nlev) ? p0b - phi0[i] : -g;
      if (k < nlev && d > -g) { d = -g; nsafe = 1; }
      dphi[i] = (k < nlev) ? d : -g;
    }
    if (nsafe) sflag = 1;
    __syncthreads();
    guess_kept = !sflag;
    if (sflag) {                                                                    // IM:254-260 (rare, whole element)
      acc = 0.0;
#pragma unroll
      for (int i = ITEMS - 1; i >= 0; i--) { acc = acc - ((k0 + i < nlev) ? dphi[i] : 0.0); suf[i] = acc; }
      const double below2 = phis + excl_sum_above<TPC>(acc, j);
#pragma unroll
      for (int i = 0; i < ITEMS; i++) ph[i] = below2 + suf[i];
    }
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      w[i] = (k0 + i < nlev) ? div_fast(ph[i] - phi0[i], dtg) : 0.0;
      s_w0[i * NT] = (k0 + i < nlev) ? w0[i] : dtg;
      s_phi0[i * NT] = phi0[i];
    }
  }
#if DIRK_EOS_SERIES
  if (guess_kept) {
    // IM:261 at the hydrostatic guess: dphi was built so that the EOS returns the hydrostatic mid-level pressure
    // (EOS:286-302 inverts EOS:175-183), so pnh = pmid up to the rounding of the interface sums (~1e-14); the Newton
    // iteration does not depend on where it starts
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      if (k0 + i < nlev && dphi[i] > 0) bad = 1;
      cc[i] = pnh[i] * rcp_fast(dphi[i]);
    }
  } else
#endif
  {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      if (k0 + i < nlev && dphi[i] > 0) bad = 1;
      eos_level(rth[i], dphi[i], pnh[i], cc[i]);                                    // IM:261
    }
  }

  // ---- Newton iteration (IM:268-336) ----
  int itercount = 0;
  double deltaerr = 0.0;
  const int maxiter = 20;
  const double ajac = dtg * dtg / (1 - KAPPA);                                      // IM:414
  while (itercount < maxiter) {
    // rows of J x = -F for this lane's levels (IM:414-433, 263-264)
    double sub[ITEMS], dg[ITEMS], sup[ITEMS], rhs[ITEMS];
    {
      double pm = __shfl_up_sync(FULLMASK, pnh[ITEMS - 1], 1, TPC);
      double cm = __shfl_up_sync(FULLMASK, cc[ITEMS - 1], 1, TPC);
#pragma unroll
      for (int i = 0; i < ITEMS; i++) {
        const bool first = (i == 0) && (j == 0);                  // level 1: IM:419-422, EOS:198
        const double bk = ajac * rdpi[i];
        const double mu = ((first ? 2.0 : 1.0) * (pnh[i] - (first ? ptop : pm))) * rdpi[i];
        const double Fn = w[i] - (s_w0[i * NT] + dtg * (mu - 1));
        const double lo = first ? 0.0 : bk * cm;
        const double up = (first ? 2.0 : 1.0) * (bk * cc[i]);
        sub[i] = lo;
        dg[i] = 1 - lo - up;
        // the super-diagonal of level nlev (IM:430-432 has none) multiplies the unknown of the padded row
        // behind it, which is exactly zero; only a column that fills every row needs the explicit zero
        sup[i] = (i == ITEMS - 1 && j == TPC - 1) ? 0.0 : up;
        rhs[i] = -Fn;
        pm = pnh[i]; cm = cc[i];
      }
    }
    // 1. forward elimination inside the lane: row i becomes  f_i s_prev + dg_i x_i + sup_i x_{i+1} = rhs_i
    double f[ITEMS], rinv[ITEMS - 1], gq[ITEMS - 1];
    f[0] = sub[0];
#pragma unroll
    for (int i = 1; i < M; i++) {
      rinv[i - 1] = rcp_fast(dg[i - 1]);
      const double l = sub[i] * rinv[i - 1];
      dg[i] = fma(-l, sup[i - 1], dg[i]);
      rhs[i] = fma(-l, rhs[i - 1], rhs[i]);
      f[i] = -(l * f[i - 1]);
    }
    //    backward elimination: rows i <= M-2 become  f_i s_prev + dg_i x_i + gq_i s_own = rhs_i
    gq[M - 2] = sup[M - 2];
#pragma unroll
    for (int i = M - 3; i >= 0; i--) {
      const double t = sup[i] * rinv[i + 1];
      f[i] = fma(-t, f[i + 1], f[i]);
      gq[i] = -(t * gq[i + 1]);
      rhs[i] = fma(-t, rhs[i + 1], rhs[i]);
    }
    // 2. reduced equation of the lane: the first row of the lane below removes x_{first of next lane}
    double Ar, Cr, Rr;
    {
      const double F0 = f[0] * rinv[0], G0 = gq[0] * rinv[0], R0 = rhs[0] * rinv[0];
      const double Fd = __shfl_down_sync(FULLMASK, F0, 1, TPC);
      const double Gd = __shfl_down_sync(FULLMASK, G0, 1, TPC);
      const double Rd = __shfl_down_sync(FULLMASK, R0, 1, TPC);
      const double cl = sup[M - 1];                   // zero in the last lane (last level or padding)
      const double D = fma(-cl, Fd, dg[M - 1]);
      const double rD = rcp_fast(D);
      Ar = f[M - 1] * rD;
      Cr = -(cl * Gd) * rD;
      Rr = fma(-cl, Rd, rhs[M - 1]) * rD;
    }
    //    parallel cyclic reduction over the TPC lanes of the column, rows normalised to unit diagonal
#pragma unroll
    for (int s = 1; s < TPC; s <<= 1) {
      // lanes j < s have Ar = 0 and lanes j >= TPC - s have Cr = 0 at this point (induction from A = 0 in
      // lane 0 and C = 0 in the last lane), so the values a shuffle returns for a missing partner (the
      // lane's own, finite) are multiplied by zero: no range selects
      const double Cu = __shfl_up_sync(FULLMASK, Cr, s, TPC), Ru = __shfl_up_sync(FULLMASK, Rr, s, TPC);
      const double Ad = __shfl_down_sync(FULLMASK, Ar, s, TPC), Rd = __shfl_down_sync(FULLMASK, Rr, s, TPC);
      const double Dn = fma(-Cr, Ad, fma(-Ar, Cu, 1.0));
      const double rD = rcp_fast(Dn);
      Rr = fma(-Cr, Rd, fma(-Ar, Ru, Rr)) * rD;
      if (2 * s < TPC) {            // the last round only needs the right-hand side
        const double Au = __shfl_up_sync(FULLMASK, Ar, s, TPC), Cd = __shfl_down_sync(FULLMASK, Cr, s, TPC);
        const double An = -(Ar * Au) * rD, Cn = -(Cr * Cd) * rD;
        Ar = An; Cr = Cn;
      }
    }
    // 3. back-substitution inside the lane
    double x[ITEMS];
    x[M - 1] = Rr;
    {
      const double sp = __shfl_up_sync(FULLMASK, Rr, 1, TPC);      // lane 0: f = 0
#pragma unroll
      for (int i = 0; i < M - 1; i++) x[i] = fma(-gq[i], Rr, fma(-f[i], sp, rhs[i])) * rinv[i];
    }
    // dphi update (IM:278-283), positivity check, w update (IM:320-322)
    // level nlev uses  dphi0 - dt2 g (w + x)  (IM:282): the same expression with w = x = 0 below it, which is
    // what the padded row behind it holds; the one slot with no row behind it gets explicit zeros
    double x_next = __shfl_down_sync(FULLMASK, x[0], 1, TPC);
    double w_next = __shfl_down_sync(FULLMASK, w[0], 1, TPC);
    if (j == TPC - 1) { x_next = 0.0; w_next = 0.0; }
    int anypos = 0;
    double xmax = 0.0;
#if DIRK_EOS_SERIES
    double dold[ITEMS];          // the dphi that pnh and cc were evaluated at
#pragma unroll
    for (int i = 0; i < ITEMS; i++) dold[i] = dphi[i];
#endif
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const bool inlane = (i + 1 < ITEMS);
      const double xb = inlane ? x[inlane ? i + 1 : i] : x_next, wb = inlane ? w[inlane ? i + 1 : i] : w_next;
      const double d = s_dphi0[i * NT] + dtg * ((wb - w[i]) + (xb - x[i]));
      dphi[i] = d;
      if (d >= 0) anypos = 1;
      xmax = fmax(xmax, fabs(x[i]));
    }
    double alpha = 1.0;
    if (__any_sync(FULLMASK, anypos)) {                                             // IM:285-319 (rare), per column
      const unsigned bal = __ballot_sync(FULLMASK, anypos);
      const unsigned gmask = (TPC == 32) ? FULLMASK : (0xffffu << ((tid & 16)));
      const bool colpos = (bal & gmask) != 0;
      double amin = 1.0;
#pragma unroll
      for (int i = 0; i < ITEMS; i++) {
        const int k = k0 + i;
        const bool inlane = (i + 1 < ITEMS);
        const double xb = inlane ? x[inlane ? i + 1 : i] : x_next, wb = inlane ? w[inlane ? i + 1 : i] : w_next;
        const double dx = xb - x[i], dw = wb - w[i];
        if (k < nlev && dx != 0) {
          const double alpha_k = -(s_dphi0[i * NT] + dtg * dw) / (dtg * dx);
          if (alpha_k >= 0) amin = fmin(amin, alpha_k);
        }
      }
      amin = group_min<TPC>(amin);
      if (colpos) {
        alpha = amin / 2;
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
          const int k = k0 + i;
          const bool inlane = (i + 1 < ITEMS);
          const double xb = inlane ? x[inlane ? i + 1 : i] : x_next, wb = inlane ? w[inlane ? i + 1 : i] : w_next;
          if (k < nlev) dphi[i] = s_dphi0[i * NT] + dtg * ((wb - w[i]) + alpha * (xb - x[i]));
        }
      }
    }
#pragma unroll
    for (int i = 0; i < ITEMS; i++) w[i] = w[i] + alpha * x[i];
    // per-element convergence test (IM:330-335): one barrier, double-buffered partial maxima
    xmax = group_max<32>(xmax);
    double *rb = red[itercount & 1];
    if ((tid & 31) == 0) rb[tid >> 5] = xmax;
    __syncthreads();
    xmax = rb[0];
#pragma unroll
    for (int i = 1; i < NW; i++) xmax = fmax(xmax, rb[i]);
    deltaerr = div_fast(xmax, wmax);                                                // IM:330
    itercount++;
    if (deltaerr < 1.0e-11) break;                                                  // IM:335
#if DIRK_EOS_SERIES
    double rr[ITEMS], ee[ITEMS], emax = 0.0;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      if (k0 + i < nlev && dphi[i] > 0) bad = 1;
      rr[i] = rcp_fast(dphi[i]);
      ee[i] = (dold[i] - dphi[i]) * rr[i];
      emax = fmax(emax, fabs(ee[i]));
    }
    if (__all_sync(FULLMASK, emax < EOS_SERIES_TINY)) {
#pragma unroll
      for (int i = 0; i < ITEMS; i++) eos_level_tiny_step(ee[i], rr[i], pnh[i], cc[i]);
    } else if (__all_sync(FULLMASK, emax < EOS_SERIES_MAX)) {
#pragma unroll
      for (int i = 0; i < ITEMS; i++) eos_level_small_step(ee[i], rr[i], pnh[i], cc[i]);
    } else
#endif
    {
#pragma unroll
      for (int i = 0; i < ITEMS; i++) {
        if (k0 + i < nlev && dphi[i] > 0) bad = 1;
        eos_level(rth[i], dphi[i], pnh[i], cc[i]);                                  // IM:324
      }
    }
  }
  // ---- phi update (IM:339) and write back ----
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const int k = k0 + i, q = k * 16 + p;
    if (k < nlev) {
      gw[q] = w[i];
      gphi[q] = s_phi0[i * NT] + dtg * w[i];
    }
  }
  if (tid == 0) {
    atomicMax(A.max_iter, itercount);
    atomicMax(A.max_deltaerr_bits, (unsigned long long)__double_as_longlong(deltaerr));
    if (itercount >= maxiter) atomicOr(A.err, ERR_DIRK_NOT_CONVERGED);
  }
  }   // element loop
  if (bad) atomicOr(A.err, ERR_EOS_BAD_STATE);
}

template <int ITEMS, int TPC, bool DSS = false>
static void launch_dirk_col_t(const DirkArgs &A, cudaStream_t st)
{
  using Cfg = DirkColCfg<ITEMS, TPC>;
  static int grid_cap = 0;
  if (!grid_cap) {
    cudaFuncSetAttribute(dirk_col_kernel<ITEMS, TPC, DSS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
    int dev = 0, nsm = 148, per_sm = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dirk_col_kernel<ITEMS, TPC, DSS>, Cfg::NT, Cfg::smem_bytes);
    grid_cap = nsm * (per_sm > 0 ? per_sm : 1);      // persistent: every CTA resident, elements strided over them
  }
  static const int grid_env = [] { const char *e = getenv("THETAL_B200_DIRK_GRID"); return e ? atoi(e) : -1; }();
  int grid = A.prm.nelem < grid_cap ? A.prm.nelem : grid_cap;
  if (grid_env == 0) grid = A.prm.nelem;                       // tuning: one element per CTA
  else if (grid_env > 0 && grid_env < A.prm.nelem) grid = grid_env;
  dirk_col_kernel<ITEMS, TPC, DSS><<<grid, Cfg::NT, Cfg::smem_bytes, st>>>(A);
}
