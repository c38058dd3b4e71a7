// Branch-free x^g for finite, normal x > 0 and a compile-time exponent g (the two powers of the
// theta-l equation of state: pnh = p0 (x/p0)^(1/(1-kappa)), eos.F90:178, and (p/p0)^(kappa-1),
// eos.F90:296).  The library pow() is accurate but full of special-case branches, which stop the
// compiler from interleaving the independent point-levels a thread owns; this one is straight-line
// code (~60 FP64 instructions), so several evaluations overlap in the FP64 pipe.
//
//   x = 2^e m, m in [sqrt(1/2), sqrt(2));  log m = 2 atanh(f), f = (m-1)/(m+1) kept as a
//   double-double;  y = g e + (g log2 e) log m as a double-double;  x^g = 2^n exp((y-n) ln 2).
// Measured on 2e7 log-uniform x in [1e-7, 1e4] per exponent against __float128 powq: max 1.56 ulp
// (glibc pow: 0.51 ulp; CUDA pow: documented 2 ulp); div_fast: 0.5 ulp on 2e7 random quotients
// (oracle/check_pow.cpp is the harness).
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define THETAL_HD __host__ __device__ __forceinline__
#else
#define THETAL_HD static inline
#endif

namespace thetal_pow {

struct PowConst {          // g and g*log2(e) as double-doubles
  double g_hi, g_lo, k_hi, k_lo;
};

THETAL_HD uint64_t d2u(double x)
{
#if defined(__CUDA_ARCH__)
  return (uint64_t)__double_as_longlong(x);
#else
  uint64_t u; memcpy(&u, &x, 8); return u;
#endif
}
THETAL_HD double u2d(uint64_t u)
{
#if defined(__CUDA_ARCH__)
  return __longlong_as_double((long long)u);
#else
  double x; memcpy(&x, &u, 8); return x;
#endif
}

// 1/d to 1 ulp (2^-52 measured on B200 over 3e8 arguments, tools/ubench_fp64.cu), no special cases, any
// normal d.  One MUFU.RCP64H (seed good to 2^-20) and a cubic correction: e = 1 - d r, r (1 + e + e^2).
// Measured against the float-seed version this replaces (F2F + MUFU.RCP + F2F + one Newton step, 2^-46):
// 2.3 instead of 6.3 cycles per warp and SM at saturation (the F2F conversions and MUFU share the
// 16-lane XU pipe) and 48 instead of 76 cycles of dependent latency.
THETAL_HD double rcp_fast(double d)
{
#if defined(__CUDA_ARCH__)
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
#else
  double r = u2d(d2u(1.0 / d) & 0xffffffff00000000ULL);   // host stand-in of the 20-bit seed
#endif
  double e = fma(-d, r, 1.0);
  e = fma(e, e, e);
  return fma(r, e, r);
}

// a/d to <= 1 ulp (not correctly rounded), branch-free
THETAL_HD double div_fast(double a, double d)
{
  const double r = rcp_fast(d);
  const double q = a * r;
  const double rem = fma(-q, d, a);
  return fma(rem, r, q);
}

THETAL_HD double pow_pos(double x, const PowConst &C)
{
  // 1. exponent / mantissa split, m in [sqrt(1/2), sqrt(2))
  const uint64_t ix = d2u(x);
  int e = (int)(ix >> 52) - 1023;
  double m = u2d((ix & 0x000fffffffffffffULL) | 0x3ff0000000000000ULL);
  const bool big = m > 1.4142135623730951;
  m = big ? 0.5 * m : m;
  e += big ? 1 : 0;
  // 2. f = (m-1)/(m+1) as f_hi + f_lo
  const double t = m - 1.0;                       // exact
  const double d_hi = 2.0 + t;
  const double d_lo = (2.0 - d_hi) + t;           // exact (Fast2Sum)
  const double rd = rcp_fast(d_hi);
  const double f_hi = t * rd;
  double rem = fma(-f_hi, d_hi, t);
  rem = fma(-f_hi, d_lo, rem);
  const double f_lo = rem * rd;
  // 3. log m = 2 f + f^3 (2/3 + 2/5 z + ... + 2/23 z^10), z = f^2 <= 0.0295
  const double z = f_hi * f_hi;
  double P = 2.0 / 23.0;
  P = fma(P, z, 2.0 / 21.0);
  P = fma(P, z, 2.0 / 19.0);
  P = fma(P, z, 2.0 / 17.0);
  P = fma(P, z, 2.0 / 15.0);
  P = fma(P, z, 2.0 / 13.0);
  P = fma(P, z, 2.0 / 11.0);
  P = fma(P, z, 2.0 / 9.0);
  P = fma(P, z, 2.0 / 7.0);
  P = fma(P, z, 2.0 / 5.0);
  P = fma(P, z, 2.0 / 3.0);
  const double L_hi = 2.0 * f_hi;
  const double L_lo = fma(2.0, f_lo, (f_hi * z) * P);
  // 4. y = g e + (g log2 e)(L_hi + L_lo) as y_hi + y_lo
  const double a_hi = C.k_hi * L_hi;
  double a_lo = fma(C.k_hi, L_hi, -a_hi);
  a_lo = fma(C.k_hi, L_lo, a_lo);
  a_lo = fma(C.k_lo, L_hi, a_lo);
  const double ed = (double)e;
  const double b_hi = C.g_hi * ed;
  double b_lo = fma(C.g_hi, ed, -b_hi);
  b_lo = fma(C.g_lo, ed, b_lo);
  const double y_hi = b_hi + a_hi;
  const double bb = y_hi - b_hi;
  const double err = (b_hi - (y_hi - bb)) + (a_hi - bb);
  const double y_lo = err + (a_lo + b_lo);
  // 5. x^g = 2^n exp(r ln 2), |r| <= 1/2
  const double n = rint(y_hi);
  const double r = (y_hi - n) + y_lo;
  const double w = fma(r, 0.6931471805599453094, r * 2.3190468138462996e-17);
  double Q = 1.0 / 6227020800.0;                  // 1/13!
  Q = fma(Q, w, 1.0 / 479001600.0);
  Q = fma(Q, w, 1.0 / 39916800.0);
  Q = fma(Q, w, 1.0 / 3628800.0);
  Q = fma(Q, w, 1.0 / 362880.0);
  Q = fma(Q, w, 1.0 / 40320.0);
  Q = fma(Q, w, 1.0 / 5040.0);
  Q = fma(Q, w, 1.0 / 720.0);
  Q = fma(Q, w, 1.0 / 120.0);
  Q = fma(Q, w, 1.0 / 24.0);
  Q = fma(Q, w, 1.0 / 6.0);
  Q = fma(Q, w, 0.5);
  const double ex = fma(w * w, Q, w) + 1.0;       // exp(w) in (0.70, 1.42)
  return u2d(d2u(ex) + ((uint64_t)(int64_t)n << 52));
}

}  // namespace thetal_pow
