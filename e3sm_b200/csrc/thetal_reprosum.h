// Order-independent global sum of doubles (the role of repro_sum_mod.F90 / wrap_repro_sum behind
// global_integral, global_norms_mod.F90:47-90): every addend is cut, relative to the exponent of the
// largest |addend|, into three signed 32-bit fixed-point limbs held in 64-bit integers.  Integer
// addition is associative, so atomics in any order, any element partition and any rank count give
// the same three integers, and therefore the same double.  96 bits below the largest exponent are
// kept (an addend loses bits only below 2^-96 of the largest one); the limbs overflow only beyond
// 2^31 addends.
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define THETAL_RS_HD __host__ __device__ __forceinline__
#else
#define THETAL_RS_HD static inline
#endif

namespace thetal_reprosum {

constexpr int NLIMB = 3;

// unbiased exponent of a non-negative double given as its bit pattern (0 and subnormals -> -1023)
THETAL_RS_HD int exponent_of_bits(uint64_t bits) { return (int)((bits >> 52) & 0x7ff) - 1023; }

// x -> limb[0..2]; requires |x| < 2^(emax+1)
THETAL_RS_HD void split(double x, int emax, long long limb[NLIMB])
{
  const double two32 = 4294967296.0;
  double r = ldexp(x, -(emax + 1)) * two32;      // |r| < 2^32, exact
  for (int i = 0; i < NLIMB; i++) {
    const double t = trunc(r);
    limb[i] = (long long)t;
    r = (r - t) * two32;                           // exact: fractional part, then a power-of-two scale
  }
}

// limbs (sums of split() outputs) -> double, one fixed evaluation order
THETAL_RS_HD double combine(const long long limb_in[NLIMB], int emax)
{
  long long l[NLIMB];
  for (int i = 0; i < NLIMB; i++) l[i] = limb_in[i];
  for (int i = NLIMB - 1; i > 0; i--) {            // carry so that the low limbs sit in [0, 2^32)
    const long long carry = l[i] >> 32;            // arithmetic shift = floor division
    l[i] -= carry * 4294967296LL;
    l[i - 1] += carry;
  }
  const double two_m32 = 1.0 / 4294967296.0;
  const double lo = ((double)l[2] * two_m32 + (double)l[1]) * two_m32;
  return ldexp((double)l[0] + lo, emax + 1 - 32);
}

}  // namespace thetal_reprosum
