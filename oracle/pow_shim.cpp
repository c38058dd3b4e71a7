// TEST INFRASTRUCTURE: host build of the branch-free pow/div of e3sm_b200/csrc/thetal_pow.h so that the CPU
// test-suite can pin their accuracy (tests/test_pow.py), and of the fixed-point reproducible sum of
// thetal_reprosum.h (tests/test_reprosum.py).  Not part of the product.
#include "../e3sm_b200/csrc/thetal_pow.h"
#include "../e3sm_b200/csrc/thetal_reprosum.h"
using namespace thetal_pow;
static const double KAPPA = 287.04 / 1005.0;
static const PowConst C_GAMMA = {1.0 / (1.0 - KAPPA), 0.0, 2.0194836983862725, 1.3732662420330506e-16};
static const PowConst C_KM1 = {KAPPA - 1.0, 0.0, -1.0306441110016318, -8.6951363107339526e-17};
extern "C" {
void shim_pow_gamma(const double *x, double *y, long n) { for (long i = 0; i < n; i++) y[i] = pow_pos(x[i], C_GAMMA); }
void shim_pow_km1(const double *x, double *y, long n) { for (long i = 0; i < n; i++) y[i] = pow_pos(x[i], C_KM1); }
void shim_div(const double *a, const double *d, double *y, long n) { for (long i = 0; i < n; i++) y[i] = div_fast(a[i], d[i]); }
// sum of x[0..n) in the order given, the way diag_accumulate_kernel + thetal_b200_global_diagnostics do it
double shim_reprosum(const double *x, long n)
{
  double mx = 0;
  for (long i = 0; i < n; i++) mx = fabs(x[i]) > mx ? fabs(x[i]) : mx;
  const int emax = thetal_reprosum::exponent_of_bits(d2u(mx));
  long long acc[thetal_reprosum::NLIMB] = {0, 0, 0}, l[thetal_reprosum::NLIMB];
  for (long i = 0; i < n; i++) {
    thetal_reprosum::split(x[i], emax, l);
    for (int j = 0; j < thetal_reprosum::NLIMB; j++) acc[j] += l[j];
  }
  return thetal_reprosum::combine(acc, emax);
}
}
