// TEST INFRASTRUCTURE: accuracy harness of the branch-free pow/div in e3sm_b200/csrc/thetal_pow.h against
// __float128 (g++ -O2 -mfma -ffp-contract=off check_pow.cpp -lquadmath; ~3 min).  Not part of the product.
#include "../e3sm_b200/csrc/thetal_pow.h"
#include <quadmath.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
using namespace thetal_pow;
static PowConst mk(__float128 g){ PowConst c; c.g_hi=(double)g; c.g_lo=(double)(g-(__float128)c.g_hi); __float128 k=g*M_LOG2Eq; c.k_hi=(double)k; c.k_lo=(double)(k-(__float128)c.k_hi); return c; }
int main(){
  const double RG=287.04, CPP=1005.0; const double KAPPA=RG/CPP;
  double gs[2]={1.0/(1.0-KAPPA), KAPPA-1.0};
  for(int ig=0;ig<2;ig++){
    double gd=gs[ig]; PowConst C=mk((__float128)gd);
    printf("g=%.17g consts %.17g %.17g %.17g %.17g\n",gd,C.g_hi,C.g_lo,C.k_hi,C.k_lo);
    double maxu=0,maxg=0; unsigned long long s=88172645463325252ULL; double worstx=0;
    for(long i=0;i<20000000;i++){
      s^=s<<13; s^=s>>7; s^=s<<17;
      double u=(s>>11)*(1.0/9007199254740992.0);
      double x=exp(log(1e-7)+u*(log(1e4)-log(1e-7)));
      __float128 ref=powq((__float128)x,(__float128)gd);
      double rd=(double)ref; double ulp=nextafter(fabs(rd),INFINITY)-fabs(rd);
      double mine=pow_pos(x,C); double eu=fabs((double)(((__float128)mine-ref)/(__float128)ulp));
      double lib=pow(x,gd); double eg=fabs((double)(((__float128)lib-ref)/(__float128)ulp));
      if(eu>maxu){maxu=eu;worstx=x;} if(eg>maxg)maxg=eg;
    }
    printf("  max ulp err: mine %.3f (x=%.17g)  glibc %.3f\n",maxu,worstx,maxg);
    // div_fast
  }
  double maxd=0; unsigned long long s=123456789ULL;
  for(long i=0;i<20000000;i++){ s^=s<<13; s^=s>>7; s^=s<<17; double a=(s>>11)*(1.0/9007199254740992.0)*1e5+1e-3; s^=s<<13; s^=s>>7; s^=s<<17; double d=-((s>>11)*(1.0/9007199254740992.0)*1e4+1e-2);
    __float128 ref=(__float128)a/(__float128)d; double rd=(double)ref; double ulp=nextafter(fabs(rd),INFINITY)-fabs(rd);
    double e=fabs((double)(((__float128)div_fast(a,d)-ref)/(__float128)ulp)); if(e>maxd)maxd=e; }
  printf("div_fast max ulp %.3f\n",maxd);
}
