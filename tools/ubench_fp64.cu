// FP64 micro-benchmarks on B200 (sm_100a): dependent-issue latency and per-SM throughput of DFMA, the two
// reciprocal seeds (MUFU.RCP via float, MUFU.RCP64H), F2F, warp shuffles of doubles and pow_pos.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_fp64 ubench_fp64.cu ; run on one GPU.
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>
#include "../e3sm_b200/csrc/thetal_pow.h"

__device__ __forceinline__ double rcp64h_seed(double d)
{
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  return r;
}
__device__ __forceinline__ double rcp_h1(double d) { double r = rcp64h_seed(d); double e = fma(-d, r, 1.0); return fma(r, e, r); }
__device__ __forceinline__ double rcp_h2(double d) { double r = rcp64h_seed(d); double e = fma(-d, r, 1.0); e = fma(e, e, e); return fma(r, e, r); }

template <int OP, int ILP>
__global__ void kern(double *out, long long *cyc, int n, double seed)
{
  double a[ILP];
#pragma unroll
  for (int i = 0; i < ILP; i++) a[i] = seed + 0.001 * (threadIdx.x + i);
  const double b = 1.0000001, c = 1e-9;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < n; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      if (OP == 0) a[i] = fma(a[i], b, c);
      if (OP == 1) a[i] = thetal_pow::rcp_fast(a[i]) + 1.0;        // float-seed reciprocal + add (keeps values near 1.x)
      if (OP == 2) a[i] = rcp_h1(a[i]) + 1.0;
      if (OP == 3) a[i] = rcp_h2(a[i]) + 1.0;
      if (OP == 4) a[i] = __shfl_xor_sync(0xffffffffu, a[i], 1) + c;
      if (OP == 5) a[i] = thetal_pow::pow_pos(a[i], thetal_pow::PowConst{1.3997994317232159, 0.0, 2.0194836983862725, 1.3732662420330506e-16}) * 0.5 + 1.0;
      if (OP == 6) a[i] = (double)(float)a[i] + c;              // F2F pair
      if (OP == 7) a[i] = a[i] * b;                              // DMUL
      if (OP == 8) a[i] = rcp64h_seed(a[i]) + 1.0;
      if (OP == 9) a[i] = 1.0 / a[i] + 1.0;                       // IEEE division
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

__global__ void accuracy(double *maxerr)
{
  double e1 = 0, e2 = 0, e3 = 0, e0 = 0;
  unsigned long long s = 88172645463325252ULL + threadIdx.x * 977 + blockIdx.x * 7919;
  for (int i = 0; i < 20000; i++) {
    s ^= s << 13; s ^= s >> 7; s ^= s << 17;
    double x = 1.0 + (double)(s >> 11) * (1.0 / 9007199254740992.0);      // [1,2)
    x = ldexp(x, (int)(s & 63) - 32);
    const double ex = 1.0 / x;
    e0 = fmax(e0, fabs(rcp64h_seed(x) - ex) / ex);
    e1 = fmax(e1, fabs(rcp_h1(x) - ex) / ex);
    e2 = fmax(e2, fabs(rcp_h2(x) - ex) / ex);
    e3 = fmax(e3, fabs(thetal_pow::rcp_fast(x) - ex) / ex);
  }
  atomicMax((unsigned long long *)&maxerr[0], (unsigned long long)__double_as_longlong(e0));
  atomicMax((unsigned long long *)&maxerr[1], (unsigned long long)__double_as_longlong(e1));
  atomicMax((unsigned long long *)&maxerr[2], (unsigned long long)__double_as_longlong(e2));
  atomicMax((unsigned long long *)&maxerr[3], (unsigned long long)__double_as_longlong(e3));
}

template <int OP, int ILP>
void run(const char *name, int ops_per, int blocks, int threads)
{
  double *out; long long *cyc;
  cudaMalloc(&out, sizeof(double) * blocks * threads); cudaMalloc(&cyc, sizeof(long long) * blocks);
  const int n = 2000;
  kern<OP, ILP><<<blocks, threads>>>(out, cyc, n, 1.5);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  kern<OP, ILP><<<blocks, threads>>>(out, cyc, n, 1.5);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h; cudaMemcpy(&h, cyc, sizeof h, cudaMemcpyDeviceToHost);
  const double per = (double)h / ((double)n * ILP);
  const double warps = (double)threads / 32;
  printf("%-34s blocks %4d thr %4d ILP %d : %7.2f cyc per op per warp-chain; %6.2f cyc per warp-op per SM-CTA (%.0f warps); %.3f ms\n",
         name, blocks, threads, ILP, per * ILP, (double)h / ((double)n * ILP * warps), warps, ms);
  cudaFree(out); cudaFree(cyc);
}

int main()
{
  // latency: one warp, one chain
  run<0, 1>("DFMA latency", 1, 1, 32);
  run<7, 1>("DMUL latency", 1, 1, 32);
  run<6, 1>("F2F f64->f32->f64 + DADD latency", 1, 1, 32);
  run<8, 1>("MUFU.RCP64H + DADD latency", 1, 1, 32);
  run<1, 1>("rcp_fast (float seed) + DADD latency", 1, 1, 32);
  run<2, 1>("rcp64h + 1 NR + DADD latency", 1, 1, 32);
  run<3, 1>("rcp64h + cubic + DADD latency", 1, 1, 32);
  run<9, 1>("IEEE 1/x + DADD latency", 1, 1, 32);
  run<4, 1>("SHFL.64 + DADD latency", 1, 1, 32);
  run<5, 1>("pow_pos latency (1 chain)", 1, 1, 32);
  // throughput: one CTA per SM, many warps, ILP 4
  run<0, 4>("DFMA throughput", 1, 148, 1024);
  run<7, 4>("DMUL throughput", 1, 148, 1024);
  run<6, 4>("F2F pair + DADD throughput", 1, 148, 1024);
  run<8, 4>("RCP64H + DADD throughput", 1, 148, 1024);
  run<1, 4>("rcp_fast + DADD throughput", 1, 148, 1024);
  run<2, 4>("rcp64h+1NR + DADD throughput", 1, 148, 1024);
  run<3, 4>("rcp64h+cubic + DADD throughput", 1, 148, 1024);
  run<4, 4>("SHFL.64 + DADD throughput", 1, 148, 1024);
  run<5, 2>("pow_pos throughput 32 warps", 1, 148, 1024);
  run<5, 5>("pow_pos 16 warps ILP5", 1, 148, 512);
  run<5, 5>("pow_pos 8 warps ILP5", 1, 148, 256);
  run<5, 1>("pow_pos 16 warps ILP1", 1, 148, 512);
  run<0, 1>("DFMA 16 warps ILP1", 1, 148, 512);
  run<0, 1>("DFMA 8 warps ILP1", 1, 148, 256);
  run<0, 1>("DFMA 4 warps ILP1", 1, 148, 128);
  double *me; cudaMalloc(&me, 32); cudaMemset(me, 0, 32);
  accuracy<<<64, 256>>>(me);
  double h[4]; cudaMemcpy(h, me, 32, cudaMemcpyDeviceToHost);
  printf("max rel err: rcp64h seed %.3e (2^%.1f), +1NR %.3e (2^%.1f), +cubic %.3e (2^%.1f), rcp_fast float seed %.3e (2^%.1f)\n",
         h[0], log2(h[0]), h[1], log2(h[1]), h[2], log2(h[2]), h[3], log2(h[3]));
  return 0;
}
