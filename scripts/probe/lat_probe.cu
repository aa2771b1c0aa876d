// Dependent-chain latencies of the fp64 instructions the factorisation's sequential parts use (one warp).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* t, double seed) {
  double x = seed + threadIdx.x * 1e-9, y = 1.0000001;
  long long c0, c1;
  const int N = 256;
  c0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) x = fma(x, y, 1e-9);
  c1 = clock64(); if (threadIdx.x == 0) t[0] = (c1 - c0);
  c0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) x = x + y;
  c1 = clock64(); if (threadIdx.x == 0) t[1] = (c1 - c0);
  c0 = clock64();
#pragma unroll 4
  for (int i = 0; i < N; ++i) x = sqrt(x + 2.0);
  c1 = clock64(); if (threadIdx.x == 0) t[2] = (c1 - c0);
  c0 = clock64();
#pragma unroll 4
  for (int i = 0; i < N; ++i) x = 1.0 / (x + 2.0);
  c1 = clock64(); if (threadIdx.x == 0) t[3] = (c1 - c0);
  c0 = clock64();
#pragma unroll 4
  for (int i = 0; i < N; ++i) x = rsqrt(x + 2.0);
  c1 = clock64(); if (threadIdx.x == 0) t[4] = (c1 - c0);
  c0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) x = __shfl_sync(0xffffffffu, x, (i * 7) & 31);
  c1 = clock64(); if (threadIdx.x == 0) t[5] = (c1 - c0);
  float f = (float)x;
  c0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) f = fmaf(f, 1.0000001f, 1e-9f);
  c1 = clock64(); if (threadIdx.x == 0) t[6] = (c1 - c0);
  // 8 independent DFMA chains (throughput of one warp)
  double z[8];
  for (int q = 0; q < 8; ++q) z[q] = x + q;
  c0 = clock64();
#pragma unroll 4
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int q = 0; q < 8; ++q) z[q] = fma(z[q], y, 1e-9);
  c1 = clock64(); if (threadIdx.x == 0) t[7] = (c1 - c0);
  for (int q = 0; q < 8; ++q) x += z[q];
  out[threadIdx.x] = x + f;
}
int main() {
  double* o; long long* t; cudaMalloc(&o, 32 * 8); cudaMalloc(&t, 8 * 8);
  for (int rep = 0; rep < 2; ++rep) k<<<1, 32>>>(o, t, 1.5);
  long long h[8]; cudaMemcpy(h, t, 64, cudaMemcpyDeviceToHost);
  const char* nm[8] = {"DFMA dep", "DADD dep", "sqrt dep", "div dep", "rsqrt dep", "SHFL.f64 dep", "FFMA dep", "DFMA x8 indep (per 8)"};
  for (int i = 0; i < 8; ++i) printf("%-24s %.1f clk/op\n", nm[i], h[i] / 256.0);
  return 0;
}
