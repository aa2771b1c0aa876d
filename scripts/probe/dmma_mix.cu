// Does mixing DFMA/DMUL into a DMMA stream cost more than its own pipe cycles?  Per 16 DMMAs: X DFMAs.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
template <int X, bool SPREAD>
__global__ void __launch_bounds__(256) k(double* out, int iters, const double* in) {
  double c[4][4][2];
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { c[i][j][0] = i; c[i][j][1] = j; }
  double a[4], b[4], z[8];
  for (int i = 0; i < 4; ++i) { a[i] = in[threadIdx.x + i]; b[i] = in[threadIdx.x + 7 + i]; }
  for (int i = 0; i < 8; ++i) z[i] = in[i] + i;
  const double y = in[9] + 1.0000001;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        dmma(c[i][j], a[i], b[j]);
        if (SPREAD && (i * 4 + j) < X) asm volatile("fma.rn.f64 %0, %0, %1, %1;" : "+d"(z[(i * 4 + j) & 7]) : "d"(y));
      }
    if (!SPREAD) {
#pragma unroll
      for (int q = 0; q < X; ++q) asm volatile("fma.rn.f64 %0, %0, %1, %1;" : "+d"(z[q & 7]) : "d"(y));
    }
  }
  double s = 0;
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) s += c[i][j][0] + c[i][j][1];
  for (int i = 0; i < 8; ++i) s += z[i];
  out[blockIdx.x * 256 + threadIdx.x] = s;
}
template <int X, bool SPREAD> void run(double* o, double* in) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000, blocks = 148 * 2; float ms;
  k<X, SPREAD><<<blocks, 256>>>(o, 100, in);
  cudaEventRecord(e0); k<X, SPREAD><<<blocks, 256>>>(o, iters, in); cudaEventRecord(e1); cudaEventSynchronize(e1);
  cudaEventElapsedTime(&ms, e0, e1);
  const double dm = 2.0 * 256 * 16 * iters * (double)blocks * 8 / (ms * 1e-3) / 1e12;
  const double ideal = 37.1 * (16.0 * 16) / (16.0 * 16 + 2.0 * X);
  printf("%2d DFMA per 16 DMMA (%s): DMMA rate %.2f TFLOP/s (pipe-cycle model %.2f)\n", X, SPREAD ? "spread" : "bunched", dm, ideal);
}
int main() {
  double *o, *in; cudaMalloc(&o, 8 * 148 * 16 * 256); cudaMalloc(&in, 8 * 4096); cudaMemset(in, 0, 8 * 4096);
  run<0, false>(o, in); run<1, false>(o, in); run<4, false>(o, in); run<8, false>(o, in); run<16, false>(o, in);
  run<1, true>(o, in); run<4, true>(o, in); run<8, true>(o, in); run<16, true>(o, in);
  return 0;
}
