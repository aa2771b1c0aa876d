// Does the DMMA rate depend on operand registers?  4 x 4 tiles from a[4], b[4] (the kernel's pattern) vs one shared (a, b);
// and with the operands reloaded from shared memory every k-step (fragment loads in the loop).
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
template <int MODE>
__global__ void __launch_bounds__(256) k(double* out, int iters, const double* in) {
  __shared__ double sh[2048];
  for (int i = threadIdx.x; i < 2048; i += 256) sh[i] = in[i];
  __syncthreads();
  double c[4][4][2];
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { c[i][j][0] = i; c[i][j][1] = j; }
  double a[4], b[4];
  for (int i = 0; i < 4; ++i) { a[i] = in[threadIdx.x + i]; b[i] = in[threadIdx.x + 7 + i]; }
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 2) {
      const double* As = &sh[((it & 3) * 8 + g) * 20 + t];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[i * 160]; b[i] = As[i * 160 + 640 + 4]; }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) dmma(c[i][j], MODE == 0 ? a[0] : a[i], MODE == 0 ? b[0] : b[j]);
  }
  double s = 0;
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) s += c[i][j][0] + c[i][j][1];
  out[blockIdx.x * 256 + threadIdx.x] = s;
}
template <int MODE> void run(const char* nm, int blocks, double* o, double* in) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000; float ms;
  k<MODE><<<blocks, 256>>>(o, 100, in);
  cudaEventRecord(e0); k<MODE><<<blocks, 256>>>(o, iters, in); cudaEventRecord(e1); cudaEventSynchronize(e1);
  cudaEventElapsedTime(&ms, e0, e1);
  printf("%-28s %3d warps/SM: %.2f TFLOP/s\n", nm, blocks * 8 / 148, 2.0 * 256 * 16 * iters * (double)blocks * 8 / (ms * 1e-3) / 1e12);
}
int main() {
  double *o, *in; cudaMalloc(&o, 8 * 148 * 8 * 256); cudaMalloc(&in, 8 * 4096); cudaMemset(in, 0, 8 * 4096);
  for (int wps : {8, 16}) {
    run<0>("shared a,b", 148 * wps / 8, o, in);
    run<1>("a[4] x b[4]", 148 * wps / 8, o, in);
    run<2>("a[4] x b[4] + LDS per step", 148 * wps / 8, o, in);
  }
  return 0;
}
