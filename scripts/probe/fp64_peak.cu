// Peak fp64 throughput of one B200 as this box delivers it: DFMA (vector pipe) and DMMA (mma.sync m8n8k4 f64),
// all SMs busy, register-resident operands.  Prints TFLOP/s and the SM clock seen through clock64 / wall time.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters, double y) {
  double z[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) z[q] = threadIdx.x * 1e-9 + q;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int q = 0; q < 16; ++q) z[q] = fma(z[q], y, 1e-9);
  }
  double s = 0;
#pragma unroll
  for (int q = 0; q < 16; ++q) s += z[q];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) dmma_kernel(double* out, int iters, double a, double b) {
  double c[16][2];
#pragma unroll
  for (int q = 0; q < 16; ++q) { c[q][0] = threadIdx.x; c[q][1] = q; }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int q = 0; q < 16; ++q)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[q][0]), "+d"(c[q][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int q = 0; q < 16; ++q) s += c[q][0] + c[q][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* o; cudaMalloc(&o, sizeof(double) * sms * 8 * 256);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int warps_per_sm : {8, 16, 32, 64}) {
    const int blocks = sms * warps_per_sm / 8, iters = 20000;
    float ms;
    dfma_kernel<<<blocks, 256>>>(o, 1000, 1.0000001);
    cudaEventRecord(e0); dfma_kernel<<<blocks, 256>>>(o, iters, 1.0000001); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    const double f1 = 2.0 * 16 * iters * (double)blocks * 256 / (ms * 1e-3) / 1e12;
    dmma_kernel<<<blocks, 256>>>(o, 1000, 1.0, 1e-9);
    cudaEventRecord(e0); dmma_kernel<<<blocks, 256>>>(o, iters, 1.0, 1e-9); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    const double f2 = 2.0 * 256 * 16 * iters * (double)blocks * 8 / (ms * 1e-3) / 1e12;
    printf("%2d warps/SM: DFMA %.2f TFLOP/s, DMMA m8n8k4 %.2f TFLOP/s (%s)\n", warps_per_sm, f1, f2, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
