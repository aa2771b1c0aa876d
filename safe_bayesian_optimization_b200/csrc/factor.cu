// factor.cu -- fp64 factor state of the GP on the GPU: K = k(X,X) + noise I = L L^T by a
// blocked right-looking Cholesky, L^-1 by blocked forward substitution, alpha = K^-1 r, the
// rank-k append, and the fp16 hi/lo operand planes of L^-1 for the tcgen05 contraction.
//
// This is the "fit" half of the external GP node the library replaces (reference
// launch/safe_bayesian_optimization.launch.py:111-117 mapping_package/terrain_mapping_node;
// hyper-parameters config/lpsc.yaml:35-37).  Everything is hand-written: no cuBLAS/cuSOLVER.
//
// Layout: row-major, leading dimension n_cap.  Blocks are NB = 128 square.  All O(n^3) work is
// in one fp64 tile routine (128x128 output tile, 256 threads, 8x8 register tile per thread,
// K staged through shared memory 16 deep):
//   panel k:  L_kk = chol(A_kk), D_k = L_kk^-1            (one CTA, shared memory)
//             L_ik = A_ik D_k^T                 i > k      (tile NT)
//             A_ij -= L_ik L_jk^T           i >= j > k     (tile NT, lower blocks only)
//             X_kc = D_k B_kc                  c <= k      (tile NN)   -- L^-1, row block k
//             B_ic -= L_ik X_kc         i > k, c <= k      (tile NN)
// with B (the running right-hand side, initially I) stored in place in L^-1.
#include <cuda_fp16.h>

#include <cmath>

#include "common.cuh"

namespace sbo {
namespace {

constexpr int NB = kFactorNB;  // 128
constexpr int TK = 16;         // K depth per shared-memory stage
constexpr int kGemmThreads = 256;

// ---------------------------------------------------------------------------------------------
// 128x128 fp64 tile: acc[8][8] per thread.  Thread (ty, tx) = (tid / 16, tid % 16) owns rows
// {g*32 + ty*2 + v} and columns {g*32 + tx*2 + v}, g in 0..3, v in 0..1, so every shared-memory
// read is a 16-byte vector and the 16 lanes of a half-warp read 256 contiguous bytes.
// ---------------------------------------------------------------------------------------------
struct TileSmem {
  double a[TK][NB];  // a[kk][row]
  double b[TK][NB];  // b[kk][col]
};

__device__ __forceinline__ void tile_zero(double (&acc)[8][8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.0;
}

// Stage a 128 x TK chunk of a K-contiguous operand (rows r0.., columns k0..) transposed into
// s[kk][row].  Thread t reads 8 consecutive doubles of row t/2.
__device__ __forceinline__ void stage_kmajor(double (*s)[NB], const double* __restrict__ src, int ld, int r0, int k0) {
  const int t = threadIdx.x;
  const int row = t >> 1, kq = (t & 1) * 8;
  const double2* p = reinterpret_cast<const double2*>(src + (size_t)(r0 + row) * ld + k0 + kq);
  double2 v0 = p[0], v1 = p[1], v2 = p[2], v3 = p[3];
  s[kq + 0][row] = v0.x; s[kq + 1][row] = v0.y;
  s[kq + 2][row] = v1.x; s[kq + 3][row] = v1.y;
  s[kq + 4][row] = v2.x; s[kq + 5][row] = v2.y;
  s[kq + 6][row] = v3.x; s[kq + 7][row] = v3.y;
}

// Stage a TK x 128 chunk of a row-major operand (rows k0.., columns c0..) as is: s[kk][col].
__device__ __forceinline__ void stage_rowmajor(double (*s)[NB], const double* __restrict__ src, int ld, int k0, int c0) {
  const int t = threadIdx.x;
  const int kk = t >> 4, cq = (t & 15) * 8;
  const double2* p = reinterpret_cast<const double2*>(src + (size_t)(k0 + kk) * ld + c0 + cq);
  double2* d = reinterpret_cast<double2*>(&s[kk][cq]);
  d[0] = p[0]; d[1] = p[1]; d[2] = p[2]; d[3] = p[3];
}

__device__ __forceinline__ void tile_mac(const TileSmem& sm, double (&acc)[8][8]) {
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
  for (int kk = 0; kk < TK; ++kk) {
    double a[8], b[8];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const double2 av = *reinterpret_cast<const double2*>(&sm.a[kk][g * 32 + ty * 2]);
      const double2 bv = *reinterpret_cast<const double2*>(&sm.b[kk][g * 32 + tx * 2]);
      a[2 * g] = av.x; a[2 * g + 1] = av.y;
      b[2 * g] = bv.x; b[2 * g + 1] = bv.y;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
  }
}

// acc = A[r0.., 0..K) * B^T (kNT: B rows c0.., K contiguous) or A * B (B rows 0..K, columns c0..).
template <bool kNT>
__device__ __forceinline__ void tile_gemm(TileSmem& sm, const double* A, int lda, int r0, int ka0, const double* B,
                                          int ldb, int c0, int kb0, int K, double (&acc)[8][8]) {
  for (int k = 0; k < K; k += TK) {
    stage_kmajor(sm.a, A, lda, r0, ka0 + k);
    if (kNT) stage_kmajor(sm.b, B, ldb, c0, kb0 + k);
    else stage_rowmajor(sm.b, B, ldb, kb0 + k, c0);
    __syncthreads();
    tile_mac(sm, acc);
    __syncthreads();
  }
}

// C[r0.., c0..] = sign * acc (+ C if kAdd)
template <bool kAdd>
__device__ __forceinline__ void tile_store(double* C, int ldc, int r0, int c0, const double (&acc)[8][8], double sign) {
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = r0 + (i >> 1) * 32 + ty * 2 + (i & 1);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      double2* p = reinterpret_cast<double2*>(C + (size_t)r * ldc + c0 + g * 32 + tx * 2);
      double2 v = kAdd ? *p : make_double2(0.0, 0.0);
      v.x = fma(sign, acc[i][2 * g], v.x);
      v.y = fma(sign, acc[i][2 * g + 1], v.y);
      *p = v;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// K matrix
// ---------------------------------------------------------------------------------------------
__global__ void build_K_kernel(const double* __restrict__ xy, int n, int np, int ld, double sf2, double inv2l2,
                               double noise, double* __restrict__ A) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= np || j >= np) return;
  double v;
  if (i < n && j < n) {
    const double dx = xy[2 * i] - xy[2 * j], dy = xy[2 * i + 1] - xy[2 * j + 1];
    v = sf2 * exp(-(dx * dx + dy * dy) * inv2l2);
    if (i == j) v += noise;
  } else {
    v = (i == j) ? 1.0 : 0.0;  // identity padding up to the block boundary
  }
  A[(size_t)i * ld + j] = v;
}

// ---------------------------------------------------------------------------------------------
// Diagonal block: Cholesky + inverse of an nb x nb (nb <= 128) block in shared memory.
// Lower triangle of S holds L, strict upper triangle holds X^T (X = L^-1), diagonals apart.
// ---------------------------------------------------------------------------------------------
constexpr int kDiagThreads = 512;
constexpr int kDiagLd = NB + 1;
constexpr size_t kDiagSmem = sizeof(double) * ((size_t)NB * kDiagLd + 2 * NB);

__global__ void __launch_bounds__(kDiagThreads)
diag_chol_inv_kernel(double* __restrict__ Ablk, int ld, int nb, double* __restrict__ Dinv, int ldd, int row_base,
                     int* __restrict__ info) {
  extern __shared__ double smem_d[];
  double* S = smem_d;                       // [NB][kDiagLd]
  double* dX = smem_d + (size_t)NB * kDiagLd;  // 1 / L_ii
  double* dL = dX + NB;                        // L_ii
  __shared__ int bad;
  const int tid = threadIdx.x;
  if (tid == 0) bad = 0;
  for (int e = tid; e < nb * nb; e += kDiagThreads) {
    const int i = e / nb, j = e - i * nb;
    S[i * kDiagLd + j] = (j <= i) ? Ablk[(size_t)i * ld + j] : 0.0;
  }
  __syncthreads();
  const int row = tid & (NB - 1), cg = tid >> 7;  // 4 column groups
  // Right-looking column Cholesky.  The pivot sqrt goes to dL; S[j][j] keeps the pre-sqrt value so
  // that no thread overwrites what another is still reading (two barriers per column).
  for (int j = 0; j < nb; ++j) {
    __syncthreads();  // trailing updates of column j-1 are complete
    const double d = S[j * kDiagLd + j];
    if (!(d > 0.0)) {  // uniform: every thread reads the same d
      if (tid == 0) { bad = 1; atomicCAS(info, 0, row_base + j + 1); }
      break;
    }
    const double piv = sqrt(d);
    const double rinv = 1.0 / piv;
    if (tid == 0) dL[j] = piv;
    if (cg == 0 && row > j && row < nb) S[row * kDiagLd + j] *= rinv;
    __syncthreads();
    if (row > j && row < nb) {
      const double lij = S[row * kDiagLd + j];
      for (int c = j + 1 + cg; c <= row; c += 4) S[row * kDiagLd + c] -= lij * S[c * kDiagLd + j];
    }
  }
  __syncthreads();
  if (bad) return;
  // inverse: thread c owns column c of X; loops are uniform so L reads broadcast
  if (tid < nb) dX[tid] = 1.0 / dL[tid];
  __syncthreads();
  if (tid < nb) {
    const int c = tid;
    for (int i = 1; i < nb; ++i) {
      double s0 = 0.0, s1 = 0.0;
      if (i > c) {
        s0 = S[i * kDiagLd + c] * dX[c];  // p == c term: L[i][c] * X[c][c]
        int p = c + 1;
        for (; p + 1 < i; p += 2) {
          s0 = fma(S[i * kDiagLd + p], S[c * kDiagLd + p], s0);
          s1 = fma(S[i * kDiagLd + p + 1], S[c * kDiagLd + p + 1], s1);
        }
        if (p < i) s0 = fma(S[i * kDiagLd + p], S[c * kDiagLd + p], s0);
        S[c * kDiagLd + i] = -(s0 + s1) * dX[i];  // X[i][c] stored transposed in the upper triangle
      }
      // X[.][c] is private to this thread: no barrier needed between rows
    }
  }
  __syncthreads();
  for (int e = tid; e < nb * nb; e += kDiagThreads) {
    const int i = e / nb, j = e - i * nb;
    Ablk[(size_t)i * ld + j] = (j < i) ? S[i * kDiagLd + j] : (j == i ? dL[i] : 0.0);
    Dinv[(size_t)i * ldd + j] = (j < i) ? S[j * kDiagLd + i] : (j == i ? dX[i] : 0.0);
  }
}

// L_ik = A_ik D_k^T for row blocks i = k+1+blockIdx.x (in place).
__global__ void __launch_bounds__(kGemmThreads)
panel_trsm_kernel(double* __restrict__ A, int ld, int k, const double* __restrict__ Dk) {
  __shared__ TileSmem sm;
  const int r0 = (k + 1 + blockIdx.x) * NB, c0 = k * NB;
  double acc[8][8];
  tile_zero(acc);
  tile_gemm<true>(sm, A, ld, r0, c0, Dk, NB, 0, 0, NB, acc);  // C[r][j] = sum_p A[r][c0+p] D[j][p]
  tile_store<false>(A, ld, r0, c0, acc, 1.0);
}

// A_ij -= L_ik L_jk^T over the lower block pairs i >= j > k.
__global__ void __launch_bounds__(kGemmThreads)
syrk_kernel(double* __restrict__ A, int ld, int k) {
  __shared__ TileSmem sm;
  // blockIdx.x enumerates pairs (a >= b) of the trailing blocks: t = a(a+1)/2 + b
  int a = (int)((sqrt(8.0 * (double)blockIdx.x + 1.0) - 1.0) * 0.5);
  while ((a + 1) * (a + 2) / 2 <= (int)blockIdx.x) ++a;
  while (a * (a + 1) / 2 > (int)blockIdx.x) --a;
  const int b = blockIdx.x - a * (a + 1) / 2;
  const int r0 = (k + 1 + a) * NB, c0 = (k + 1 + b) * NB, kc = k * NB;
  double acc[8][8];
  tile_zero(acc);
  tile_gemm<true>(sm, A, ld, r0, kc, A, ld, c0, kc, NB, acc);
  tile_store<true>(A, ld, r0, c0, acc, -1.0);
}

// X_kc = D_k B_kc for c = blockIdx.x < k (in place in Linv); X_kk = D_k.
__global__ void __launch_bounds__(kGemmThreads)
trtri_row_kernel(double* __restrict__ X, int ld, int k, const double* __restrict__ Dk) {
  __shared__ TileSmem sm;
  const int c = blockIdx.x;
  const int r0 = k * NB, c0 = c * NB;
  if (c == k) {
    for (int e = threadIdx.x; e < NB * NB; e += kGemmThreads) {
      const int i = e / NB, j = e - i * NB;
      X[(size_t)(r0 + i) * ld + c0 + j] = Dk[(size_t)i * NB + j];
    }
    return;
  }
  double acc[8][8];
  tile_zero(acc);
  tile_gemm<false>(sm, Dk, NB, 0, 0, X, ld, c0, r0, NB, acc);  // C[r][j] = sum_p D[r][p] B[r0+p][c0+j]
  tile_store<false>(X, ld, r0, c0, acc, 1.0);
}

// B_ic -= L_ik X_kc for i = k+1+blockIdx.y, c = blockIdx.x <= k.
__global__ void __launch_bounds__(kGemmThreads)
trtri_update_kernel(double* __restrict__ X, int ld, const double* __restrict__ A, int k) {
  __shared__ TileSmem sm;
  const int r0 = (k + 1 + blockIdx.y) * NB, c0 = blockIdx.x * NB, kc = k * NB;
  double acc[8][8];
  tile_zero(acc);
  tile_gemm<false>(sm, A, ld, r0, kc, X, ld, c0, kc, NB, acc);
  tile_store<true>(X, ld, r0, c0, acc, -1.0);
}

// Zero the identity padding of L^-1 (rows/cols in [n, np)) so padded rows contribute nothing.
__global__ void clear_pad_kernel(double* __restrict__ X, int ld, int n, int np) {
  const int i = n + blockIdx.x * blockDim.x + threadIdx.x;
  if (i < np) X[(size_t)i * ld + i] = 0.0;
}

// ---------------------------------------------------------------------------------------------
// alpha = L^-T (L^-1 r)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gemv_lower_kernel(const double* __restrict__ X, int ld, int n, const double* __restrict__ r, double* __restrict__ w) {
  const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  const double* row = X + (size_t)warp * ld;
  double s = 0.0;
  for (int p = lane; p <= warp; p += 32) s = fma(row[p], r[p], s);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
  if (lane == 0) w[warp] = s;
}

__global__ void __launch_bounds__(256)
gemv_lower_t_kernel(const double* __restrict__ X, int ld, int n, const double* __restrict__ w, double* __restrict__ out,
                    float* __restrict__ out_f32) {
  __shared__ double red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  double s = 0.0;
  if (c < n)
    for (int i = blockIdx.x * 32 + ty; i < n; i += 8)
      if (i >= c) s = fma(X[(size_t)i * ld + c], w[i], s);
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < n) {
    double t = 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q) t += red[q][tx];
    out[c] = t;
    if (out_f32) out_f32[c] = (float)t;
  }
}

__global__ void fill_kernel(double* p, double v, int count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) p[i] = v;
}
__global__ void fill_f32_kernel(float* p, float v, int count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) p[i] = v;
}

// ---------------------------------------------------------------------------------------------
// fp16 operand planes of L^-1 for the tcgen05 contraction: xs = 2^shift x, hi = rn_f16(xs),
// lo = rn_f16(xs - hi); rows [r0, r1), all n_cap columns.  |xs| <= 2^15 by tc_plane_shifts' bound.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
planes_kernel(const double* __restrict__ X, int ld, int r0, int r1, int ncols, __half* __restrict__ planes,
              size_t lo_offset, double scale) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  const int r = r0 + blockIdx.y;
  if (r >= r1 || c >= ncols) return;
  const double x = X[(size_t)r * ld + c] * scale;
  const __half hi = __float2half_rn((float)x);
  const __half lo = __float2half_rn((float)(x - (double)__half2float(hi)));
  planes[(size_t)r * ld + c] = hi;
  planes[lo_offset + (size_t)r * ld + c] = lo;
}

// ---------------------------------------------------------------------------------------------
// rank-k append
// ---------------------------------------------------------------------------------------------
// K21[i][p] (k x n, ld n_cap) and S[i][j] = K22 + noise I (k x k, ld kMaxAppend)
__global__ void append_cov_kernel(const double* __restrict__ xy, int n, int k, double sf2, double inv2l2, double noise,
                                  double* __restrict__ K21, int ld, double* __restrict__ S) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (p >= n + k) return;
  const double dx = xy[2 * (n + i)] - xy[2 * p], dy = xy[2 * (n + i) + 1] - xy[2 * p + 1];
  double v = sf2 * exp(-(dx * dx + dy * dy) * inv2l2);
  if (p < n) {
    K21[(size_t)i * ld + p] = v;
  } else {
    const int j = p - n;
    if (j == i) v += noise;
    S[i * kMaxAppend + j] = v;
  }
}

// L21[i][j] = sum_{p<=j} K21[i][p] Linv[j][p]; one warp per row j of Linv, 8 new rows at a time.
__global__ void __launch_bounds__(256)
append_l21_kernel(const double* __restrict__ X, int ld, int n, int k, const double* __restrict__ K21,
                  double* __restrict__ L21) {
  const int j = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (j >= n) return;
  const double* row = X + (size_t)j * ld;
  for (int i0 = 0; i0 < k; i0 += 8) {
    double s[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) s[q] = 0.0;
    for (int p = lane; p <= j; p += 32) {
      const double x = row[p];
#pragma unroll
      for (int q = 0; q < 8; ++q)
        if (i0 + q < k) s[q] = fma(K21[(size_t)(i0 + q) * ld + p], x, s[q]);
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) s[q] += __shfl_down_sync(0xffffffffu, s[q], d);
      if (lane == 0 && i0 + q < k) L21[(size_t)(i0 + q) * ld + j] = s[q];
    }
  }
}

// S[i][j] -= L21[i] . L21[j] for j <= i; one warp per pair.
__global__ void __launch_bounds__(256)
append_schur_kernel(const double* __restrict__ L21, int ld, int n, int k, double* __restrict__ S) {
  const int w = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= k * k) return;
  const int i = w / k, j = w - i * k;
  if (j > i) return;
  double s = 0.0;
  for (int p = lane; p < n; p += 32) s = fma(L21[(size_t)i * ld + p], L21[(size_t)j * ld + p], s);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
  if (lane == 0) S[i * kMaxAppend + j] -= s;
}

// T[i][c] += sum_{p in split, p >= c} L21[i][p] Linv[p][c]; thread per column, gridDim.y row splits.
__global__ void __launch_bounds__(128)
append_t_kernel(const double* __restrict__ X, int ld, int n, int k, const double* __restrict__ L21,
                double* __restrict__ T) {
  const int c = blockIdx.x * 128 + threadIdx.x;
  const int per = (n + gridDim.y - 1) / gridDim.y;
  const int p_lo = blockIdx.y * per, p_hi = min(n, p_lo + per);
  if (c >= n || p_hi <= c) return;
  for (int i0 = 0; i0 < k; i0 += 8) {
    double s[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) s[q] = 0.0;
    for (int p = max(p_lo, c); p < p_hi; ++p) {
      const double x = X[(size_t)p * ld + c];
#pragma unroll
      for (int q = 0; q < 8; ++q)
        if (i0 + q < k) s[q] = fma(L21[(size_t)(i0 + q) * ld + p], x, s[q]);
    }
#pragma unroll
    for (int q = 0; q < 8; ++q)
      if (i0 + q < k) atomicAdd(&T[(size_t)(i0 + q) * ld + c], s[q]);
  }
}

// New rows of L^-1: X[n+i][c] = -sum_{q<=i} X22[i][q] T[q][c] (c < n), X[n+i][n+q] = X22[i][q].
__global__ void __launch_bounds__(256)
append_rows_kernel(double* __restrict__ X, int ld, int n, int k, const double* __restrict__ T,
                   const double* __restrict__ X22) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  const int i = blockIdx.y;
  if (c >= n + k) return;
  double v;
  if (c < n) {
    double s = 0.0;
    for (int q = 0; q <= i; ++q) s = fma(X22[i * kMaxAppend + q], T[(size_t)q * ld + c], s);
    v = -s;
  } else {
    const int q = c - n;
    v = (q <= i) ? X22[i * kMaxAppend + q] : 0.0;
  }
  X[(size_t)(n + i) * ld + c] = v;
}

int32_t check_info(sbo_ctx* ctx, const char* what) {
  int info = 0;
  SBO_CUDA(ctx, cudaMemcpyAsync(&info, ctx->d_info, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (info != 0)
    return fail(ctx, SBO_ERR_NOT_POSDEF,
                std::string(what) + ": covariance not positive definite at row " + std::to_string(info - 1));
  return SBO_OK;
}

int32_t compute_alpha(sbo_ctx* ctx, int n) {
  const int ld = ctx->n_cap;
  gemv_lower_kernel<<<(n * 32 + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_Linv, ld, n, ctx->d_r, ctx->d_w);
  SBO_LAUNCH_CHECK(ctx, "gemv_lower_kernel");
  gemv_lower_t_kernel<<<(n + 31) / 32, 256, 0, ctx->stream>>>(ctx->d_Linv, ld, n, ctx->d_w, ctx->d_alpha,
                                                              ctx->d_alpha_f32);
  SBO_LAUNCH_CHECK(ctx, "gemv_lower_t_kernel");
  return SBO_OK;
}

}  // namespace

int32_t refresh_planes(sbo_ctx* ctx, int row_begin, int row_end) {
  if (!ctx->d_planes || row_end <= row_begin) return SBO_OK;
  const int ld = ctx->n_cap;
  dim3 grid((ld + 255) / 256, row_end - row_begin);
  planes_kernel<<<grid, 256, 0, ctx->stream>>>(ctx->d_Linv, ld, row_begin, row_end, ld,
                                               reinterpret_cast<__half*>(ctx->d_planes), (size_t)ld * ld,
                                               std::ldexp(1.0, ctx->shift_b));
  SBO_LAUNCH_CHECK(ctx, "planes_kernel");
  return SBO_OK;
}

int32_t factor_full(sbo_ctx* ctx) {
  const int n = ctx->n, ld = ctx->n_cap;
  const int np = (n + NB - 1) / NB * NB;
  const int nblk = np / NB;
  const double sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  const double inv2l2 = 1.0 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale);
  cudaStream_t st = ctx->stream;

  static bool attr_done = false;
  if (!attr_done) {
    SBO_CUDA(ctx, cudaFuncSetAttribute(diag_chol_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kDiagSmem));
    attr_done = true;
  }
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_info, 0, sizeof(int), st));
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_Linv, 0, sizeof(double) * (size_t)ld * ld, st));
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_alpha, 0, sizeof(double) * (size_t)ld, st));
  if (ctx->d_alpha_f32) SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_alpha_f32, 0, sizeof(float) * (size_t)ld, st));
  if (n == 0) {
    SBO_TRY(refresh_planes(ctx, 0, ld));
    return SBO_OK;
  }
  {
    dim3 b(32, 8), g((np + 31) / 32, (np + 7) / 8);
    build_K_kernel<<<g, b, 0, st>>>(ctx->d_xy, n, np, ld, sf2, inv2l2, ctx->cfg.noise_var, ctx->d_A);
    SBO_LAUNCH_CHECK(ctx, "build_K_kernel");
  }
  for (int k = 0; k < nblk; ++k) {
    double* Dk = ctx->d_Dinv + (size_t)k * NB * NB;
    diag_chol_inv_kernel<<<1, kDiagThreads, kDiagSmem, st>>>(ctx->d_A + (size_t)k * NB * ld + (size_t)k * NB, ld, NB,
                                                             Dk, NB, k * NB, ctx->d_info);
    SBO_LAUNCH_CHECK(ctx, "diag_chol_inv_kernel");
    const int rem = nblk - k - 1;
    if (rem > 0) {
      panel_trsm_kernel<<<rem, kGemmThreads, 0, st>>>(ctx->d_A, ld, k, Dk);
      SBO_LAUNCH_CHECK(ctx, "panel_trsm_kernel");
      syrk_kernel<<<rem * (rem + 1) / 2, kGemmThreads, 0, st>>>(ctx->d_A, ld, k);
      SBO_LAUNCH_CHECK(ctx, "syrk_kernel");
    }
    trtri_row_kernel<<<k + 1, kGemmThreads, 0, st>>>(ctx->d_Linv, ld, k, Dk);
    SBO_LAUNCH_CHECK(ctx, "trtri_row_kernel");
    if (rem > 0) {
      dim3 g(k + 1, rem);
      trtri_update_kernel<<<g, kGemmThreads, 0, st>>>(ctx->d_Linv, ld, ctx->d_A, k);
      SBO_LAUNCH_CHECK(ctx, "trtri_update_kernel");
    }
  }
  if (np > n) {
    clear_pad_kernel<<<(np - n + 127) / 128, 128, 0, st>>>(ctx->d_Linv, ld, n, np);
    SBO_LAUNCH_CHECK(ctx, "clear_pad_kernel");
  }
  SBO_TRY(check_info(ctx, "sbo_set_observations"));
  SBO_TRY(compute_alpha(ctx, n));
  SBO_TRY(refresh_planes(ctx, 0, ld));
  return SBO_OK;
}

int32_t factor_append(sbo_ctx* ctx, int k) {
  const int n = ctx->n, ld = ctx->n_cap;
  const double sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  const double inv2l2 = 1.0 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale);
  cudaStream_t st = ctx->stream;
  double* K21 = ctx->d_small;                                   // [kMaxAppend][ld]; overwritten by T later
  double* L21 = ctx->d_small + (size_t)kMaxAppend * ld;         // [kMaxAppend][ld]
  double* S = ctx->d_small + (size_t)2 * kMaxAppend * ld;       // [kMaxAppend][kMaxAppend] -> L22
  double* X22 = S + (size_t)kMaxAppend * kMaxAppend;            // [kMaxAppend][kMaxAppend]

  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_info, 0, sizeof(int), st));
  {
    dim3 g((n + k + 127) / 128, k);
    append_cov_kernel<<<g, 128, 0, st>>>(ctx->d_xy, n, k, sf2, inv2l2, ctx->cfg.noise_var, K21, ld, S);
    SBO_LAUNCH_CHECK(ctx, "append_cov_kernel");
  }
  if (n > 0) {
    append_l21_kernel<<<(n * 32 + 255) / 256, 256, 0, st>>>(ctx->d_Linv, ld, n, k, K21, L21);
    SBO_LAUNCH_CHECK(ctx, "append_l21_kernel");
    append_schur_kernel<<<(k * k * 32 + 255) / 256, 256, 0, st>>>(L21, ld, n, k, S);
    SBO_LAUNCH_CHECK(ctx, "append_schur_kernel");
  }
  diag_chol_inv_kernel<<<1, kDiagThreads, kDiagSmem, st>>>(S, kMaxAppend, k, X22, kMaxAppend, n, ctx->d_info);
  SBO_LAUNCH_CHECK(ctx, "diag_chol_inv_kernel(append)");
  SBO_TRY(check_info(ctx, "sbo_append_observations"));
  double* T = K21;
  if (n > 0) {
    SBO_CUDA(ctx, cudaMemsetAsync(T, 0, sizeof(double) * (size_t)k * ld, st));
    int splits = (ctx->sm_count * 4) / ((n + 127) / 128);
    if (splits < 1) splits = 1;
    if (splits > 64) splits = 64;
    dim3 g((n + 127) / 128, splits);
    append_t_kernel<<<g, 128, 0, st>>>(ctx->d_Linv, ld, n, k, L21, T);
    SBO_LAUNCH_CHECK(ctx, "append_t_kernel");
  }
  {
    dim3 g((n + k + 255) / 256, k);
    append_rows_kernel<<<g, 256, 0, st>>>(ctx->d_Linv, ld, n, k, T, X22);
    SBO_LAUNCH_CHECK(ctx, "append_rows_kernel");
  }
  SBO_TRY(compute_alpha(ctx, n + k));
  SBO_TRY(refresh_planes(ctx, n, n + k));
  return SBO_OK;
}

}  // namespace sbo
