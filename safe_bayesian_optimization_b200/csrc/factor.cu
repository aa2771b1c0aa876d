// factor.cu -- fp64 factor state of the GP on the GPU: K = k(X,X) + noise I = L L^T by a
// blocked right-looking Cholesky, L^-1 by blocked forward substitution, alpha = K^-1 r, the
// rank-k append, and the fp16 hi/lo operand planes of L^-1 for the tcgen05 contraction.
//
// This is the "fit" half of the external GP node the library replaces (reference
// launch/safe_bayesian_optimization.launch.py:111-117 mapping_package/terrain_mapping_node;
// hyper-parameters config/lpsc.yaml:35-37).  Everything is hand-written: no cuBLAS/cuSOLVER.
//
// Layout: row-major, leading dimension n_cap.  Blocks are NB = 128 square.  All O(n^3) work is
// in one fp64 tile routine per regime: small fits (n <= 2944) 64x64 output tiles on the FMA pipe (256 threads,
// 4x4 register tile per thread, K staged through shared memory 16 deep, four CTAs per 128-block); large fits
// 128x128 tiles on DMMA with cp.async double buffering (factor_dmma.cuh) under the look-ahead schedule of
// factor_dist.cuh, on one GPU or shared between several:
//   panel k:  L_kk = chol(A_kk), D_k = L_kk^-1            (one CTA, shared memory)
//             L_ik = A_ik D_k^T                 i > k      (tile NT)
//             A_ij -= L_ik L_jk^T           i >= j > k     (tile NT, lower blocks only)
//             X_kc = D_k B_kc                  c <= k      (tile NN)   -- L^-1, row block k
//             B_ic -= L_ik X_kc         i > k, c <= k      (tile NN)
// with B (the running right-hand side, initially I) stored in place in L^-1.
#include <cuda_fp16.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <mutex>
#include <new>

#include "common.cuh"

namespace sbo {
namespace {

constexpr int NB = kFactorNB;  // 128
constexpr int TK = 16;         // K depth per shared-memory stage
constexpr int kGemmThreads = 256;

// ---------------------------------------------------------------------------------------------
// TM x TM fp64 tile (TM = 128: acc[8][8] per thread, TM = 64: acc[4][4]), 256 threads.  Thread
// (ty, tx) = (tid / 16, tid % 16) owns rows {g*32 + ty*2 + v} and columns {g*32 + tx*2 + v},
// g in 0..TM/32-1, v in 0..1, so every shared-memory read is a 16-byte vector and the 16 lanes of a
// half-warp read 256 contiguous bytes.  The 64-tile is for small n, where a launch has fewer
// 128-tiles than the GPU has SMs and the step time is one thread's serial FMA count: a quarter of
// the work per thread, four times the CTAs.
// ---------------------------------------------------------------------------------------------
template <int TM> struct TileSmem {
  double a[TK][TM];  // a[kk][row]
  double b[TK][TM];  // b[kk][col]
};

template <int R> __device__ __forceinline__ void tile_zero(double (&acc)[R][R]) {
#pragma unroll
  for (int i = 0; i < R; ++i)
#pragma unroll
    for (int j = 0; j < R; ++j) acc[i][j] = 0.0;
}

// Stage a TM x TK chunk of a K-contiguous operand (rows r0.., columns k0..) transposed into
// s[kk][row].  Thread t reads 2048 / 256 / (128 / TM) consecutive doubles of one row.
template <int TM>
__device__ __forceinline__ void stage_kmajor(double (*s)[TM], const double* __restrict__ src, int ld, int r0, int k0) {
  const int t = threadIdx.x;
  if (TM == 128) {
    const int row = t >> 1, kq = (t & 1) * 8;
    const double2* p = reinterpret_cast<const double2*>(src + (size_t)(r0 + row) * ld + k0 + kq);
    double2 v0 = p[0], v1 = p[1], v2 = p[2], v3 = p[3];
    s[kq + 0][row] = v0.x; s[kq + 1][row] = v0.y;
    s[kq + 2][row] = v1.x; s[kq + 3][row] = v1.y;
    s[kq + 4][row] = v2.x; s[kq + 5][row] = v2.y;
    s[kq + 6][row] = v3.x; s[kq + 7][row] = v3.y;
  } else {
    const int row = t >> 2, kq = (t & 3) * 4;
    const double2* p = reinterpret_cast<const double2*>(src + (size_t)(r0 + row) * ld + k0 + kq);
    double2 v0 = p[0], v1 = p[1];
    s[kq + 0][row] = v0.x; s[kq + 1][row] = v0.y;
    s[kq + 2][row] = v1.x; s[kq + 3][row] = v1.y;
  }
}

// Stage a TK x TM chunk of a row-major operand (rows k0.., columns c0..) as is: s[kk][col].
template <int TM>
__device__ __forceinline__ void stage_rowmajor(double (*s)[TM], const double* __restrict__ src, int ld, int k0, int c0) {
  const int t = threadIdx.x;
  if (TM == 128) {
    const int kk = t >> 4, cq = (t & 15) * 8;
    const double2* p = reinterpret_cast<const double2*>(src + (size_t)(k0 + kk) * ld + c0 + cq);
    double2* d = reinterpret_cast<double2*>(&s[kk][cq]);
    d[0] = p[0]; d[1] = p[1]; d[2] = p[2]; d[3] = p[3];
  } else {
    const int kk = t >> 4, cq = (t & 15) * 4;
    const double2* p = reinterpret_cast<const double2*>(src + (size_t)(k0 + kk) * ld + c0 + cq);
    double2* d = reinterpret_cast<double2*>(&s[kk][cq]);
    d[0] = p[0]; d[1] = p[1];
  }
}

template <int TM>
__device__ __forceinline__ void tile_mac(const TileSmem<TM>& sm, double (&acc)[TM / 16][TM / 16]) {
  constexpr int G = TM / 32;
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
  for (int kk = 0; kk < TK; ++kk) {
    double a[2 * G], b[2 * G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const double2 av = *reinterpret_cast<const double2*>(&sm.a[kk][g * 32 + ty * 2]);
      const double2 bv = *reinterpret_cast<const double2*>(&sm.b[kk][g * 32 + tx * 2]);
      a[2 * g] = av.x; a[2 * g + 1] = av.y;
      b[2 * g] = bv.x; b[2 * g + 1] = bv.y;
    }
#pragma unroll
    for (int i = 0; i < 2 * G; ++i)
#pragma unroll
      for (int j = 0; j < 2 * G; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
  }
}

// acc = A[r0.., 0..K) * B^T (kNT: B rows c0.., K contiguous) or A * B (B rows 0..K, columns c0..).
template <bool kNT, int TM>
__device__ __forceinline__ void tile_gemm(TileSmem<TM>& sm, const double* A, int lda, int r0, int ka0, const double* B,
                                          int ldb, int c0, int kb0, int K, double (&acc)[TM / 16][TM / 16]) {
  for (int k = 0; k < K; k += TK) {
    stage_kmajor<TM>(sm.a, A, lda, r0, ka0 + k);
    if (kNT) stage_kmajor<TM>(sm.b, B, ldb, c0, kb0 + k);
    else stage_rowmajor<TM>(sm.b, B, ldb, kb0 + k, c0);
    __syncthreads();
    tile_mac<TM>(sm, acc);
    __syncthreads();
  }
}

// C[r0.., c0..] = sign * acc (+ C if kAdd)
template <bool kAdd, int TM>
__device__ __forceinline__ void tile_store(double* C, int ldc, int r0, int c0, const double (&acc)[TM / 16][TM / 16],
                                           double sign) {
  constexpr int G = TM / 32;
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
  for (int i = 0; i < 2 * G; ++i) {
    const int r = r0 + (i >> 1) * 32 + ty * 2 + (i & 1);
#pragma unroll
    for (int g = 0; g < G; ++g) {
      double2* p = reinterpret_cast<double2*>(C + (size_t)r * ldc + c0 + g * 32 + tx * 2);
      double2 v = kAdd ? *p : make_double2(0.0, 0.0);
      v.x = fma(sign, acc[i][2 * g], v.x);
      v.y = fma(sign, acc[i][2 * g + 1], v.y);
      *p = v;
    }
  }
}

// sub-tile offsets of a TM-tile inside its 128-block: blockIdx.z enumerates the (128 / TM)^2 sub-tiles
template <int TM> __device__ __forceinline__ int sub_r() { return TM == 128 ? 0 : (int)(blockIdx.z >> 1) * TM; }
template <int TM> __device__ __forceinline__ int sub_c() { return TM == 128 ? 0 : (int)(blockIdx.z & 1) * TM; }

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
// Diagonal block: Cholesky + inverse of an nb x nb (nb <= 128) block, one CTA, blocked by 32.
//   for each 32-column panel kb:  warp 0 factors and inverts the 32x32 diagonal sub-block in
//   registers (lane = row, shuffles broadcast the pivot column: no barrier inside the 32 steps);
//   all warps then form the panel below it (A_ik X_kk^T) and the trailing update.
//   The inverse of the whole block follows from the 32x32 inverses by block forward substitution.
// Shared layout: lower triangle of S holds L, strict upper triangle holds X^T (X = L^-1), the two
// diagonals live in dL / dX.  Blocks smaller than 128 are padded with the identity.
// ---------------------------------------------------------------------------------------------
constexpr int kDiagThreads = 256;
constexpr int kDiagLd = NB + 1;
constexpr int SB = 32;            // sub-block
constexpr int kPanelLd = NB - SB; // 96: rows below the first sub-block
struct DiagSmem {
  double S[NB][kDiagLd];          // lower: L, strict upper: X^T
  double P[SB][kPanelLd];         // current panel, transposed: P[t][i - (base + 32)] = L[i][base + t]
  double Xd[NB / SB][SB][SB + 1]; // dense 32x32 inverses of the diagonal sub-blocks (zeros above the diagonal)
  double Lt[SB][SB + 2];          // current sub-block's L^T (row stride even: 16-byte pairs)
  double cb[2][SB];               // pivot-column broadcast, double-buffered
  double dX[NB], dL[NB];
  int bad;
};
constexpr size_t kDiagSmem = sizeof(DiagSmem);

#ifdef SBO_DIAG_TIMING
__device__ long long g_diag_t[16];
#define DIAG_T(i) do { if (threadIdx.x == 0) g_diag_t[i] = clock64(); } while (0)
#else
#define DIAG_T(i) do {} while (0)
#endif

// Warp-level 32x32 Cholesky + inverse.  Lane r holds row r in registers; the pivot column is broadcast
// through shared memory (one STS + a few LDS.128 per column: shuffles of doubles cost 4-8 issue cycles each
// and 1000 of them paced the first version of this kernel).  Entries right of the diagonal hold garbage that
// is never read.  Column j:  a[c] -= a[j] cb[c] / d  (d = pivot, cb = column j before scaling), then
// a[j] *= d^-1/2.
__device__ __forceinline__ void warp_chol_inv32(DiagSmem& sm, int base, int row_base, int* info, double min_pivot) {
  const int lane = threadIdx.x & 31;
  double a[SB], rinv[SB];
#pragma unroll
  for (int c = 0; c < SB; ++c) a[c] = sm.S[base + lane][base + c];
#pragma unroll
  for (int j = 0; j < SB; ++j) {
    double* cb = sm.cb[j & 1];
    cb[lane] = a[j];
    __syncwarp();
    double d = cb[j];
    if (!(d > min_pivot)) {  // uniform
      if (lane == 0) { sm.bad = 1; atomicCAS(info, 0, row_base + base + j + 1); }
      d = 1.0;
    }
    const double ri = rsqrt(d);
    rinv[j] = ri;
    const double t = a[j] * (ri * ri);
#pragma unroll
    for (int c = j + 1; c < SB; ++c) a[c] = fma(-t, cb[c], a[c]);
    a[j] = lane == j ? d * ri : a[j] * ri;
  }
  // L to shared memory: the lower triangle of S, the diagonal, and L^T for the inverse below
#pragma unroll
  for (int c = 0; c < SB; ++c) {
    if (c < lane) sm.S[base + lane][base + c] = a[c];
    if (c == lane) sm.dL[base + lane] = a[c];
    sm.Lt[c][lane] = a[c];  // only [c][i >= c] is read
  }
  __syncwarp();
  // X = L^-1 by column sweep: lane c owns column c; after x[p] is final, s[i] -= L[i][p] x[p] for i > p
  double x[SB];
#pragma unroll
  for (int i = 0; i < SB; ++i) x[i] = (i == lane) ? 1.0 : 0.0;
#pragma unroll
  for (int q = 0; q < SB; ++q) {
    x[q] *= rinv[q];
#pragma unroll
    for (int i = q + 1; i < SB; ++i) x[i] = fma(-sm.Lt[q][i], x[q], x[i]);
  }
#pragma unroll
  for (int c = 0; c < SB; ++c) {
    if (c > lane) sm.S[base + lane][base + c] = x[c];  // X[c][lane], transposed into the upper triangle
    if (c == lane) sm.dX[base + lane] = x[c];
    sm.Xd[base / SB][c][lane] = x[c];
  }
}

__global__ void __launch_bounds__(kDiagThreads)
diag_chol_inv_kernel(double* __restrict__ Ablk, int ld, int nb, double* __restrict__ Dinv, int ldd, int row_base,
                     int* __restrict__ info, double min_pivot = 0.0) {   // a pivot must exceed min_pivot (0: LAPACK's test)
  extern __shared__ __align__(16) unsigned char smem_diag[];
  DiagSmem& sm = *reinterpret_cast<DiagSmem*>(smem_diag);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) sm.bad = 0;
  if (nb <= 8) {
    // A handful of rows (the node's steady state appends one observation at a time): one thread, same operations
    // as the warp routine (pivot d, ri = rsqrt(d), column scaled by ri, updates with the unscaled column).
    if (tid == 0) {
      double a[8][8], rinv[8];
      for (int i = 0; i < nb; ++i)
        for (int j = 0; j <= i; ++j) a[i][j] = Ablk[(size_t)i * ld + j];
      bool bad = false;
      for (int j = 0; j < nb && !bad; ++j) {
        double d = a[j][j];
        if (!(d > min_pivot)) {
          atomicCAS(info, 0, row_base + j + 1);
          bad = true;
          break;
        }
        const double ri = rsqrt(d);
        rinv[j] = ri;
        for (int i = j + 1; i < nb; ++i) {
          const double t = a[i][j] * (ri * ri);
          for (int c = j + 1; c <= i; ++c) a[i][c] = fma(-t, a[c][j], a[i][c]);
        }
        for (int i = j + 1; i < nb; ++i) a[i][j] *= ri;
        a[j][j] = d * ri;
      }
      if (!bad) {
        for (int c = 0; c < nb; ++c) {   // column c of X = L^-1
          double x[8];
          for (int i = 0; i < nb; ++i) x[i] = (i == c) ? 1.0 : 0.0;
          for (int q = c; q < nb; ++q) {
            x[q] *= rinv[q];
            for (int i = q + 1; i < nb; ++i) x[i] = fma(-a[i][q], x[q], x[i]);
          }
          for (int i = 0; i < nb; ++i) Dinv[(size_t)i * ldd + c] = (i >= c) ? x[i] : 0.0;
        }
      }
    }
    return;
  }
  if (nb <= SB) {
    // One sub-block (the Schur complement of a rank-k append, k <= 32): the warp-level routine is the whole job;
    // loading, factoring and inverting the other 15 sub-blocks of identity padding cost 60 of this kernel's 69 us.
    for (int e = tid; e < SB * SB; e += kDiagThreads) {
      const int i = e >> 5, j = e & (SB - 1);
      double v = (i == j) ? 1.0 : 0.0;
      if (i < nb && j < nb) v = (j <= i) ? Ablk[(size_t)i * ld + j] : 0.0;
      sm.S[i][j] = v;
    }
    __syncthreads();
    if (warp == 0) warp_chol_inv32(sm, 0, row_base, info, min_pivot);
    __syncthreads();
    if (sm.bad) return;
    for (int e = tid; e < nb * nb; e += kDiagThreads) {
      const int i = e / nb, j = e - i * nb;
      Dinv[(size_t)i * ldd + j] = (j < i) ? sm.S[j][i] : (j == i ? sm.dX[i] : 0.0);
    }
    return;
  }
#pragma unroll 4
  for (int e = tid; e < NB * NB; e += kDiagThreads) {
    const int i = e >> 7, j = e & (NB - 1);
    double v = (i == j) ? 1.0 : 0.0;  // identity padding of a short block
    if (i < nb && j < nb) v = (j <= i) ? Ablk[(size_t)i * ld + j] : 0.0;
    sm.S[i][j] = v;
  }
  DIAG_T(0);
  __syncthreads();
  DIAG_T(1);

  for (int kb = 0; kb < NB / SB; ++kb) {
    const int base = kb * SB;
    if (kb == 1) DIAG_T(6);
    if (warp == 0) warp_chol_inv32(sm, base, row_base, info, min_pivot);
    if (kb == 0) DIAG_T(3);
    __syncthreads();
    if (kb == 0) DIAG_T(4);
    const int below = NB - base - SB;  // rows under the sub-block
    if (below > 0) {
      // ---- panel: L[i][base+c] = sum_t A[i][base+t] X[c][t], rows i = base+32+warp+8q ----
      const int nq = below / 8;
      double out[12];
#pragma unroll
      for (int q = 0; q < 12; ++q) out[q] = 0.0;
#pragma unroll 4
      for (int t = 0; t < SB; ++t) {
        const double xv = sm.Xd[kb][lane][t];
#pragma unroll
        for (int q = 0; q < 12; ++q)
          if (q < nq) out[q] = fma(sm.S[base + SB + warp + 8 * q][base + t], xv, out[q]);
      }
      __syncthreads();
#pragma unroll
      for (int q = 0; q < 12; ++q)
        if (q < nq) {
          sm.S[base + SB + warp + 8 * q][base + lane] = out[q];
          sm.P[lane][warp + 8 * q] = out[q];
        }
      __syncthreads();
      if (kb == 0) DIAG_T(5);
      // ---- trailing update: A[i][j] -= L[i][base..] . L[j][base..], sub-blocks ib >= jb; lane = column,
      //      warp = 4 rows of every sub-block row ib ----
      const int nsub = below / SB;
      for (int jb = 0; jb < nsub; ++jb) {
        double acc[3][4];
#pragma unroll
        for (int u = 0; u < 3; ++u)
#pragma unroll
          for (int v = 0; v < 4; ++v) acc[u][v] = 0.0;
#pragma unroll 4
        for (int t = 0; t < SB; ++t) {
          const double lj = sm.P[t][jb * SB + lane];
#pragma unroll
          for (int u = 0; u < 3; ++u) {
            if (jb + u < nsub) {
              const double2 r01 = *reinterpret_cast<const double2*>(&sm.P[t][(jb + u) * SB + warp * 4]);
              const double2 r23 = *reinterpret_cast<const double2*>(&sm.P[t][(jb + u) * SB + warp * 4 + 2]);
              acc[u][0] = fma(r01.x, lj, acc[u][0]);
              acc[u][1] = fma(r01.y, lj, acc[u][1]);
              acc[u][2] = fma(r23.x, lj, acc[u][2]);
              acc[u][3] = fma(r23.y, lj, acc[u][3]);
            }
          }
        }
        const int j = base + SB + jb * SB + lane;
#pragma unroll
        for (int u = 0; u < 3; ++u)
          if (jb + u < nsub) {
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              const int i = base + SB + (jb + u) * SB + warp * 4 + v;
              if (j <= i) sm.S[i][j] -= acc[u][v];  // the upper triangle holds X^T and must not be touched
            }
          }
      }
      __syncthreads();
    }
  }
  DIAG_T(7);
  if (sm.bad) return;

  // ---- inverse of the whole block: X_ij = -X_ii sum_{p=j..i-1} L_ip X_pj for block row bi, block column bj < bi.
  //      lane = column of block bj, warp + 8q = row of block bi ----
  for (int bi = 1; bi < NB / SB; ++bi) {
    double t[3][4];
#pragma unroll
    for (int bj = 0; bj < 3; ++bj)
#pragma unroll
      for (int q = 0; q < 4; ++q) t[bj][q] = 0.0;
#pragma unroll
    for (int bj = 0; bj < 3; ++bj) {
      if (bj < bi) {
        // k in the diagonal block bj: dense copy, no masks
#pragma unroll 4
        for (int k = 0; k < SB; ++k) {
          const double xv = sm.Xd[bj][k][lane];
#pragma unroll
          for (int q = 0; q < 4; ++q) t[bj][q] = fma(sm.S[bi * SB + warp + 8 * q][bj * SB + k], xv, t[bj][q]);
        }
        // k in the blocks between: X[k][col] lives transposed in the upper triangle
#pragma unroll 4
        for (int k = (bj + 1) * SB; k < bi * SB; ++k) {
          const double xv = sm.S[bj * SB + lane][k];
#pragma unroll
          for (int q = 0; q < 4; ++q) t[bj][q] = fma(sm.S[bi * SB + warp + 8 * q][k], xv, t[bj][q]);
        }
      }
    }
    __syncthreads();
    // park T (as T^T) where X_ij will live: S[col][row] of the upper triangle, still unused
#pragma unroll
    for (int bj = 0; bj < 3; ++bj)
      if (bj < bi) {
#pragma unroll
        for (int q = 0; q < 4; ++q) sm.S[bj * SB + lane][bi * SB + warp + 8 * q] = t[bj][q];
      }
    __syncthreads();
    double o[3][4];
#pragma unroll
    for (int bj = 0; bj < 3; ++bj)
#pragma unroll
      for (int q = 0; q < 4; ++q) o[bj][q] = 0.0;
#pragma unroll 4
    for (int k = 0; k < SB; ++k) {
      double xr[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) xr[q] = sm.Xd[bi][warp + 8 * q][k];  // X_ii[r][k], zero for k > r
#pragma unroll
      for (int bj = 0; bj < 3; ++bj)
        if (bj < bi) {
          const double tv = sm.S[bj * SB + lane][bi * SB + k];  // T[k][col]
#pragma unroll
          for (int q = 0; q < 4; ++q) o[bj][q] = fma(xr[q], tv, o[bj][q]);
        }
    }
    __syncthreads();
#pragma unroll
    for (int bj = 0; bj < 3; ++bj)
      if (bj < bi) {
#pragma unroll
        for (int q = 0; q < 4; ++q) sm.S[bj * SB + lane][bi * SB + warp + 8 * q] = -o[bj][q];
      }
    __syncthreads();
  }

  DIAG_T(8);
  // Only the inverse leaves the kernel: nothing reads L_kk again (the panel and the rows of L^-1 multiply by D_k),
  // and writing it cost a third of this phase.  Two columns per thread, 16-byte stores.
#pragma unroll 4
  for (int e = tid; e < NB * NB / 2; e += kDiagThreads) {
    const int i = e >> 6, j = (e & (NB / 2 - 1)) * 2;
    if (i < nb && j < nb) {
      const double x0 = (j < i) ? sm.S[j][i] : (j == i ? sm.dX[i] : 0.0);
      const double x1 = (j + 1 < i) ? sm.S[j + 1][i] : (j + 1 == i ? sm.dX[i] : 0.0);
      if (j + 1 < nb && (ldd & 1) == 0) *reinterpret_cast<double2*>(&Dinv[(size_t)i * ldd + j]) = make_double2(x0, x1);
      else {
        Dinv[(size_t)i * ldd + j] = x0;
        if (j + 1 < nb) Dinv[(size_t)i * ldd + j + 1] = x1;
      }
    }
  }
  DIAG_T(9);
}

// L_ik = A_ik D_k^T for row blocks i = k+1+blockIdx.x (in place).
// In place: a TM=64 sub-tile reads the whole 128-wide row panel of A_ik and writes half of it, so the
// result goes to `out` (a scratch panel) and is copied back by panel_copy_kernel.
template <int TM>
__global__ void __launch_bounds__(kGemmThreads)
panel_trsm_kernel(const double* __restrict__ A, int ld, int k, const double* __restrict__ Dk, double* __restrict__ out,
                  int ldo) {
  __shared__ TileSmem<TM> sm;
  const int r0 = (k + 1 + blockIdx.x) * NB + sub_r<TM>(), c0 = k * NB, cs = sub_c<TM>();
  double acc[TM / 16][TM / 16];
  tile_zero(acc);
  tile_gemm<true, TM>(sm, A, ld, r0, c0, Dk, NB, cs, 0, NB, acc);  // C[r][j] = sum_p A[r][c0+p] D[j][p]
  tile_store<false, TM>(out, ldo, (int)blockIdx.x * NB + sub_r<TM>(), cs, acc, 1.0);
}

// A[(k+1) NB + r][k NB + c] = src[r][c]
__global__ void panel_copy_kernel(double* __restrict__ A, int ld, int k, const double* __restrict__ src, int lds, int rows) {
  const int c = threadIdx.x, r = blockIdx.x * blockDim.y + threadIdx.y;
  if (r < rows) A[(size_t)((k + 1) * NB + r) * ld + k * NB + c] = src[(size_t)r * lds + c];
}

// A_ij -= L_ik L_jk^T over the lower block pairs i >= j > k.
template <int TM>
__global__ void __launch_bounds__(kGemmThreads)
syrk_kernel(double* __restrict__ A, int ld, int k) {
  __shared__ TileSmem<TM> sm;
  // blockIdx.x enumerates pairs (a >= b) of the trailing blocks: t = a(a+1)/2 + b
  int a = (int)((sqrt(8.0 * (double)blockIdx.x + 1.0) - 1.0) * 0.5);
  while ((a + 1) * (a + 2) / 2 <= (int)blockIdx.x) ++a;
  while (a * (a + 1) / 2 > (int)blockIdx.x) --a;
  const int b = blockIdx.x - a * (a + 1) / 2;
  const int r0 = (k + 1 + a) * NB + sub_r<TM>(), c0 = (k + 1 + b) * NB + sub_c<TM>(), kc = k * NB;
  double acc[TM / 16][TM / 16];
  tile_zero(acc);
  tile_gemm<true, TM>(sm, A, ld, r0, kc, A, ld, c0, kc, NB, acc);
  tile_store<true, TM>(A, ld, r0, c0, acc, -1.0);
}

// X_kc = D_k B_kc for c = blockIdx.x < k (in place in Linv: reads rows of block row k, columns of its own
// sub-tile only, and writes that sub-tile... after reading ALL rows of the block row in those columns, so a
// TM=64 sub-tile (upper half rows) would overwrite what the lower-half sub-tile still reads: `out` scratch
// + row_copy_kernel, like the panel); X_kk = D_k.
template <int TM>
__global__ void __launch_bounds__(kGemmThreads)
trtri_row_kernel(const double* __restrict__ X, int ld, int k, const double* __restrict__ Dk, double* __restrict__ out,
                 int ldo) {
  __shared__ TileSmem<TM> sm;
  const int c = blockIdx.x;
  const int r0 = k * NB, c0 = c * NB + sub_c<TM>(), rs = sub_r<TM>();
  if (c == k) {
    for (int e = threadIdx.x; e < TM * TM; e += kGemmThreads) {
      const int i = e / TM, j = e - i * TM;
      out[(size_t)(rs + i) * ldo + c0 + j] = Dk[(size_t)(rs + i) * NB + sub_c<TM>() + j];
    }
    return;
  }
  double acc[TM / 16][TM / 16];
  tile_zero(acc);
  tile_gemm<false, TM>(sm, Dk, NB, rs, 0, X, ld, c0, r0, NB, acc);  // C[r][j] = sum_p D[r][p] B[r0+p][c0+j]
  tile_store<false, TM>(out, ldo, rs, c0, acc, 1.0);
}

// X[k NB + r][c] = src[r][c] for c < cols
__global__ void row_copy_kernel(double* __restrict__ X, int ld, int k, const double* __restrict__ src, int lds, int cols) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (c < cols) X[(size_t)(k * NB + r) * ld + c] = src[(size_t)r * lds + c];
}

// B_ic -= L_ik X_kc for i = k+1+blockIdx.y, c = blockIdx.x <= k.
template <int TM>
__global__ void __launch_bounds__(kGemmThreads)
trtri_update_kernel(double* __restrict__ X, int ld, const double* __restrict__ A, int k) {
  __shared__ TileSmem<TM> sm;
  const int r0 = (k + 1 + blockIdx.y) * NB + sub_r<TM>(), c0 = blockIdx.x * NB + sub_c<TM>(), kc = k * NB;
  double acc[TM / 16][TM / 16];
  tile_zero(acc);
  tile_gemm<false, TM>(sm, A, ld, r0, kc, X, ld, c0, kc, NB, acc);
  tile_store<true, TM>(X, ld, r0, c0, acc, -1.0);
}

#include "factor_dmma.cuh"

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

// S[i][j] -= L21[i] . L21[j] for j <= i; one CTA per pair (one warp over n = 16 384 entries took 114 us).
__global__ void __launch_bounds__(256)
append_schur_kernel(const double* __restrict__ L21, int ld, int n, int k, double* __restrict__ S) {
  __shared__ double red[8];
  const int i = blockIdx.x / k, j = blockIdx.x - i * k;
  if (j > i) return;
  const double* a = L21 + (size_t)i * ld;
  const double* b = L21 + (size_t)j * ld;
  double s = 0.0;
  for (int p = threadIdx.x; p < n; p += 256) s = fma(a[p], b[p], s);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q) t += red[q];
    S[i * kMaxAppend + j] -= t;
  }
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
#pragma unroll 4
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

// One new row (the node's steady state appends one observation per step): the two passes over L^-1 above carry eight
// predicated accumulators (78-80 registers, four 8-byte loads in flight per thread: ~24 KB per SM, 2.5 TB/s at
// n = 16 384).  These keep eight 16-byte loads in flight per thread.  Entries beyond the diagonal are masked, not trusted.
__global__ void __launch_bounds__(256)
append_l21_one_kernel(const double* __restrict__ X, int ld, int n, const double* __restrict__ K21, double* __restrict__ L21) {
  const int j = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (j >= n) return;
  const double2* row = reinterpret_cast<const double2*>(X + (size_t)j * ld);
  const double2* kk = reinterpret_cast<const double2*>(K21);
  const int npairs = (j + 2) >> 1;   // double2 pairs that hold entries 0 .. j
  double s0 = 0.0, s1 = 0.0;
  int q = lane;
  for (; q + 96 < npairs; q += 128) {   // four pairs of loads issued before the first use
    double2 x[4], c[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { x[u] = __ldcs(row + q + 32 * u); c[u] = __ldg(kk + q + 32 * u); }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const bool in = 2 * (q + 32 * u) + 1 <= j;
      s0 = fma(c[u].x, x[u].x, s0);
      s1 = fma(in ? c[u].y : 0.0, in ? x[u].y : 0.0, s1);
    }
  }
  for (; q < npairs; q += 32) {
    const double2 x = __ldcs(row + q), c = __ldg(kk + q);
    const bool in = 2 * q + 1 <= j;
    s0 = fma(c.x, x.x, s0);
    s1 = fma(in ? c.y : 0.0, in ? x.y : 0.0, s1);
  }
  double s = s0 + s1;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
  if (lane == 0) L21[j] = s;
}

// T[c] += sum_{p in split, p >= c} L21[p] Linv[p][c]; a thread owns two neighbouring columns, gridDim.y row splits.
__global__ void __launch_bounds__(128)
append_t_one_kernel(const double* __restrict__ X, int ld, int n, const double* __restrict__ L21, double* __restrict__ T) {
  const int c0 = 2 * (blockIdx.x * 128 + threadIdx.x);
  const int per = (n + gridDim.y - 1) / gridDim.y;
  const int p_lo = blockIdx.y * per, p_hi = min(n, p_lo + per);
  if (c0 >= n || p_hi <= c0) return;
  const double* col = X + c0;
  double s0 = 0.0, s1 = 0.0;
  int p = max(p_lo, c0);
  for (; p + 3 < p_hi; p += 4) {   // four rows' loads issued before the first use
    double2 x[4];
    double l[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      x[u] = __ldcs(reinterpret_cast<const double2*>(col + (size_t)(p + u) * ld));
      l[u] = __ldg(L21 + p + u);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      s0 = fma(l[u], x[u].x, s0);
      s1 = fma(l[u], p + u > c0 ? x[u].y : 0.0, s1);   // (row c0 holds nothing in column c0 + 1)
    }
  }
  for (; p < p_hi; ++p) {
    const double2 x = __ldcs(reinterpret_cast<const double2*>(col + (size_t)p * ld));
    const double l = __ldg(L21 + p);
    s0 = fma(l, x.x, s0);
    s1 = fma(l, p > c0 ? x.y : 0.0, s1);
  }
  atomicAdd(&T[c0], s0);
  if (c0 + 1 < n) atomicAdd(&T[c0 + 1], s1);
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

// alpha after a rank-k append without the two O(n^2) products: with w = L^-1 r,
//   w_new[i] = X[n+i][0 .. n+i] . r           (the old entries of w do not change: L^-1 only gains rows)
//   alpha    = L^-T w  =  alpha_old (0 beyond n)  +  sum_i X[n+i][:]^T w_new[i].
__global__ void __launch_bounds__(256)
append_w_kernel(const double* __restrict__ X, int ld, int n, int k, const double* __restrict__ r, double* __restrict__ w) {
  __shared__ double red[8];
  const int i = blockIdx.x;
  const double* row = X + (size_t)(n + i) * ld;
  double s = 0.0;
  for (int c = threadIdx.x; c <= n + i; c += 256) s = fma(row[c], r[c], s);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q) t += red[q];
    w[n + i] = t;
  }
}

__global__ void __launch_bounds__(256)
append_alpha_kernel(const double* __restrict__ X, int ld, int n, int k, const double* __restrict__ w,
                    double* __restrict__ alpha, float* __restrict__ alpha_f32) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= n + k) return;
  double a = c < n ? alpha[c] : 0.0;
  for (int i = 0; i < k; ++i) a = fma(X[(size_t)(n + i) * ld + c], w[n + i], a);   // X[n+i][c] = 0 for c > n + i
  alpha[c] = a;
  if (alpha_f32) alpha_f32[c] = (float)a;
}

// max over rows [row_begin, row_end) of sum_c |X[row][c]| (c <= row): one warp per row, max by atomicMax on the bit
// pattern (non-negative doubles order like their bits).  Feeds the cut-off of the block-sparse walk (plan.cu).
__global__ void __launch_bounds__(256)
row_l1_max_kernel(const double* __restrict__ X, int ld, int row_begin, int row_end, double* __restrict__ out) {
  const int row = row_begin + ((blockIdx.x * 256 + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (row >= row_end) return;
  const double* x = X + (size_t)row * ld;
  double s = 0.0;
  for (int c = lane; c <= row; c += 32) s += fabs(x[c]);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
  if (lane == 0) atomicMax(reinterpret_cast<unsigned long long*>(out), (unsigned long long)__double_as_longlong(s));
}

// The same for the few rows of an append: one CTA per row.
__global__ void __launch_bounds__(256)
row_l1_max_rows_kernel(const double* __restrict__ X, int ld, int row_begin, double* __restrict__ out) {
  __shared__ double red[8];
  const int row = row_begin + blockIdx.x;
  const double* x = X + (size_t)row * ld;
  double s = 0.0;
  for (int c = threadIdx.x; c <= row; c += 256) s += fabs(x[c]);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q) t += red[q];
    atomicMax(reinterpret_cast<unsigned long long*>(out), (unsigned long long)__double_as_longlong(t));
  }
}

// d_info: [0] Cholesky status (int), [8] max row 1-norm of L^-1 over the rows of a full fit, [16] over the rows of an append
struct InfoRecord {
  int info, pad;
  double r1_full, r1_new;
};

int32_t check_info(sbo_ctx* ctx, const char* what, InfoRecord* rec_out = nullptr) {
  InfoRecord rec;
  SBO_CUDA(ctx, cudaMemcpyAsync(&rec, ctx->d_info, sizeof(rec), cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (rec_out) *rec_out = rec;
  const int info = rec.info;
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

// the shared-memory opt-in is per device: kept per context, not in a process-wide flag
static int32_t ensure_factor_attr(sbo_ctx* ctx) {
  if (ctx->attr_factor) return SBO_OK;
  SBO_CUDA(ctx, cudaFuncSetAttribute(diag_chol_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDiagSmem));
  SBO_CUDA(ctx, cudaFuncSetAttribute(panel_trsm_dm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDmSmem));
  SBO_CUDA(ctx, cudaFuncSetAttribute(syrk_dm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDmSmem));
  SBO_CUDA(ctx, cudaFuncSetAttribute(trtri_row_dm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDmSmem));
  SBO_CUDA(ctx, cudaFuncSetAttribute(trtri_update_dm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDmSmem));
  ctx->attr_factor = true;
  return SBO_OK;
}

static int32_t ensure_factor_streams(sbo_ctx* ctx, int nblk) {
  if (!ctx->aux_stream) SBO_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking));
  while ((int)ctx->ev_factor.size() < 1 + 2 * nblk) {
    cudaEvent_t e;
    SBO_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ctx->ev_factor.push_back(e);
  }
  return SBO_OK;
}

int32_t refresh_planes(sbo_ctx* ctx, int row_begin, int row_end, cudaStream_t stream) {
  if (!ctx->d_planes || row_end <= row_begin) return SBO_OK;
  if (!stream) stream = ctx->stream;
  const int ld = ctx->n_cap;
  dim3 grid((ld + 255) / 256, row_end - row_begin);
  planes_kernel<<<grid, 256, 0, stream>>>(ctx->d_Linv, ld, row_begin, row_end, ld,
                                               reinterpret_cast<__half*>(ctx->d_planes), (size_t)ld * ld,
                                               std::ldexp(1.0, ctx->shift_b));
  SBO_LAUNCH_CHECK(ctx, "planes_kernel");
  return SBO_OK;
}

#include "factor_dist.cuh"

static int32_t factor_full_timed(sbo_ctx* ctx);

int32_t factor_full(sbo_ctx* ctx) {
  // device time of the fit for sbo_fit_info (events on the context's stream; the side streams join it before the end)
  if (!ctx->ev_fit[0]) {
    SBO_CUDA(ctx, cudaEventCreate(&ctx->ev_fit[0]));
    SBO_CUDA(ctx, cudaEventCreate(&ctx->ev_fit[1]));
  }
  ctx->fit_ms = 0.f;
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev_fit[0], ctx->stream));
  const int32_t st = factor_full_timed(ctx);
  if (st == SBO_OK) {
    SBO_CUDA(ctx, cudaEventRecord(ctx->ev_fit[1], ctx->stream));
    SBO_CUDA(ctx, cudaEventSynchronize(ctx->ev_fit[1]));
    SBO_CUDA(ctx, cudaEventElapsedTime(&ctx->fit_ms, ctx->ev_fit[0], ctx->ev_fit[1]));
  }
  return st;
}

static int32_t factor_full_timed(sbo_ctx* ctx) {
  if (ctx->n > 0) {
    int32_t st = SBO_OK;
    if (factor_full_dist(ctx, &st)) return st;   // shared between the ranks / group members (factor_dist.cuh)
  }
  ctx->fit_dist_used = 0;
  const int n = ctx->n, ld = ctx->n_cap;
  const int np = (n + NB - 1) / NB * NB;
  const int nblk = np / NB;
  const double sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  const double inv2l2 = 1.0 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale);
  cudaStream_t st = ctx->stream;

  SBO_TRY(ensure_factor_attr(ctx));
  ctx->linv_r1 = 0.0;
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_info, 0, sizeof(InfoRecord), st));
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
  // Small n only (a launch has at most 2 x SMs 128-tiles: n <= 2944 on 148 SMs; larger fits ran in factor_full_dist above):
  // 64 x 64 FMA tiles, four CTAs per 128-block -- a launch's duration is then one thread's serial FMA count.
  // Two chains: the Cholesky steps (diag -> panel -> trailing update) on the context's stream, the rows of
  // L^-1 (which only need D_k and the finished panel k) on an auxiliary stream, joined at the end.
  SBO_TRY(ensure_factor_streams(ctx, nblk));
  cudaStream_t sb = ctx->aux_stream;
  double* scr_panel = ctx->d_small;                          // [rem * 128][128]
  double* scr_row = ctx->d_small + (size_t)kMaxAppend * ld;  // [128][ld]
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev_factor[0], st));
  SBO_CUDA(ctx, cudaStreamWaitEvent(sb, ctx->ev_factor[0], 0));
  // the first row-block pairs of the posterior start as soon as their rows of L^-1 are final (posterior_tc.cu)
  const int eager_pairs = tc_eager_pairs(ctx);
  const int eager_rows = eager_pairs * 512;
  int planes_done = 0;
  for (int k = 0; k < nblk; ++k) {
    double* Dk = ctx->d_Dinv + (size_t)k * NB * NB;
    cudaEvent_t ev_d = ctx->ev_factor[1 + 2 * k], ev_t = ctx->ev_factor[2 + 2 * k];
    diag_chol_inv_kernel<<<1, kDiagThreads, kDiagSmem, st>>>(ctx->d_A + (size_t)k * NB * ld + (size_t)k * NB, ld, NB,
                                                             Dk, NB, k * NB, ctx->d_info);
    SBO_LAUNCH_CHECK(ctx, "diag_chol_inv_kernel");
    SBO_CUDA(ctx, cudaEventRecord(ev_d, st));
    const int rem = nblk - k - 1;
    if (rem > 0) {
      // (a 64-sub-tile reads the whole 128-wide row panel and writes half of it: result to scratch, then copied back)
      panel_trsm_kernel<64><<<dim3(rem, 1, 4), kGemmThreads, 0, st>>>(ctx->d_A, ld, k, Dk, scr_panel, NB);
      SBO_LAUNCH_CHECK(ctx, "panel_trsm_kernel");
      panel_copy_kernel<<<(rem * NB + 7) / 8, dim3(NB, 8), 0, st>>>(ctx->d_A, ld, k, scr_panel, NB, rem * NB);
      SBO_LAUNCH_CHECK(ctx, "panel_copy_kernel");
      SBO_CUDA(ctx, cudaEventRecord(ev_t, st));
      syrk_kernel<64><<<dim3(rem * (rem + 1) / 2, 1, 4), kGemmThreads, 0, st>>>(ctx->d_A, ld, k);
      SBO_LAUNCH_CHECK(ctx, "syrk_kernel");
    }
    SBO_CUDA(ctx, cudaStreamWaitEvent(sb, ev_d, 0));
    trtri_row_kernel<64><<<dim3(k + 1, 1, 4), kGemmThreads, 0, sb>>>(ctx->d_Linv, ld, k, Dk, scr_row, ld);
    SBO_LAUNCH_CHECK(ctx, "trtri_row_kernel");
    row_copy_kernel<<<dim3(((k + 1) * NB + 255) / 256, NB), 256, 0, sb>>>(ctx->d_Linv, ld, k, scr_row, ld, (k + 1) * NB);
    SBO_LAUNCH_CHECK(ctx, "row_copy_kernel");
    if (eager_rows > 0 && (k + 1) * NB == eager_rows) {   // rows [0, eager_rows) of L^-1 are final on sb
      SBO_TRY(refresh_planes(ctx, 0, eager_rows, sb));
      cudaEvent_t ev_ready = ctx->ev_factor[2 + 2 * (nblk - 1)];   // the last block has no panel: its "panel done" event is free
      SBO_CUDA(ctx, cudaEventRecord(ev_ready, sb));
      if (tc_eager_launch(ctx, eager_pairs, ev_ready) != SBO_OK) {   // best effort: the fit does not depend on it
        cudaGetLastError();
        ctx->eager_inflight = false;
      }
      planes_done = eager_rows;   // converted above either way
    }
    if (rem > 0) {
      SBO_CUDA(ctx, cudaStreamWaitEvent(sb, ev_t, 0));
      trtri_update_kernel<64><<<dim3(k + 1, rem, 4), kGemmThreads, 0, sb>>>(ctx->d_Linv, ld, ctx->d_A, k);
      SBO_LAUNCH_CHECK(ctx, "trtri_update_kernel");
    }
  }
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev_factor[0], sb));
  SBO_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_factor[0], 0));
  if (np > n) {
    clear_pad_kernel<<<(np - n + 127) / 128, 128, 0, st>>>(ctx->d_Linv, ld, n, np);
    SBO_LAUNCH_CHECK(ctx, "clear_pad_kernel");
  }
  row_l1_max_kernel<<<(n * 32 + 255) / 256, 256, 0, st>>>(ctx->d_Linv, ld, 0, n, reinterpret_cast<double*>(ctx->d_info) + 1);
  SBO_LAUNCH_CHECK(ctx, "row_l1_max_kernel");
  InfoRecord rec;
  SBO_TRY(check_info(ctx, "sbo_set_observations", &rec));
  ctx->linv_r1 = rec.r1_full;
  SBO_TRY(compute_alpha(ctx, n));
  SBO_TRY(refresh_planes(ctx, planes_done, ld));   // (rows the eager launch is reading were converted for it above)
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

  SBO_TRY(ensure_factor_attr(ctx));   // (an append can be the first factor call of a context)
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_info, 0, sizeof(int), st));
  SBO_CUDA(ctx, cudaMemsetAsync(reinterpret_cast<double*>(ctx->d_info) + 2, 0, sizeof(double), st));
  {
    dim3 g((n + k + 127) / 128, k);
    append_cov_kernel<<<g, 128, 0, st>>>(ctx->d_xy, n, k, sf2, inv2l2, ctx->cfg.noise_var, K21, ld, S);
    SBO_LAUNCH_CHECK(ctx, "append_cov_kernel");
  }
  if (n > 0) {
    if (k == 1) append_l21_one_kernel<<<(n * 32 + 255) / 256, 256, 0, st>>>(ctx->d_Linv, ld, n, K21, L21);
    else append_l21_kernel<<<(n * 32 + 255) / 256, 256, 0, st>>>(ctx->d_Linv, ld, n, k, K21, L21);
    SBO_LAUNCH_CHECK(ctx, "append_l21_kernel");
    append_schur_kernel<<<k * k, 256, 0, st>>>(L21, ld, n, k, S);
    SBO_LAUNCH_CHECK(ctx, "append_schur_kernel");
  }
  // A Schur-complement pivot K22 - |l21|^2 below a few hundred roundings of K22 is cancellation noise, whatever its
  // sign: the new observation is numerically a combination of those held (a duplicate under a vanishing noise_var),
  // and the rows of L^-1 it would produce are garbage.  Rejected as not positive definite, independent of the summation
  // order of the passes above.  (noise_var >= 1e-13 sigma_f^2 never comes near: the pivot is >= noise_var.)
  const double min_pivot = 256.0 * 2.220446049250313e-16 * (sf2 + ctx->cfg.noise_var);
  diag_chol_inv_kernel<<<1, kDiagThreads, kDiagSmem, st>>>(S, kMaxAppend, k, X22, kMaxAppend, n, ctx->d_info, min_pivot);
  SBO_LAUNCH_CHECK(ctx, "diag_chol_inv_kernel(append)");
  // The positive-definiteness verdict is read once, at the end (one host synchronisation per append instead of two):
  // until then only rows >= n of L^-1 (not part of the state) and alpha are written, and alpha is restored below.
  double* T = K21;
  if (n > 0) {
    SBO_CUDA(ctx, cudaMemsetAsync(T, 0, sizeof(double) * (size_t)k * ld, st));
    int splits = (ctx->sm_count * 16) / ((n + 127) / 128);   // ~2048 resident threads per SM: one load in flight each
    if (splits < 1) splits = 1;
    if (splits > 64) splits = 64;
    if (k == 1) {   // two columns per thread: half the column blocks, the same number of CTAs
      const int bx = (n + 255) / 256;
      splits = std::max(1, std::min(64, (ctx->sm_count * 16) / bx));
      append_t_one_kernel<<<dim3(bx, splits), 128, 0, st>>>(ctx->d_Linv, ld, n, L21, T);
    } else {
      dim3 g((n + 127) / 128, splits);
      append_t_kernel<<<g, 128, 0, st>>>(ctx->d_Linv, ld, n, k, L21, T);
    }
    SBO_LAUNCH_CHECK(ctx, "append_t_kernel");
  }
  {
    dim3 g((n + k + 255) / 256, k);
    append_rows_kernel<<<g, 256, 0, st>>>(ctx->d_Linv, ld, n, k, T, X22);
    SBO_LAUNCH_CHECK(ctx, "append_rows_kernel");
  }
  if (n > 0) {   // alpha by the O(n k) update; the first fit of an empty context goes the general way
    append_w_kernel<<<k, 256, 0, st>>>(ctx->d_Linv, ld, n, k, ctx->d_r, ctx->d_w);
    SBO_LAUNCH_CHECK(ctx, "append_w_kernel");
    append_alpha_kernel<<<(n + k + 255) / 256, 256, 0, st>>>(ctx->d_Linv, ld, n, k, ctx->d_w, ctx->d_alpha, ctx->d_alpha_f32);
    SBO_LAUNCH_CHECK(ctx, "append_alpha_kernel");
  } else {
    SBO_TRY(compute_alpha(ctx, n + k));
  }
  SBO_TRY(refresh_planes(ctx, n, n + k));
  row_l1_max_rows_kernel<<<k, 256, 0, st>>>(ctx->d_Linv, ld, n, reinterpret_cast<double*>(ctx->d_info) + 2);
  SBO_LAUNCH_CHECK(ctx, "row_l1_max_kernel");
  InfoRecord rec;
  const int32_t verdict = check_info(ctx, "sbo_append_observations", &rec);   // synchronises the stream
  if (verdict == SBO_OK && rec.r1_new > ctx->linv_r1) ctx->linv_r1 = rec.r1_new;
  if (verdict != SBO_OK) {   // not positive definite: the new rows were computed from garbage
    // rows [n, n + k) of L^-1 and their fp16 planes go back to zero (the invariant the posterior kernels rely on:
    // L^-1 is zero beyond n), alpha back to the old observations'
    const std::string why = ctx->err;
    SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_Linv + (size_t)n * ld, 0, sizeof(double) * (size_t)k * ld, st));
    SBO_TRY(refresh_planes(ctx, n, n + k));
    ctx->err = why;
    SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_alpha + n, 0, sizeof(double) * (size_t)k, st));
    if (ctx->d_alpha_f32) SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_alpha_f32 + n, 0, sizeof(float) * (size_t)k, st));
    if (n > 0) SBO_TRY(compute_alpha(ctx, n));
    SBO_CUDA(ctx, cudaStreamSynchronize(st));
  }
  return verdict;
}

}  // namespace sbo

#ifdef SBO_DIAG_TIMING
extern "C" int32_t sbo_debug_diag_timing(long long* host16) {
  return cudaMemcpyFromSymbol(host16, sbo::g_diag_t, sizeof(long long) * 16) == cudaSuccess ? 0 : -3;
}
#endif
