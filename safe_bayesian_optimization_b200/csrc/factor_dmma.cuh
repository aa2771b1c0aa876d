// factor_dmma.cuh -- the 128 x 128 x 128 fp64 tile product of the factorisation on the DMMA path (included by factor.cu,
// inside namespace sbo::{anonymous}).
//
// Round 2 measurement (profiles/launches_bench_c4_r02e.csv): at n = 16 384 the fit is 207 ms, 91 % of it in the trailing
// update and the L^-1 update, i.e. 13 TFLOP/s of fp64 from a register-tile FMA routine with an un-pipelined 16-deep
// shared-memory stage; posterior_f64.cu holds 28 TFLOP/s on the same pipe with mma.sync.m8n8k4.f64 and cp.async double
// buffering.  Same recipe here: 512 threads = 16 warps as 4 x 4, warp tile 32 x 32 = 4 x 4 DMMA tiles (32 accumulators per
// thread), K = 128 in four slabs of 32, two slab buffers, one cp.async group in flight while the other is multiplied.
// Operand A is always K-contiguous (rows of a row-major matrix); operand B is either K-contiguous too (kNT: C = A B^T) or a
// row-major K x N block (C = A B).  Slab rows are padded (36 / 132 doubles) so that the 16 fragment reads of a half-warp
// (row or column g = 0..3, k = t = 0..3) fall in 16 different 8-byte banks.
//
// Used by the large-n fit (more 128-tiles per launch than 2 x SMs), replicated or shared; small n keeps the 64 x 64 FMA
// tiles (four times the CTAs per launch, where a launch is one thread's serial FMA count anyway).

constexpr int kDmThreads = 512;
constexpr int kDmSlab = 32;                 // K per slab
constexpr int kDmLdK = kDmSlab + 4;         // padded row of a K-contiguous slab  [128][36]
constexpr int kDmLdN = NB + 4;              // padded row of a row-major K x N slab [32][132]
constexpr int kDmBufDoubles = NB * kDmLdK;  // 4608 >= 32 * 132 = 4224
constexpr size_t kDmSmem = sizeof(double) * 4 * kDmBufDoubles;   // A0 A1 B0 B1 = 147 456 B

__device__ __forceinline__ void dm_cp16(void* smem_dst, const void* gmem_src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void dm_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void dm_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void dm_mma(double (&c)[2], double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// slab of a K-contiguous operand: rows r0 .. r0 + 127 of M, columns k0 .. k0 + 31  ->  s[row][36]
__device__ __forceinline__ void dm_load_kmajor(double* s, const double* __restrict__ M, int ld, int r0, int k0) {
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int c = threadIdx.x + u * kDmThreads, row = c >> 4, kq = (c & 15) * 2;
    dm_cp16(s + row * kDmLdK + kq, M + (size_t)(r0 + row) * ld + k0 + kq);
  }
}
// slab of a row-major K x N operand: rows k0 .. k0 + 31 of M, columns c0 .. c0 + 127  ->  s[k][132]
__device__ __forceinline__ void dm_load_nmajor(double* s, const double* __restrict__ M, int ld, int k0, int c0) {
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int c = threadIdx.x + u * kDmThreads, kk = c >> 6, nq = (c & 63) * 2;
    dm_cp16(s + kk * kDmLdN + nq, M + (size_t)(k0 + kk) * ld + c0 + nq);
  }
}

// acc[mi][ni] += A[r0.., ka0 .. ka0 + 127] * op(B):  kNT: B rows c0.., columns kb0..;  else B rows kb0.., columns c0..
template <bool kNT>
__device__ __forceinline__ void dm_gemm128(double* smem, const double* __restrict__ A, int lda, int r0, int ka0,
                                           const double* __restrict__ B, int ldb, int c0, int kb0, double (&acc)[4][4][2]) {
  auto sa = [&](int buf) { return smem + buf * kDmBufDoubles; };
  auto sb = [&](int buf) { return smem + (2 + buf) * kDmBufDoubles; };
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wm = warp >> 2, wn = warp & 3, g = lane >> 2, t = lane & 3;
  auto load = [&](int s, int buf) {
    dm_load_kmajor(sa(buf), A, lda, r0, ka0 + s * kDmSlab);
    if (kNT) dm_load_kmajor(sb(buf), B, ldb, c0, kb0 + s * kDmSlab);
    else dm_load_nmajor(sb(buf), B, ldb, kb0 + s * kDmSlab, c0);
    dm_commit();
  };
  load(0, 0);
#pragma unroll 1
  for (int s = 0; s < NB / kDmSlab; ++s) {
    if (s + 1 < NB / kDmSlab) {
      load(s + 1, (s + 1) & 1);
      dm_wait<1>();
    } else {
      dm_wait<0>();
    }
    __syncthreads();
    const double* pa = sa(s & 1) + (wm * 32 + g) * kDmLdK + t;
    const double* pb = kNT ? sb(s & 1) + (wn * 32 + g) * kDmLdK + t : sb(s & 1) + t * kDmLdN + wn * 32 + g;
#pragma unroll
    for (int kk = 0; kk < kDmSlab / 4; ++kk) {
      double a[4], b[4];
#pragma unroll
      for (int mi = 0; mi < 4; ++mi) a[mi] = pa[mi * 8 * kDmLdK + kk * 4];
#pragma unroll
      for (int ni = 0; ni < 4; ++ni) b[ni] = kNT ? pb[ni * 8 * kDmLdK + kk * 4] : pb[kk * 4 * kDmLdN + ni * 8];
#pragma unroll
      for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) dm_mma(acc[mi][ni], a[mi], b[ni]);
    }
    __syncthreads();   // the buffer multiplied here is the target of the load issued in the next iteration
  }
}

__device__ __forceinline__ void dm_zero(double (&acc)[4][4][2]) {
#pragma unroll
  for (int mi = 0; mi < 4; ++mi)
#pragma unroll
    for (int ni = 0; ni < 4; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
}

// C[r0.., c0..] = sign * acc (+ C if kAdd); lane (g, t) of warp (wm, wn) holds rows wm 32 + mi 8 + g, columns wn 32 + ni 8 + 2 t, + 1
template <bool kAdd>
__device__ __forceinline__ void dm_store(double* __restrict__ C, int ldc, int r0, int c0, const double (&acc)[4][4][2], double sign) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wm = warp >> 2, wn = warp & 3, g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int mi = 0; mi < 4; ++mi)
#pragma unroll
    for (int ni = 0; ni < 4; ++ni) {
      double2* p = reinterpret_cast<double2*>(C + (size_t)(r0 + wm * 32 + mi * 8 + g) * ldc + c0 + wn * 32 + ni * 8 + 2 * t);
      double2 v = kAdd ? *p : make_double2(0.0, 0.0);
      v.x = fma(sign, acc[mi][ni][0], v.x);
      v.y = fma(sign, acc[mi][ni][1], v.y);
      *p = v;
    }
}

// snake deal of block columns to the ranks of a shared fit (factor_dist.cuh); W = 1: every column, in order
struct ColMap {
  int W, me;
  __host__ __device__ int col(int j) const { return 2 * W * (j >> 1) + ((j & 1) ? 2 * W - 1 - me : me); }   // j-th column of `me`
  __host__ __device__ static int owner(int c, int W) {
    const int t = c % (2 * W);
    return t < W ? t : 2 * W - 1 - t;
  }
  __host__ int count_le(int k) const {   // how many of my columns are <= k
    const int t = (k + 1) % (2 * W);
    return 2 * ((k + 1) / (2 * W)) + (me < t ? 1 : 0) + (2 * W - 1 - me < t ? 1 : 0);
  }
};

// L_ik = A_ik D_k^T for row blocks i = k + 1 + blockIdx.x, in place (every read of the block precedes its stores)
__global__ void __launch_bounds__(kDmThreads, 1)
panel_trsm_dm_kernel(double* __restrict__ A, int ld, int k, const double* __restrict__ Dk) {
  extern __shared__ __align__(16) double dm_smem[];
  const int r0 = (k + 1 + (int)blockIdx.x) * NB, c0 = k * NB;
  double acc[4][4][2];
  dm_zero(acc);
  dm_gemm128<true>(dm_smem, A, ld, r0, c0, Dk, NB, 0, 0, acc);   // C[r][j] = sum_p A[r][c0 + p] D[j][p]
  dm_store<false>(A, ld, r0, c0, acc, 1.0);
}

// A_ab -= L_ak L_bk^T for a = k + 1 + blockIdx.x >= b, b = the (j0 + blockIdx.y)-th block column of this rank (> k)
__global__ void __launch_bounds__(kDmThreads, 1)
syrk_dm_kernel(double* __restrict__ A, int ld, int k, ColMap map, int j0) {
  extern __shared__ __align__(16) double dm_smem[];
  const int b = map.col(j0 + (int)blockIdx.y), a = k + 1 + (int)blockIdx.x;
  if (a < b) return;
  const int r0 = a * NB, c0 = b * NB, kc = k * NB;
  double acc[4][4][2];
  dm_zero(acc);
  dm_gemm128<true>(dm_smem, A, ld, r0, kc, A, ld, c0, kc, acc);
  dm_store<true>(A, ld, r0, c0, acc, -1.0);
}

// X_kc = D_k B_kc (X_kk = D_k) for c = the blockIdx.x-th block column of this rank (<= k), in place
__global__ void __launch_bounds__(kDmThreads, 1)
trtri_row_dm_kernel(double* __restrict__ X, int ld, int k, const double* __restrict__ Dk, ColMap map) {
  extern __shared__ __align__(16) double dm_smem[];
  const int c = map.col((int)blockIdx.x);
  const int r0 = k * NB, c0 = c * NB;
  if (c == k) {
    for (int e = threadIdx.x; e < NB * NB; e += kDmThreads) {
      const int i = e / NB, j = e - i * NB;
      X[(size_t)(r0 + i) * ld + c0 + j] = Dk[(size_t)i * NB + j];
    }
    return;
  }
  double acc[4][4][2];
  dm_zero(acc);
  dm_gemm128<false>(dm_smem, Dk, NB, 0, 0, X, ld, c0, r0, acc);   // C[r][j] = sum_p D[r][p] B[r0 + p][c0 + j]
  dm_store<false>(X, ld, r0, c0, acc, 1.0);
}

// B_ic -= L_ik X_kc for i = k + 1 + blockIdx.y, c = the blockIdx.x-th block column of this rank (<= k)
__global__ void __launch_bounds__(kDmThreads, 1)
trtri_update_dm_kernel(double* __restrict__ X, int ld, const double* __restrict__ A, int k, ColMap map) {
  extern __shared__ __align__(16) double dm_smem[];
  const int r0 = (k + 1 + (int)blockIdx.y) * NB, c0 = map.col((int)blockIdx.x) * NB, kc = k * NB;
  double acc[4][4][2];
  dm_zero(acc);
  dm_gemm128<false>(dm_smem, A, ld, r0, kc, X, ld, c0, kc, acc);
  dm_store<true>(X, ld, r0, c0, acc, -1.0);
}
