// posterior_simt.cu -- GP posterior over the cells of this shard with plain FMA pipes:
//   mu(x*)     = prior_mean + k*^T alpha
//   sigma^2(x*) = sigma_f^2 - | L^-1 k* |^2
// This is the fp64-mode kernel (tcgen05 has no fp64 kind) and, instantiated for float, the
// independent cross-check of the tcgen05 kernel.  It is the "predict" half of the external GP
// node the library replaces (reference src/safe_bayesian_optimization_node.cpp:568-644 receives
// exactly these two arrays as `values` / `uncertainties`).
//
// A CTA owns 128 consecutive cells.  RBF cross-covariance entries k(x*, x_k) are generated on
// the fly into shared memory (never stored in HBM), the matching 128 x 16 slab of L^-1 is
// staged beside them, and each thread keeps an 8x8 register tile of V = K* L^-T.  When a block
// of 128 rows of L^-1 is finished the tile is squared and row-summed in fp64; V itself never
// leaves the registers.  Rows j of L^-1 only see k <= j (lower triangular), so the k loop of
// row block jb stops at its diagonal.
#include "common.cuh"

namespace sbo {
namespace {

constexpr int TM = 128, TN = 128, TK = 16, kThreads = 256;

struct SimtParams {
  const double* xy;     // [n][2]
  const double* Linv;   // [n_cap][ld]
  const double* alpha;  // [n]
  int n, ld;
  int j_lo, j_hi;       // rows of L^-1 to contract
  int accumulate;       // 1: sigma^2 -= |rows|^2 on top of the stored value
  double x0, y0, rx, ry;
  int W;
  long long cell0, m;
  double sf2, inv2l2, prior_mean;
};

template <typename T> __device__ __forceinline__ T rbf_entry(double dx, double dy, double sf2, double inv2l2);
template <> __device__ __forceinline__ double rbf_entry<double>(double dx, double dy, double sf2, double inv2l2) {
  return sf2 * exp(-(dx * dx + dy * dy) * inv2l2);
}
template <> __device__ __forceinline__ float rbf_entry<float>(double dx, double dy, double sf2, double inv2l2) {
  const float fx = (float)dx, fy = (float)dy;  // the difference is formed in fp64, so it is exact to fp32 ulp
  return (float)sf2 * expf(-(fx * fx + fy * fy) * (float)inv2l2);
}

template <typename T> struct Vec2;
template <> struct Vec2<float> { using type = float2; };
template <> struct Vec2<double> { using type = double2; };

template <typename T>
struct Smem {
  alignas(16) T a[TK][TM];  // K* chunk, a[kk][cell]
  alignas(16) T b[TK][TN];  // L^-1 chunk, b[kk][j]
  double mu_part[2][TM];
};

template <typename T>
__global__ void __launch_bounds__(kThreads)
posterior_simt_kernel(SimtParams p, T* __restrict__ mu_out, T* __restrict__ s2_out) {
  __shared__ Smem<T> sm;
  const int tid = threadIdx.x;
  const int ty = tid >> 4, tx = tid & 15;
  const int gc = tid & (TM - 1), gk = (tid >> 7) * 8;  // generator role: cell gc, k offsets gk..gk+7
  const long long ntiles = (p.m + TM - 1) / TM;
  const int jb_lo = p.j_lo / TN, jb_hi = (p.j_hi + TN - 1) / TN;  // row blocks [jb_lo, jb_hi)
  const int n_pad = (p.n + TK - 1) / TK * TK;

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long lc = tile * TM + gc;  // local cell of the generator role
    const bool cell_ok = lc < p.m;
    double cx = 0.0, cy = 0.0;
    if (cell_ok) {
      const long long cell = p.cell0 + lc;
      const int iy = (int)(cell / p.W), ix = (int)(cell - (long long)iy * p.W);
      cx = cell_coord(p.x0, p.rx, ix);
      cy = cell_coord(p.y0, p.ry, iy);
    }
    double rs[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) rs[i] = 0.0;
    double mu_acc = 0.0;

    for (int jb = jb_lo; jb < jb_hi; ++jb) {
      const int j0 = jb * TN;
      const bool last = (jb == jb_hi - 1);
      // the last block walks every k so that mu sees all observations (L^-1 is zero beyond the diagonal)
      int kd = last ? n_pad : min(n_pad, j0 + TN);
      T acc[8][8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = (T)0;

      for (int k0 = 0; k0 < kd; k0 += TK) {
        // ---- generate K*[cells, k0..k0+16) ----
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int k = k0 + gk + e;
          T v = (T)0;
          if (cell_ok && k < p.n) {
            const double2 o = *reinterpret_cast<const double2*>(p.xy + 2 * (size_t)k);
            v = rbf_entry<T>(cx - o.x, cy - o.y, p.sf2, p.inv2l2);
            if (last) mu_acc = fma((double)v, p.alpha[k], mu_acc);
          }
          sm.a[gk + e][gc] = v;
        }
        // ---- stage L^-1[j0..j0+128, k0..k0+16) transposed; rows outside [j_lo, j_hi) count as zero ----
        {
          const int row = tid >> 1, kq = (tid & 1) * 8;
          const int j = j0 + row;
          if (j >= p.j_lo && j < p.j_hi && k0 + kq <= j) {
            const double2* src = reinterpret_cast<const double2*>(p.Linv + (size_t)j * p.ld + k0 + kq);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const double2 v = src[q];
              sm.b[kq + 2 * q][row] = (T)v.x;
              sm.b[kq + 2 * q + 1][row] = (T)v.y;
            }
          } else {
#pragma unroll
            for (int q = 0; q < 8; ++q) sm.b[kq + q][row] = (T)0;
          }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < TK; ++kk) {
          T a[8], b[8];
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            using V2 = typename Vec2<T>::type;
            const V2 av = *reinterpret_cast<const V2*>(&sm.a[kk][g * 32 + ty * 2]);
            const V2 bv = *reinterpret_cast<const V2*>(&sm.b[kk][g * 32 + tx * 2]);
            a[2 * g] = av.x; a[2 * g + 1] = av.y;
            b[2 * g] = bv.x; b[2 * g + 1] = bv.y;
          }
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
      }
      // ---- fused epilogue: squared row sums of this 128 x 128 block of V, in fp64 ----
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < 8; ++j) s = fma((double)acc[i][j], (double)acc[i][j], s);
        rs[i] += s;
      }
    }

    // ---- finish the tile: reduce over the 16 lanes that share a row, write mu / sigma^2 ----
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int d = 8; d > 0; d >>= 1) rs[i] += __shfl_xor_sync(0xffffffffu, rs[i], d);
    }
    sm.mu_part[tid >> 7][gc] = mu_acc;
    __syncthreads();
    if (tx == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = (i >> 1) * 32 + ty * 2 + (i & 1);
        const long long c = tile * TM + r;
        if (c < p.m) {
          const double prev = p.accumulate ? (double)s2_out[c] : p.sf2;
          s2_out[c] = (T)(prev - rs[i]);
        }
      }
    }
    if (tid < TM && cell_ok) mu_out[lc] = (T)(p.prior_mean + sm.mu_part[0][gc] + sm.mu_part[1][gc]);
    __syncthreads();  // mu_part is reused by the next tile
  }
}

}  // namespace

int32_t launch_posterior_simt(sbo_ctx* ctx, int j_lo, int j_hi, bool accumulate) {
  SimtParams p;
  p.xy = ctx->d_xy; p.Linv = ctx->d_Linv; p.alpha = ctx->d_alpha;
  p.n = ctx->n; p.ld = ctx->n_cap; p.j_lo = j_lo; p.j_hi = j_hi; p.accumulate = accumulate ? 1 : 0;
  p.x0 = ctx->x0; p.y0 = ctx->y0; p.rx = ctx->rx; p.ry = ctx->ry; p.W = ctx->W;
  p.cell0 = ctx->cell0; p.m = ctx->m_local;
  p.sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  p.inv2l2 = 1.0 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale);
  p.prior_mean = ctx->cfg.prior_mean;
  if (j_hi <= j_lo) { p.j_lo = 0; p.j_hi = 0; }
  const long long ntiles = (p.m + TM - 1) / TM;
  if (ntiles == 0) return SBO_OK;
  // one resident CTA per SM for fp64 (register tile), two for fp32; grid = multiple of the SM count
  const int per_sm = ctx->cfg.precision == SBO_FP64 ? 1 : 2;
  long long grid = (long long)ctx->sm_count * per_sm;
  if (grid > ntiles) grid = ntiles;
  if (ctx->cfg.precision == SBO_FP64) {
    posterior_simt_kernel<double><<<(unsigned)grid, kThreads, 0, ctx->stream>>>(
        p, static_cast<double*>(ctx->d_mu), static_cast<double*>(ctx->d_s2));
  } else {
    posterior_simt_kernel<float><<<(unsigned)grid, kThreads, 0, ctx->stream>>>(
        p, static_cast<float*>(ctx->d_mu), static_cast<float*>(ctx->d_s2));
  }
  SBO_LAUNCH_CHECK(ctx, "posterior_simt_kernel");
  return SBO_OK;
}

}  // namespace sbo
