// posterior_thin.cu -- the incremental BO step after a small rank-k append (k <= 8), in O(m n k).
//
// After sbo_append_observations the factor gains k rows:  L^-1 = [[L11^-1, 0], [R, L22^-1]].  For every cell
//   sigma^2_new = sigma^2_old - sum_r u_r^2,   u_r = rows_new[r] . k*      (sequential conditioning)
//   mu_new      = prior_mean + k* . alpha_new                              (alpha changes everywhere)
// so one pass over the n observations per cell yields both.  This is what the ROS node's steady state looks
// like (one measurement per second, src/turtlesim_spatial_publisher.py:43) and what BASELINE configs C3/C5 time.
// The contraction has N = k (<= 8) columns: far too thin for a tensor-core tile (the 256-row block kernel of
// posterior_tc.cu spends 256/k times the necessary MMAs on it), and its cost is the K* entries themselves --
// one ex2 (fp32) or exp (fp64) per cell and observation.  Hence a SIMT kernel: one thread owns C cells of a grid
// row, a CTA stages 256 observations at a time (tile-relative coordinate, y factor, alpha, the k new rows of
// L^-1) in shared memory and every thread reads them as broadcast 16-byte vectors.
//
// Numerics: fp32 mode accumulates in fp32 over 64 observations and in fp64 across (an fp32 running sum over
// n = 16 384 terms would cost ~1e-5); coordinates are tile-relative fp32 (hi, lo) differences as in the
// tcgen05 kernel.  fp64 mode is plain fp64.
#include <algorithm>
#include <cmath>

#include "common.cuh"

namespace sbo {
namespace {

constexpr int kThinThreads = 256;
constexpr int kChunk = 256;   // observations staged per pass
constexpr int kSub = 64;      // fp32 accumulation span

struct ThinParams {
  const float4* obs_f32;   // fp32 mode: scaled (x_hi, x_lo, y_hi, y_lo)
  const double* xy;        // fp64 mode
  const double* Linv;      // [n_cap][ld] fp64
  const double* alpha;     // [n] fp64
  int n, ld;
  int row_lo, k_new;       // new rows [row_lo, row_lo + k_new)
  int W, rows_local, row0;
  double x0, y0, rx, ry;
  double xscale;           // fp32 mode: sqrt(log2(e) / 2 l^2)
  double sf2, inv2l2, prior_mean;
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// four consecutive staged values as 16-byte shared-memory loads (every lane reads the same address: broadcast)
__device__ __forceinline__ void load4(const float* p, float (&v)[4]) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4(const double* p, double (&v)[4]) {
  const double2 t0 = *reinterpret_cast<const double2*>(p), t1 = *reinterpret_cast<const double2*>(p + 2);
  v[0] = t0.x; v[1] = t0.y; v[2] = t1.x; v[3] = t1.y;
}

template <typename T, int KR>
struct __align__(16) ThinSmem {
  T x[kChunk];       // tile-relative x of the observation (fp32 mode: scaled)
  T e[kChunk];       // y factor: sigma_f^2 exp(-dy^2 / 2 l^2), 0 beyond n
  T a[kChunk];       // alpha
  T l[KR][kChunk];   // the new rows of L^-1
};

// K* entry from the tile-relative x difference and the staged y factor
__device__ __forceinline__ float kstar(float dx, float e, double) { return ex2f(-dx * dx) * e; }
__device__ __forceinline__ double kstar(double dx, double e, double inv2l2) { return exp(-dx * dx * inv2l2) * e; }

template <typename T, int KR, int C>
__global__ void __launch_bounds__(kThinThreads)
posterior_thin_kernel(ThinParams p, T* __restrict__ mu_out, T* __restrict__ s2_out) {
  __shared__ ThinSmem<T, KR> sm;
  constexpr bool kF32 = sizeof(T) == 4;
  const int tid = threadIdx.x;
  const int span = kThinThreads * C;
  const int tiles_per_row = (p.W + span - 1) / span;
  const long long ntiles = (long long)p.rows_local * tiles_per_row;

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int lrow = (int)(tile / tiles_per_row), col0 = (int)(tile % tiles_per_row) * span;
    // tile origin; fp32 mode works in scaled units with (hi, lo) pairs, fp64 mode in plain coordinates
    const double ox = cell_coord(p.x0, p.rx, col0), oy = cell_coord(p.y0, p.ry, p.row0 + lrow);
    const double oxs = ox * p.xscale, oys = oy * p.xscale;
    const float oxh = (float)oxs, oxl = (float)(oxs - (double)oxh);
    const float oyh = (float)oys, oyl = (float)(oys - (double)oyh);
    T cx[C];
#pragma unroll
    for (int q = 0; q < C; ++q) {
      const double off = (double)(q * kThinThreads + tid) * p.rx;
      cx[q] = kF32 ? (T)(off * p.xscale) : (T)off;
    }
    double mu[C], u[C][KR];
#pragma unroll
    for (int q = 0; q < C; ++q) {
      mu[q] = 0.0;
#pragma unroll
      for (int r = 0; r < KR; ++r) u[q][r] = 0.0;
    }

    for (int k0 = 0; k0 < p.n; k0 += kChunk) {
      // ---- stage observations k0 .. k0+255 ----
      {
        const int j = k0 + tid;
        T tx = (T)0, te = (T)0, ta = (T)0;
        if (j < p.n) {
          if (kF32) {
            const float4 o = p.obs_f32[j];
            const float yr = (o.z - oyh) + (o.w - oyl);
            tx = (T)((o.x - oxh) + (o.y - oxl));
            te = (T)(ex2f(-yr * yr) * (float)p.sf2);
          } else {
            const double2 o = *reinterpret_cast<const double2*>(p.xy + 2 * (size_t)j);
            const double dy = o.y - oy;
            tx = (T)(o.x - ox);
            te = (T)(p.sf2 * exp(-dy * dy * p.inv2l2));
          }
          ta = (T)p.alpha[j];
        }
        sm.x[tid] = tx; sm.e[tid] = te; sm.a[tid] = ta;
#pragma unroll
        for (int r = 0; r < KR; ++r)
          sm.l[r][tid] = (j < p.n && r < p.k_new) ? (T)p.Linv[(size_t)(p.row_lo + r) * p.ld + j] : (T)0;
      }
      __syncthreads();
      const int cnt = min(kChunk, p.n - k0);
      for (int s0 = 0; s0 < cnt; s0 += kSub) {
        T fmu[C], fu[C][KR];
#pragma unroll
        for (int q = 0; q < C; ++q) {
          fmu[q] = (T)0;
#pragma unroll
          for (int r = 0; r < KR; ++r) fu[q][r] = (T)0;
        }
#pragma unroll 2
        for (int j4 = s0; j4 < s0 + kSub; j4 += 4) {  // the chunk is zero-padded: no bound check inside
          T xs[4], es[4], as[4], ls[KR][4];
          load4(&sm.x[j4], xs);
          load4(&sm.e[j4], es);
          load4(&sm.a[j4], as);
#pragma unroll
          for (int r = 0; r < KR; ++r) load4(&sm.l[r][j4], ls[r]);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int q = 0; q < C; ++q) {
              const T kv = kstar(xs[i] - cx[q], es[i], p.inv2l2);
              fmu[q] = fma(kv, as[i], fmu[q]);
#pragma unroll
              for (int r = 0; r < KR; ++r) {
                fu[q][r] = fma(kv, ls[r][i], fu[q][r]);
              }
            }
          }
        }
#pragma unroll
        for (int q = 0; q < C; ++q) {
          mu[q] += (double)fmu[q];
#pragma unroll
          for (int r = 0; r < KR; ++r) u[q][r] += (double)fu[q][r];
        }
      }
      __syncthreads();
    }

#pragma unroll
    for (int q = 0; q < C; ++q) {
      const int col = col0 + q * kThinThreads + tid;
      if (col < p.W) {
        const long long lc = (long long)lrow * p.W + col;
        double red = 0.0;
#pragma unroll
        for (int r = 0; r < KR; ++r) red = fma(u[q][r], u[q][r], red);
        s2_out[lc] = (T)((double)s2_out[lc] - red);
        mu_out[lc] = (T)(p.prior_mean + mu[q]);
      }
    }
  }
}

template <typename T, int KR, int C>
void launch_thin(sbo_ctx* ctx, const ThinParams& p) {
  const int span = kThinThreads * C;
  const long long ntiles = (long long)p.rows_local * ((p.W + span - 1) / span);
  const unsigned grid = (unsigned)std::min<long long>(ntiles, (long long)ctx->sm_count * 8);
  posterior_thin_kernel<T, KR, C><<<grid, kThinThreads, 0, ctx->stream>>>(p, static_cast<T*>(ctx->d_mu), static_cast<T*>(ctx->d_s2));
}

}  // namespace

// Incremental posterior for new rows [row_lo, n) of L^-1, n - row_lo <= 8.
int32_t launch_posterior_thin(sbo_ctx* ctx, int row_lo) {
  if (ctx->m_local == 0) return SBO_OK;
  const int k_new = ctx->n - row_lo;
  if (k_new < 0 || k_new > kThinMaxRows) return fail(ctx, SBO_ERR_STATE, "posterior_thin: more than 8 new rows");
  ThinParams p;
  p.obs_f32 = reinterpret_cast<const float4*>(ctx->d_obs_f32);
  p.xy = ctx->d_xy; p.Linv = ctx->d_Linv; p.alpha = ctx->d_alpha;
  p.n = ctx->n; p.ld = ctx->n_cap; p.row_lo = row_lo; p.k_new = k_new;
  p.W = ctx->W; p.rows_local = ctx->row1 - ctx->row0; p.row0 = ctx->row0;
  p.x0 = ctx->x0; p.y0 = ctx->y0; p.rx = ctx->rx; p.ry = ctx->ry;
  p.xscale = std::sqrt(1.4426950408889634 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale));
  p.sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  p.inv2l2 = 1.0 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale);
  p.prior_mean = ctx->cfg.prior_mean;
  const int kr = k_new <= 1 ? 1 : (k_new <= 2 ? 2 : (k_new <= 4 ? 4 : 8));
  if (ctx->cfg.precision == SBO_FP32) {
    SBO_TRY(tc_pack_observations(ctx));
    switch (kr) {
      case 1: launch_thin<float, 1, 4>(ctx, p); break;
      case 2: launch_thin<float, 2, 4>(ctx, p); break;
      case 4: launch_thin<float, 4, 2>(ctx, p); break;
      default: launch_thin<float, 8, 2>(ctx, p); break;
    }
  } else {
    switch (kr) {
      case 1: launch_thin<double, 1, 2>(ctx, p); break;
      case 2: launch_thin<double, 2, 2>(ctx, p); break;
      case 4: launch_thin<double, 4, 2>(ctx, p); break;
      default: launch_thin<double, 8, 1>(ctx, p); break;
    }
  }
  SBO_LAUNCH_CHECK(ctx, "posterior_thin_kernel");
  return SBO_OK;
}

}  // namespace sbo
