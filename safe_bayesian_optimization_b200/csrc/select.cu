// select.cu -- GetNextSubgoal's selection rule on the GPU.
//
// Replaces reference src/safe_bayesian_optimization_node.cpp:508-549: for a
// frontier list (cell indices from cv::findContours, :461-492) compute
//   width_i = Q1 - Q0 (:516),  dist_i = sqrt(dx^2 + dy^2) to the goal (:519-523),
// sort (dist, position) pairs ascending (:526-530), keep the first
// T = max(1, F/4) (:533) and return the entry with the strictly largest width
// scanning in sorted order from -1.0 (:536-547), i.e. the maximum width with
// ties resolved to the smallest (dist, position).
//
// The sort is not materialised: an MSB-first radix select over the 96-bit key
// (distance bits, list position) finds the T-th smallest key exactly, then a
// masked argmax runs over the keys at or below it.  F is O(perimeter) (1e2-1e5
// cells), so one 1024-thread CTA is enough; the work is a handful of passes
// over two arrays that stay in L2.
#include "common.cuh"

namespace sbo {
namespace {

constexpr int kSelThreads = 1024;

struct SelParams {
  double x0, y0, rx, ry;
  int W;
  long long cell0, m;
  double beta, noise_add, gx, gy;
};

struct Cand {
  double w;
  unsigned long long key;
  long long pos;
};

__device__ __forceinline__ bool cand_better(const Cand& a, const Cand& b) {
  // a beats b: larger width, then smaller distance key, then smaller position
  if (a.pos < 0) return false;
  if (b.pos < 0) return true;
  if (a.w != b.w) return a.w > b.w;
  if (a.key != b.key) return a.key < b.key;
  return a.pos < b.pos;
}

template <typename T, bool kStd>
__global__ void __launch_bounds__(kSelThreads)
select_kernel(const int32_t* __restrict__ frontier, long long F, SelParams p, const T* __restrict__ mu,
              const T* __restrict__ s, unsigned long long* __restrict__ keys, double* __restrict__ width,
              int32_t* __restrict__ out, int mode) {
  // mode 0: the whole rule.  Device groups (group.cu) split it: mode 1 = phase 1 only, on the cells of THIS shard (cells
  // of other shards are marked key = ~0, width = -2 and are no error); mode 2 = phases 2 and 3 on keys / widths that the
  // host has merged from all shards.
  __shared__ unsigned int hist[256];
  __shared__ unsigned long long sh_prefix;
  __shared__ long long sh_rank;
  __shared__ int sh_bad;
  __shared__ Cand sh_cand[kSelThreads / 32];
  const int tid = threadIdx.x;
  if (tid == 0) sh_bad = 0;
  __syncthreads();

  // ---- phase 1: distance keys and widths --------------------------------------
  for (long long i = tid; mode != 2 && i < F; i += kSelThreads) {
    const long long cell = frontier[i];
    const long long local = cell - p.cell0;
    if (local < 0 || local >= p.m) {
      if (mode == 0) sh_bad = 1;
      keys[i] = ~0ull;
      width[i] = -2.0;
      continue;
    }
    const int iy = (int)(cell / p.W), ix = (int)(cell - (long long)iy * p.W);
    const double dx = __dsub_rn(cell_coord(p.x0, p.rx, ix), p.gx);
    const double dy = __dsub_rn(cell_coord(p.y0, p.ry, iy), p.gy);
    const double d = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    double sd;
    if (kStd) {
      sd = (double)s[local];
    } else {
      double v = (double)s[local];
      if (v < 0.0) v = 0.0;
      sd = __dsqrt_rn(__dadd_rn(v, p.noise_add));
    }
    const double m_ = (double)mu[local];
    const double conf = __dmul_rn(p.beta, sd);
    const double q0 = __dsub_rn(m_, conf), q1 = __dadd_rn(m_, conf);
    keys[i] = (unsigned long long)__double_as_longlong(d);  // d >= 0: bit pattern is order preserving
    width[i] = __dsub_rn(q1, q0);
  }
  __syncthreads();
  if (sh_bad) {
    if (tid == 0) *out = -2;  // frontier cell outside this context's cells
    return;
  }
  if (mode == 1) return;

  // ---- phase 2: radix select of the T-th smallest (key, position) -------------
  long long top = F / 4;
  if (top < 1) top = 1;
  if (tid == 0) {
    sh_prefix = 0ull;
    sh_rank = top - 1;
  }
  __syncthreads();
  // 8 digits over the distance bits
  for (int dgt = 7; dgt >= 0; --dgt) {
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    const unsigned long long prefix = sh_prefix;
    const unsigned long long decided = (dgt == 7) ? 0ull : (~0ull << (8 * (dgt + 1)));
    for (long long i = tid; i < F; i += kSelThreads) {
      const unsigned long long k = keys[i];
      if ((k & decided) == prefix) atomicAdd(&hist[(k >> (8 * dgt)) & 255u], 1u);
    }
    __syncthreads();
    if (tid == 0) {
      long long r = sh_rank, cum = 0;
      int b = 0;
      for (; b < 256; ++b) {
        if (r < cum + (long long)hist[b]) break;
        cum += hist[b];
      }
      sh_rank = r - cum;
      sh_prefix = prefix | ((unsigned long long)b << (8 * dgt));
    }
    __syncthreads();
  }
  const unsigned long long key_T = sh_prefix;
  // 4 digits over the position among entries whose distance equals key_T
  __shared__ unsigned int sh_pos_prefix;
  if (tid == 0) sh_pos_prefix = 0u;
  __syncthreads();
  for (int dgt = 3; dgt >= 0; --dgt) {
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    const unsigned int prefix = sh_pos_prefix;
    const unsigned int decided = (dgt == 3) ? 0u : (~0u << (8 * (dgt + 1)));
    for (long long i = tid; i < F; i += kSelThreads) {
      if (keys[i] == key_T && (((unsigned int)i) & decided) == prefix)
        atomicAdd(&hist[(((unsigned int)i) >> (8 * dgt)) & 255u], 1u);
    }
    __syncthreads();
    if (tid == 0) {
      long long r = sh_rank, cum = 0;
      int b = 0;
      for (; b < 256; ++b) {
        if (r < cum + (long long)hist[b]) break;
        cum += hist[b];
      }
      sh_rank = r - cum;
      sh_pos_prefix = prefix | ((unsigned int)b << (8 * dgt));
    }
    __syncthreads();
  }
  const long long pos_T = (long long)sh_pos_prefix;

  // ---- phase 3: masked argmax of the width --------------------------------------
  Cand best{-1.0, 0ull, -1};
  for (long long i = tid; i < F; i += kSelThreads) {
    const unsigned long long k = keys[i];
    if (k < key_T || (k == key_T && i <= pos_T)) {
      const double w = width[i];
      if (w > -1.0) {  // the scan starts from best = -1.0 with a strict > (:536, :543)
        Cand c{w, k, i};
        if (cand_better(c, best)) best = c;
      }
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    Cand o;
    o.w = __shfl_down_sync(0xffffffffu, best.w, d);
    o.key = __shfl_down_sync(0xffffffffu, best.key, d);
    o.pos = __shfl_down_sync(0xffffffffu, best.pos, d);
    if (cand_better(o, best)) best = o;
  }
  if ((tid & 31) == 0) sh_cand[tid >> 5] = best;
  __syncthreads();
  if (tid == 0) {
    Cand r = sh_cand[0];
    for (int w = 1; w < kSelThreads / 32; ++w)
      if (cand_better(sh_cand[w], r)) r = sh_cand[w];
    *out = r.pos >= 0 ? frontier[r.pos] : -1;
  }
}

}  // namespace

int32_t launch_select(sbo_ctx* ctx, const int32_t* d_frontier, int64_t F, double gx, double gy, int mode) {
  SelParams p;
  p.x0 = ctx->x0; p.y0 = ctx->y0; p.rx = ctx->rx; p.ry = ctx->ry; p.W = ctx->W;
  p.beta = ctx->last_beta; p.gx = gx; p.gy = gy;
  if (ctx->source == PosteriorSource::kHostArrays) {
    p.cell0 = 0; p.m = ctx->source_m; p.noise_add = 0.0;
    select_kernel<double, true><<<1, kSelThreads, 0, ctx->stream>>>(d_frontier, F, p, ctx->d_mu64, ctx->d_std64,
                                                                     ctx->d_sel_keys, ctx->d_sel_width, ctx->d_sel_out, mode);
  } else {
    p.cell0 = ctx->cell0; p.m = ctx->m_local;
    p.noise_add = ctx->cfg.include_noise_in_sigma ? ctx->cfg.noise_var : 0.0;
    if (ctx->cfg.precision == SBO_FP64) {
      select_kernel<double, false><<<1, kSelThreads, 0, ctx->stream>>>(
          d_frontier, F, p, static_cast<const double*>(ctx->d_mu), static_cast<const double*>(ctx->d_s2),
          ctx->d_sel_keys, ctx->d_sel_width, ctx->d_sel_out, mode);
    } else {
      select_kernel<float, false><<<1, kSelThreads, 0, ctx->stream>>>(
          d_frontier, F, p, static_cast<const float*>(ctx->d_mu), static_cast<const float*>(ctx->d_s2),
          ctx->d_sel_keys, ctx->d_sel_width, ctx->d_sel_out, mode);
    }
  }
  SBO_LAUNCH_CHECK(ctx, "select_kernel");
  return SBO_OK;
}

}  // namespace sbo
