// sets.cu -- fused confidence interval + safe mask + argmax pass.
//
// Replaces OptimizerNode::ComputeConfidenceIntervals / UpdateSafeSet
// (reference src/safe_bayesian_optimization_node.cpp:409-416) and the
// rasterise loop that turns S_ into a 0/255 image (:444-456):
//     conf = beta * sigma ; Q0 = mu - conf ; Q1 = mu + conf ; S = Q0 > f_min
// All interval arithmetic is fp64 with every product/sum rounded on its own
// (__dmul_rn/__dsub_rn/__dadd_rn are never contracted into FMAs), so that on
// fp64 inputs the result is bit-identical to the reference's Eigen code built
// for baseline x86-64.
//
// HBM-bound: reads mu + sigma^2 (2*sizeof(T) per cell), writes 1 mask byte.
// Each thread handles 4 consecutive cells per access (16 B loads in fp32,
// 2x16 B in fp64, one 4-byte mask store => a warp stores 128 contiguous
// bytes), several independent accesses in flight per thread; argmax is a
// warp-shuffle -> block -> last-block-done grid reduction on (value, index)
// records ordered (value desc, index asc).
#include <algorithm>
#include <cmath>
#include <string>

#include "common.cuh"

namespace sbo {

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;  // independent quads in flight per thread

struct Best {
  double v;
  long long i;
};

__device__ __forceinline__ void merge(Best& a, double v, long long i) {
  if (better(v, i, a.v, a.i)) {
    a.v = v;
    a.i = i;
  }
}

__device__ __forceinline__ Best shfl_down(const Best& b, int d) {
  Best o;
  o.v = __shfl_down_sync(0xffffffffu, b.v, d);
  o.i = __shfl_down_sync(0xffffffffu, b.i, d);
  return o;
}

struct Acc {
  Best ucb{0.0, -1};
  Best wid{0.0, -1};
  long long safe = 0;
  long long clamped = 0;
  long long low = 0;   // sigma^2 < low_thr (0.02 sigma_f^2): cells where fp32 mode cannot hold 1e-5 of sigma^2 itself
};

// One cell.  kStd: `s` is already a standard deviation (reference's std_);
// otherwise it is a variance: clamp at 0, add the noise term, take the root.
template <bool kStd>
__device__ __forceinline__ uint32_t eval_cell(double mu, double s, double beta, double f_min, double noise_add,
                                              double low_thr, long long idx, Acc& acc, double& q0, double& q1) {
  double sd;
  if (kStd) {
    sd = s;
  } else {
    double v = s;
    if (v < low_thr) acc.low++;
    if (v < 0.0) {
      v = 0.0;
      acc.clamped++;
    }
    v = __dadd_rn(v, noise_add);
    sd = __dsqrt_rn(v);
  }
  const double conf = __dmul_rn(beta, sd);  // :412
  q0 = __dsub_rn(mu, conf);                 // :414
  q1 = __dadd_rn(mu, conf);                 // :415
  const bool safe = q0 > f_min;             // :409 (strict; NaN -> false)
  if (safe) {
    acc.safe++;
    // indices grow within a thread, so strict > keeps the lowest index on ties
    if (acc.ucb.i < 0 || q1 > acc.ucb.v) {
      acc.ucb.v = q1;
      acc.ucb.i = idx;
    }
    const double w = __dsub_rn(q1, q0);  // :516
    if (acc.wid.i < 0 || w > acc.wid.v) {
      acc.wid.v = w;
      acc.wid.i = idx;
    }
  }
  return safe ? 0xFFu : 0u;
}

template <typename T>
struct Quad;  // 4 consecutive values
template <>
struct Quad<float> {
  float4 a;
  __device__ __forceinline__ void load(const float* p) { a = __ldcs(reinterpret_cast<const float4*>(p)); }
  __device__ __forceinline__ double get(int i) const { return i == 0 ? a.x : i == 1 ? a.y : i == 2 ? a.z : a.w; }
};
template <>
struct Quad<double> {
  double2 a, b;
  __device__ __forceinline__ void load(const double* p) {
    a = __ldcs(reinterpret_cast<const double2*>(p));
    b = __ldcs(reinterpret_cast<const double2*>(p) + 1);
  }
  __device__ __forceinline__ double get(int i) const { return i == 0 ? a.x : i == 1 ? a.y : i == 2 ? b.x : b.y; }
};

__device__ __forceinline__ void store_q(double* q, long long base, double v0, double v1, double v2, double v3) {
  __stcs(reinterpret_cast<double2*>(q + base), make_double2(v0, v1));
  __stcs(reinterpret_cast<double2*>(q + base) + 1, make_double2(v2, v3));
}

// kDelta (SBO_OPT_MASK_DELTA): `mask` still holds the previous step's bytes; a thread that is about to change any of them
// raises the flag of the 64 KiB segment they lie in (dirty[cell >> kMaskSegShift]; plain stores of 1, every writer agrees).
// The host copies back only the flagged segments.  One extra byte read per cell.
template <typename T, bool kStd, bool kVec, bool kDelta = false>
__global__ void __launch_bounds__(kThreads)
sets_kernel(const T* __restrict__ mu, const T* __restrict__ s, long long m, long long cell0, double beta, double f_min,
            double noise_add, double low_thr, uint8_t* __restrict__ mask, double* __restrict__ q0out, double* __restrict__ q1out,
            SetsResult* __restrict__ partials, SetsResult* __restrict__ result, unsigned int* __restrict__ ticket,
            unsigned int* __restrict__ dirty = nullptr) {
  Acc acc;
  const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
  const long long nthreads = (long long)gridDim.x * kThreads;

  if (kVec) {
    const long long nquads = m >> 2;
    for (long long q = tid; q < nquads; q += nthreads * kUnroll) {
      Quad<T> qm[kUnroll], qs[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long qq = q + (long long)u * nthreads;
        if (qq < nquads) {
          qm[u].load(mu + (qq << 2));
          qs[u].load(s + (qq << 2));
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long qq = q + (long long)u * nthreads;
        if (qq < nquads) {
          const long long base = qq << 2;
          double a0[4], a1[4];
          uint32_t bytes = 0;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            bytes |= eval_cell<kStd>(qm[u].get(e), qs[u].get(e), beta, f_min, noise_add, low_thr, cell0 + base + e, acc, a0[e],
                                     a1[e])
                     << (8 * e);
          }
          if (kDelta) {
            if (__ldcs(reinterpret_cast<const unsigned int*>(mask + base)) != bytes) dirty[base >> kMaskSegShift] = 1u;
          }
          __stcs(reinterpret_cast<unsigned int*>(mask + base), bytes);
          if (q0out) store_q(q0out, base, a0[0], a0[1], a0[2], a0[3]);
          if (q1out) store_q(q1out, base, a1[0], a1[1], a1[2], a1[3]);
        }
      }
    }
    // tail (m % 4 cells)
    const long long tail0 = (m >> 2) << 2;
    if (tid < m - tail0) {
      const long long c = tail0 + tid;
      double a0, a1;
      const uint32_t b = eval_cell<kStd>((double)mu[c], (double)s[c], beta, f_min, noise_add, low_thr, cell0 + c, acc, a0, a1);
      if (kDelta && mask[c] != (uint8_t)b) dirty[c >> kMaskSegShift] = 1u;
      mask[c] = (uint8_t)b;
      if (q0out) q0out[c] = a0;
      if (q1out) q1out[c] = a1;
    }
  } else {
    for (long long c = tid; c < m; c += nthreads) {
      double a0, a1;
      const uint32_t b = eval_cell<kStd>((double)mu[c], (double)s[c], beta, f_min, noise_add, low_thr, cell0 + c, acc, a0, a1);
      if (kDelta && mask[c] != (uint8_t)b) dirty[c >> kMaskSegShift] = 1u;
      mask[c] = (uint8_t)b;
      if (q0out) q0out[c] = a0;
      if (q1out) q1out[c] = a1;
    }
  }

  // ---- warp -> block -> grid reduction -------------------------------------
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    Best u = shfl_down(acc.ucb, d), w = shfl_down(acc.wid, d);
    merge(acc.ucb, u.v, u.i);
    merge(acc.wid, w.v, w.i);
    acc.safe += __shfl_down_sync(0xffffffffu, acc.safe, d);
    acc.clamped += __shfl_down_sync(0xffffffffu, acc.clamped, d);
    acc.low += __shfl_down_sync(0xffffffffu, acc.low, d);
  }
  __shared__ SetsResult warp_res[kThreads / 32];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) warp_res[warp] = SetsResult{acc.ucb.v, acc.ucb.i, acc.wid.v, acc.wid.i, acc.safe, acc.clamped, acc.low};
  __syncthreads();
  if (threadIdx.x == 0) {
    SetsResult r = warp_res[0];
    for (int w = 1; w < kThreads / 32; ++w) {
      const SetsResult o = warp_res[w];
      if (better(o.ucb_val, o.ucb_idx, r.ucb_val, r.ucb_idx)) { r.ucb_val = o.ucb_val; r.ucb_idx = o.ucb_idx; }
      if (better(o.width_val, o.width_idx, r.width_val, r.width_idx)) { r.width_val = o.width_val; r.width_idx = o.width_idx; }
      r.safe_count += o.safe_count;
      r.clamped += o.clamped;
      r.low_sigma2 += o.low_sigma2;
    }
    partials[blockIdx.x] = r;
    __threadfence();
    const unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // last block: fold all per-block records
  Acc tot;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += kThreads) {
    const SetsResult o = partials[b];
    merge(tot.ucb, o.ucb_val, o.ucb_idx);
    merge(tot.wid, o.width_val, o.width_idx);
    tot.safe += o.safe_count;
    tot.clamped += o.clamped;
    tot.low += o.low_sigma2;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    Best u = shfl_down(tot.ucb, d), w = shfl_down(tot.wid, d);
    merge(tot.ucb, u.v, u.i);
    merge(tot.wid, w.v, w.i);
    tot.safe += __shfl_down_sync(0xffffffffu, tot.safe, d);
    tot.clamped += __shfl_down_sync(0xffffffffu, tot.clamped, d);
    tot.low += __shfl_down_sync(0xffffffffu, tot.low, d);
  }
  __syncthreads();  // warp_res reuse
  if (lane == 0) warp_res[warp] = SetsResult{tot.ucb.v, tot.ucb.i, tot.wid.v, tot.wid.i, tot.safe, tot.clamped, tot.low};
  __syncthreads();
  if (threadIdx.x == 0) {
    SetsResult r = warp_res[0];
    for (int w = 1; w < kThreads / 32; ++w) {
      const SetsResult o = warp_res[w];
      if (better(o.ucb_val, o.ucb_idx, r.ucb_val, r.ucb_idx)) { r.ucb_val = o.ucb_val; r.ucb_idx = o.ucb_idx; }
      if (better(o.width_val, o.width_idx, r.width_val, r.width_idx)) { r.width_val = o.width_val; r.width_idx = o.width_idx; }
      r.safe_count += o.safe_count;
      r.clamped += o.clamped;
      r.low_sigma2 += o.low_sigma2;
    }
    *result = r;
    *ticket = 0;  // ready for the next launch
  }
}

template <typename T, bool kStd>
int32_t launch(sbo_ctx* ctx, const T* mu, const T* s, int64_t m, int64_t cell0, double beta, double f_min,
               double noise_add, uint8_t* mask, double* q0, double* q1, unsigned int* dirty = nullptr) {
  const double low_thr = kStd ? 0.0 : 0.02 * ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  auto aligned = [](const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  const bool vec = aligned(mu, 16) && aligned(s, 16) && aligned(mask, 4) && (!q0 || aligned(q0, 16)) &&
                   (!q1 || aligned(q1, 16));
  const int64_t work = vec ? (m + 3) / 4 : m;
  int64_t blocks = (work + (int64_t)kThreads * kUnroll - 1) / ((int64_t)kThreads * kUnroll);
  const int64_t cap = sets_grid_size(ctx);
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if constexpr (!kStd) {
    if (dirty) {   // the step's pass with changed-segment flags (SBO_OPT_MASK_DELTA)
      if (vec)
        sets_kernel<T, kStd, true, true><<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(
            mu, s, m, cell0, beta, f_min, noise_add, low_thr, mask, q0, q1, ctx->d_partials, ctx->d_result, ctx->d_ticket, dirty);
      else
        sets_kernel<T, kStd, false, true><<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(
            mu, s, m, cell0, beta, f_min, noise_add, low_thr, mask, q0, q1, ctx->d_partials, ctx->d_result, ctx->d_ticket, dirty);
      SBO_LAUNCH_CHECK(ctx, "sets_kernel(delta)");
      return SBO_OK;
    }
  }
  if (vec) {
    sets_kernel<T, kStd, true><<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(
        mu, s, m, cell0, beta, f_min, noise_add, low_thr, mask, q0, q1, ctx->d_partials, ctx->d_result, ctx->d_ticket);
  } else {
    sets_kernel<T, kStd, false><<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(
        mu, s, m, cell0, beta, f_min, noise_add, low_thr, mask, q0, q1, ctx->d_partials, ctx->d_result, ctx->d_ticket);
  }
  SBO_LAUNCH_CHECK(ctx, "sets_kernel");
  return SBO_OK;
}

}  // namespace

int sets_grid_size(const sbo_ctx* ctx) { return ctx->sm_count * 8; }  // 8 x 256 threads resident per SM

int32_t launch_sets_step(sbo_ctx* ctx, double beta, double f_min) {
  const double noise_add = ctx->cfg.include_noise_in_sigma ? ctx->cfg.noise_var : 0.0;
  unsigned int* dirty = ctx->opt.mask_delta ? ctx->d_mask_dirty : nullptr;
  if (ctx->cfg.precision == SBO_FP64) {
    return launch<double, false>(ctx, static_cast<const double*>(ctx->d_mu), static_cast<const double*>(ctx->d_s2),
                                 ctx->m_local, ctx->cell0, beta, f_min, noise_add, ctx->d_mask, nullptr, nullptr, dirty);
  }
  return launch<float, false>(ctx, static_cast<const float*>(ctx->d_mu), static_cast<const float*>(ctx->d_s2),
                              ctx->m_local, ctx->cell0, beta, f_min, noise_add, ctx->d_mask, nullptr, nullptr, dirty);
}

int32_t launch_sets_arrays(sbo_ctx* ctx, const double* d_mu, const double* d_std, int64_t m, double beta, double f_min,
                           uint8_t* d_mask, double* d_q0, double* d_q1) {
  return launch<double, true>(ctx, d_mu, d_std, m, 0, beta, f_min, 0.0, d_mask, d_q0, d_q1);
}

}  // namespace sbo

// ---------------------------------------------------------------------------------------------------------------------
// sbo_compute_sets_noalloc -- the CI + mask + argmax pass for callers that may not allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernel as sbo_compute_sets_device (sets_kernel<double, std inputs>); its grid
// reduction scratch (one record per block + the last-block ticket) lives in a file-scope __device__ array instead of the
// context's cudaMalloc'ed buffers, so the call needs no context and allocates nothing.  Every GPU that loads libsbo.so
// holds the array (kNoAllocBlocks x 56 B + 256 B = 112.25 KiB) until the process exits.  All calls on a device share it:
// calls that overlap on the GPU (two streams, two host threads) race on it -- serialise them.
// ---------------------------------------------------------------------------------------------------------------------
namespace sbo {
namespace {

constexpr int kNoAllocBlocks = 2048;                       // per-block records the scratch holds
constexpr long long kNoAllocMaxCells = 1LL << 27;           // stated capacity: 134 217 728 cells (2 x the 8192^2 grid of C4)
struct NoAllocScratch {
  unsigned int ticket;
  unsigned int pad[63];                                     // the records start on their own 256-byte line
  SetsResult partials[kNoAllocBlocks];
};
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the device address is looked up once
// per device, with that device current, and cached here.
constexpr int kNoAllocMaxDevices = 64;
NoAllocScratch* g_noalloc_ptr[kNoAllocMaxDevices] = {};

int32_t noalloc_scratch(NoAllocScratch** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "sbo_compute_sets_noalloc: cudaGetDevice");
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_compute_sets_noalloc: device ordinal beyond 63");
  NoAllocScratch* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_noalloc_scratch);      // set-up, once per device (idempotent: a race stores the same value)
    if (e != cudaSuccess || !sym) return cuda_fail(nullptr, e, "sbo_compute_sets_noalloc: cudaGetSymbolAddress(scratch)");
    p = static_cast<NoAllocScratch*>(sym);
    __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return SBO_OK;
}

}  // namespace
}  // namespace sbo

extern "C" int32_t sbo_compute_sets_noalloc(const double* mu_dev, const double* std_dev, int64_t m, double beta, double f_min,
                                            uint8_t* mask_dev, sbo_sets_record* result_dev, void* cuda_stream) {
  using namespace sbo;
  static_assert(sizeof(sbo_sets_record) == sizeof(SetsResult), "sbo_sets_record mirrors the device record");
  if (m < 0 || !result_dev || (m > 0 && (!mu_dev || !std_dev || !mask_dev)))
    return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_compute_sets_noalloc: null pointer or m < 0");
  if (!(std::isfinite(beta) && std::isfinite(f_min))) return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_compute_sets_noalloc: beta / f_min not finite");
  if (m > kNoAllocMaxCells)
    return fail(nullptr, SBO_ERR_CAPACITY, "sbo_compute_sets_noalloc: m = " + std::to_string((long long)m) + " exceeds the fixed capacity of " +
                                               std::to_string(kNoAllocMaxCells) + " cells (use sbo_compute_sets_device)");
  NoAllocScratch* scratch = nullptr;
  SBO_TRY(noalloc_scratch(&scratch));
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  // the array keeps the last call's contents: the ticket must start at zero (a call that was aborted mid-way leaves it
  // anywhere); the per-block records are all overwritten before they are read
  cudaError_t e = cudaMemsetAsync(&scratch->ticket, 0, sizeof(unsigned int), stream);
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "sbo_compute_sets_noalloc: cudaMemsetAsync(ticket)");
  auto aligned = [](const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  const bool vec = aligned(mu_dev, 16) && aligned(std_dev, 16) && aligned(mask_dev, 4);
  const int64_t work = vec ? (m + 3) / 4 : m;
  int64_t blocks = (work + (int64_t)kThreads * kUnroll - 1) / ((int64_t)kThreads * kUnroll);
  blocks = std::max<int64_t>(1, std::min<int64_t>(blocks, kNoAllocBlocks));
  SetsResult* result = reinterpret_cast<SetsResult*>(result_dev);
  if (vec)
    sets_kernel<double, true, true><<<(unsigned)blocks, kThreads, 0, stream>>>(mu_dev, std_dev, m, 0, beta, f_min, 0.0, 0.0, mask_dev, nullptr,
                                                                                nullptr, scratch->partials, result, &scratch->ticket);
  else
    sets_kernel<double, true, false><<<(unsigned)blocks, kThreads, 0, stream>>>(mu_dev, std_dev, m, 0, beta, f_min, 0.0, 0.0, mask_dev, nullptr,
                                                                                 nullptr, scratch->partials, result, &scratch->ticket);
  e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "sbo_compute_sets_noalloc: sets_kernel launch");
  return SBO_OK;   // asynchronous: the record and the mask are complete when `cuda_stream` reaches this point
}
