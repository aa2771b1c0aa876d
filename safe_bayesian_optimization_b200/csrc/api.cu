// api.cu -- the C-ABI of include/sbo.h: context life cycle, host<->device plumbing and the
// sequencing of the kernels of one BO step.  No arithmetic of the path lives here.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "common.cuh"

namespace sbo {

static thread_local std::string g_create_error;

int32_t fail(sbo_ctx* ctx, int32_t status, const std::string& msg) {
  if (ctx) ctx->err = msg;
  else g_create_error = msg;
  return status;
}

int32_t cuda_fail(sbo_ctx* ctx, cudaError_t e, const char* what) {
  return fail(ctx, SBO_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

namespace {

size_t elem_size(const sbo_ctx* ctx) { return ctx->cfg.precision == SBO_FP64 ? sizeof(double) : sizeof(float); }

bool all_finite(const double* p, size_t count) {
  for (size_t i = 0; i < count; ++i)
    if (!std::isfinite(p[i])) return false;
  return true;
}

void free_all(sbo_ctx* c) {
  tc_destroy(c);
  void* dev[] = {c->d_xy, c->d_r, c->d_A, c->d_Linv, c->d_Dinv, c->d_alpha, c->d_w, c->d_small, c->d_info,
                 c->d_planes, c->d_alpha_f32, c->d_obs_f32, c->d_chunk_bbox, c->d_exec_cols, c->d_fr_img, c->d_fr_lut, c->d_cp_idx, c->d_cp_counts, c->d_mu, c->d_s2, c->d_mask, c->d_mu64, c->d_std64, c->d_q0, c->d_q1,
                 c->d_partials, c->d_result, c->d_ticket, c->d_frontier, c->d_sel_keys, c->d_sel_width, c->d_sel_out,
                 c->d_gather, c->d_chunk_bbox64, c->d_drift_mu, c->d_drift_s2, c->d_drift_out, c->d_fit_stage, c->d_fit_send, c->d_fit_recv,
                 c->d_mask_dirty};
  for (void* p : dev)
    if (p) cudaFree(p);
  if (c->h_result) cudaFreeHost(c->h_result);
  if (c->h_sel_out) cudaFreeHost(c->h_sel_out);
  if (c->h_mask_dirty) cudaFreeHost(c->h_mask_dirty);
  for (auto& e : c->ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : c->ev_factor) cudaEventDestroy(e);
  c->ev_factor.clear();
  for (auto& e : c->ev_fit)
    if (e) cudaEventDestroy(e);
  for (cudaStream_t* q : {&c->fit_hi, &c->fit_lo, &c->fit_cm})
    if (*q) cudaStreamDestroy(*q);
  if (c->aux_stream) cudaStreamDestroy(c->aux_stream);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
}

// Which posterior kernel runs for this context.
int32_t pick_kernel(sbo_ctx* ctx, int* which) {
  if (ctx->cfg.kernel == SBO_KERNEL_SIMT) {
    *which = SBO_KERNEL_SIMT;
    return SBO_OK;
  }
  if (ctx->cfg.precision == SBO_FP64) {
    *which = SBO_KERNEL_DMMA;
    return SBO_OK;
  }
  if (!ctx->tc_ok)
    return fail(ctx, SBO_ERR_STATE, "tcgen05 posterior kernel unavailable on this device/driver (no fallback in fp32 mode; "
                                    "set kernel = SBO_KERNEL_SIMT explicitly to use the FMA kernel)");
  *which = SBO_KERNEL_TCGEN05;
  return SBO_OK;
}

void fill_out(const sbo_ctx* ctx, const SetsResult& r, int kernel_used, const float* ms, sbo_step_out* out) {
  std::memset(out, 0, sizeof(*out));
  out->n_cells = ctx->source == PosteriorSource::kHostArrays ? ctx->source_m : ctx->m_local;
  out->safe_count = r.safe_count;
  out->argmax_ucb.idx = r.ucb_idx;
  out->argmax_ucb.value = r.ucb_idx >= 0 ? r.ucb_val : 0.0;
  out->argmax_width.idx = r.width_idx;
  out->argmax_width.value = r.width_idx >= 0 ? r.width_val : 0.0;
  out->clamped_sigma2 = r.clamped;
  out->low_sigma2 = r.low_sigma2;
  out->refreshed = ctx->inc_refreshed;
  out->drift = ctx->inc_drift;
  out->n_obs = ctx->n;
  out->kernel_used = kernel_used;
  out->ms_posterior = ms[0];
  out->ms_sets = ms[1];
  out->ms_comm = ms[2];
  out->ms_total = ms[3];
}

// Drift estimate of the incremental chain: `drift_rows` grid rows of this shard are re-evaluated from scratch with the
// full-refresh kernel of the context's precision (O(W n^2) each) into scratch buffers and compared with the stored,
// incrementally updated values.  The result is read with the step's record (finish_step); above tol / 4 the next
// incremental step becomes a full refresh.
template <typename T>
__global__ void drift_compare_kernel(const T* __restrict__ mu, const T* __restrict__ s2, const T* __restrict__ mu_x,
                                     const T* __restrict__ s2_x, int W, long long off, double inv_sf2, double* __restrict__ out) {
  double dm = 0.0, ds = 0.0;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < W; c += gridDim.x * blockDim.x) {
    const double m0 = (double)mu_x[c], m1 = (double)mu[off + c];
    dm = fmax(dm, fabs(m1 - m0) / fmax(1.0, fabs(m0)));
    ds = fmax(ds, fabs((double)s2[off + c] - (double)s2_x[c]) * inv_sf2);
  }
  for (int d = 16; d > 0; d >>= 1) {
    dm = fmax(dm, __shfl_xor_sync(0xffffffffu, dm, d));
    ds = fmax(ds, __shfl_xor_sync(0xffffffffu, ds, d));
  }
  if ((threadIdx.x & 31) == 0) {   // non-negative doubles order like their bit patterns
    atomicMax(reinterpret_cast<unsigned long long*>(out), (unsigned long long)__double_as_longlong(dm));
    atomicMax(reinterpret_cast<unsigned long long*>(out + 1), (unsigned long long)__double_as_longlong(ds));
  }
}

int32_t drift_check(sbo_ctx* ctx) {
  const int rows = ctx->row1 - ctx->row0, W = ctx->W;
  if (rows <= 0 || ctx->n == 0) return SBO_OK;
  int which = 0;
  SBO_TRY(pick_kernel(ctx, &which));
  // A sample = a block of consecutive grid rows: one row for the tcgen05 / SIMT kernels (their tiles are pieces of one
  // row), 8 for the DMMA kernel, whose tile is a 16 x 8 patch of cells -- a single row would cost it the same as eight.
  const int rb = std::min(rows, which == SBO_KERNEL_DMMA ? 8 : 1);
  const int dr = std::min(std::max(ctx->opt.drift_rows, 1), rows / rb);
  const size_t es = elem_size(ctx);
  const int64_t need = (int64_t)W * rb;
  if (ctx->drift_cap < need) {
    for (void** p : {&ctx->d_drift_mu, &ctx->d_drift_s2}) {
      if (*p) cudaFree(*p);
      *p = nullptr;
      SBO_CUDA(ctx, cudaMalloc(p, es * (size_t)need));
    }
    if (!ctx->d_drift_out) SBO_CUDA(ctx, cudaMalloc(&ctx->d_drift_out, 2 * sizeof(double)));
    ctx->drift_cap = need;
  }
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_drift_out, 0, 2 * sizeof(double), ctx->stream));
  // the launchers read the shard geometry and the output buffers from the context: point them at one row + the scratch
  const int row0 = ctx->row0, row1 = ctx->row1;
  const int64_t m_local = ctx->m_local, cell0 = ctx->cell0;
  void* mu = ctx->d_mu;
  void* s2 = ctx->d_s2;
  int32_t st = SBO_OK;
  for (int i = 0; i < dr && st == SBO_OK; ++i) {
    const int lr = std::min((int)(((2LL * i + 1) * rows) / (2LL * dr)), rows - rb);   // spread over the shard
    ctx->row0 = row0 + lr; ctx->row1 = row0 + lr + rb;
    ctx->m_local = need; ctx->cell0 = (int64_t)ctx->row0 * W;
    ctx->d_mu = ctx->d_drift_mu; ctx->d_s2 = ctx->d_drift_s2;
    if (which == SBO_KERNEL_TCGEN05) st = launch_posterior_tc(ctx, 0, ctx->n, false);
    else if (which == SBO_KERNEL_DMMA) st = launch_posterior_f64(ctx, 0, ctx->n, false);
    else st = launch_posterior_simt(ctx, 0, ctx->n, false);
    ctx->row0 = row0; ctx->row1 = row1; ctx->m_local = m_local; ctx->cell0 = cell0; ctx->d_mu = mu; ctx->d_s2 = s2;
    if (st != SBO_OK) break;
    const double inv_sf2 = 1.0 / (ctx->cfg.sigma_f * ctx->cfg.sigma_f);
    const long long off = (long long)lr * W;
    if (ctx->cfg.precision == SBO_FP64)
      drift_compare_kernel<double><<<(int)std::min<int64_t>((need + 255) / 256, 64), 256, 0, ctx->stream>>>(
          static_cast<const double*>(mu), static_cast<const double*>(s2), static_cast<const double*>(ctx->d_drift_mu),
          static_cast<const double*>(ctx->d_drift_s2), (int)need, off, inv_sf2, ctx->d_drift_out);
    else
      drift_compare_kernel<float><<<(int)std::min<int64_t>((need + 255) / 256, 64), 256, 0, ctx->stream>>>(
          static_cast<const float*>(mu), static_cast<const float*>(s2), static_cast<const float*>(ctx->d_drift_mu),
          static_cast<const float*>(ctx->d_drift_s2), (int)need, off, inv_sf2, ctx->d_drift_out);
    SBO_LAUNCH_CHECK(ctx, "drift_compare_kernel");
  }
  SBO_TRY(st);
  ctx->drift_pending = true;
  return SBO_OK;
}

// posterior (full or incremental) + sets + collective, all on ctx->stream; events around each phase.
int32_t run_step(sbo_ctx* ctx, double beta, double f_min, bool incremental, int* kernel_used) {
  if (!ctx->grid_set) return fail(ctx, SBO_ERR_STATE, "sbo_step: grid not set (call sbo_set_grid)");
  if (!(std::isfinite(beta) && std::isfinite(f_min))) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_step: beta/f_min not finite");
  int which = 0;
  SBO_TRY(pick_kernel(ctx, &which));
  int j_lo = 0, j_hi = ctx->n;
  bool accumulate = false;
  ctx->inc_refreshed = 0;
  ctx->inc_drift = -1.f;
  if (incremental) {
    if (!ctx->posterior_valid || ctx->posterior_n > ctx->n)
      return fail(ctx, SBO_ERR_STATE, "sbo_step_incremental: no previous posterior on this grid for a prefix of the observations");
    // Refresh policy: the incremental update subtracts from the stored sigma^2 and adds to mu, so rounding
    // accumulates; the reference rebuilds all state every step (:633-644) and cannot drift.  A full posterior replaces the
    // update every `refresh_every` incremental steps, or when the last drift check (below) came out above tol / 4.
    int every = ctx->opt.refresh_every;
    if (every < 0) every = ctx->cfg.precision == SBO_FP64 ? 0 : 256;
    if (ctx->force_refresh || (every > 0 && ctx->inc_steps + 1 >= every)) {
      incremental = false;
      ctx->inc_refreshed = 1;
    } else {
      j_lo = ctx->posterior_n;
      accumulate = true;
    }
  }
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
  // a few new rows: the thin SIMT kernel (O(m n k)); otherwise the row blocks that hold them
  const bool thin_off = !ctx->opt.thin;
  if (incremental && which == SBO_KERNEL_TCGEN05 && tc_inc_applicable(ctx, j_hi - j_lo)) {
    // large grid, fp32: the k projections as a tensor-core GEMM through the separable kernel (posterior_tc_inc.cuh)
    SBO_TRY(launch_posterior_tc_inc(ctx, j_lo));
  } else if (incremental && !thin_off && j_hi - j_lo <= kThinMaxRows) {
    SBO_TRY(launch_posterior_thin(ctx, j_lo));
    which = SBO_KERNEL_THIN;
  } else if (which == SBO_KERNEL_TCGEN05) {
    SBO_TRY(launch_posterior_tc(ctx, j_lo, j_hi, accumulate));
  } else if (which == SBO_KERNEL_DMMA) {
    SBO_TRY(launch_posterior_f64(ctx, j_lo, j_hi, accumulate));
  } else {
    SBO_TRY(launch_posterior_simt(ctx, j_lo, j_hi, accumulate));
  }
  if (incremental) {
    ++ctx->inc_steps;
    int check_every = ctx->opt.drift_check_every;
    if (check_every < 0) check_every = ctx->cfg.precision == SBO_FP64 ? 256 : 16;   // the documented defaults (sbo.h)
    if (check_every > 0 && ctx->inc_steps % check_every == 0) SBO_TRY(drift_check(ctx));
  } else {
    ctx->inc_steps = 0;
    ctx->force_refresh = false;
  }
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
  SBO_TRY(launch_sets_step(ctx, beta, f_min));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
  SBO_TRY(comm_gather_reduce(ctx));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));
  ctx->source = PosteriorSource::kStep;
  ctx->last_beta = beta;
  ctx->posterior_valid = true;
  ctx->posterior_n = ctx->n;
  *kernel_used = which;
  return SBO_OK;
}

int32_t finish_step(sbo_ctx* ctx, int kernel_used, sbo_step_out* out) {
  SBO_CUDA(ctx, cudaMemcpyAsync(ctx->h_result, ctx->d_result, sizeof(SetsResult), cudaMemcpyDeviceToHost, ctx->stream));
  double drift[2] = {0.0, 0.0};
  const bool drift_read = ctx->drift_pending;
  if (drift_read)
    SBO_CUDA(ctx, cudaMemcpyAsync(drift, ctx->d_drift_out, sizeof(drift), cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[4], ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (drift_read) {
    ctx->drift_pending = false;
    ctx->inc_drift = (float)std::max(drift[0], drift[1]);
    const double tol = ctx->cfg.precision == SBO_FP64 ? 1e-10 : 1e-5;
    if (std::max(drift[0], drift[1]) > 0.25 * tol) ctx->force_refresh = true;   // the next incremental step refreshes
  }
  if (out) {
    float ms[4] = {0, 0, 0, 0};
    cudaEventElapsedTime(&ms[0], ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&ms[1], ctx->ev[1], ctx->ev[2]);
    cudaEventElapsedTime(&ms[2], ctx->ev[2], ctx->ev[3]);
    cudaEventElapsedTime(&ms[3], ctx->ev[0], ctx->ev[4]);
    fill_out(ctx, *ctx->h_result, kernel_used, ms, out);
  }
  return SBO_OK;
}

}  // namespace
}  // namespace sbo

using namespace sbo;

extern "C" {

int32_t sbo_abi_version(void) { return SBO_ABI_VERSION; }

const char* sbo_status_string(int32_t s) {
  switch (s) {
    case SBO_OK: return "ok";
    case SBO_ERR_INVALID_ARG: return "invalid argument";
    case SBO_ERR_NO_DEVICE: return "no sm_100 CUDA device";
    case SBO_ERR_CUDA: return "CUDA error";
    case SBO_ERR_NOT_POSDEF: return "covariance not positive definite";
    case SBO_ERR_CAPACITY: return "capacity exceeded";
    case SBO_ERR_STATE: return "call out of order";
    case SBO_ERR_COMM: return "communicator error";
    default: return "unknown status";
  }
}

const char* sbo_last_error(const sbo_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int32_t sbo_create(const sbo_config* cfg, sbo_ctx** out) {
  if (!cfg || !out) return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_create: null argument");
  *out = nullptr;
  if (!(cfg->sigma_f > 0.0 && cfg->length_scale > 0.0 && cfg->noise_var >= 0.0 && std::isfinite(cfg->prior_mean) &&
        std::isfinite(cfg->sigma_f) && std::isfinite(cfg->length_scale) && std::isfinite(cfg->noise_var)))
    return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_create: sigma_f, length_scale must be > 0, noise_var >= 0, all finite");
  if (cfg->precision != SBO_FP32 && cfg->precision != SBO_FP64)
    return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_create: precision must be SBO_FP32 or SBO_FP64");
  if (cfg->kernel < SBO_KERNEL_AUTO || cfg->kernel > SBO_KERNEL_TCGEN05 ||
      (cfg->kernel == SBO_KERNEL_TCGEN05 && cfg->precision != SBO_FP32))
    return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_create: bad kernel choice (tcgen05 is fp32-mode only)");
  if (cfg->max_obs < 1 || cfg->max_cells < 1) return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_create: max_obs and max_cells must be >= 1");
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
    return fail(nullptr, SBO_ERR_NO_DEVICE, "sbo_create: no CUDA device (this library has no CPU fallback)");
  if (cfg->device < 0 || cfg->device >= count) return fail(nullptr, SBO_ERR_INVALID_ARG, "sbo_create: device ordinal out of range");
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, cfg->device) != cudaSuccess || prop.major != 10)
    return fail(nullptr, SBO_ERR_NO_DEVICE, "sbo_create: device is not compute capability 10.x (built for sm_100a only)");

  sbo_ctx* c = new (std::nothrow) sbo_ctx();
  if (!c) return fail(nullptr, SBO_ERR_CUDA, "sbo_create: out of host memory");
  c->cfg = *cfg;
  c->device = cfg->device;
  c->sm_count = prop.multiProcessorCount;
  c->n_cap = (cfg->max_obs + kObsPad - 1) / kObsPad * kObsPad;
  const size_t nc = (size_t)c->n_cap, es = elem_size(c), mc = (size_t)cfg->max_cells;

#define CREATE_CUDA(call)                                               \
  do {                                                                  \
    cudaError_t e__ = (call);                                           \
    if (e__ != cudaSuccess) {                                           \
      fail(nullptr, SBO_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
      free_all(c);                                                      \
      delete c;                                                         \
      return SBO_ERR_CUDA;                                              \
    }                                                                   \
  } while (0)

  CREATE_CUDA(cudaSetDevice(c->device));
  CREATE_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  c->stream = c->own_stream;
  for (auto& e : c->ev) CREATE_CUDA(cudaEventCreate(&e));
  CREATE_CUDA(cudaMalloc(&c->d_xy, sizeof(double) * 2 * nc));
  CREATE_CUDA(cudaMalloc(&c->d_r, sizeof(double) * nc));
  CREATE_CUDA(cudaMalloc(&c->d_A, sizeof(double) * nc * nc));
  CREATE_CUDA(cudaMalloc(&c->d_Linv, sizeof(double) * nc * nc));
  CREATE_CUDA(cudaMalloc(&c->d_Dinv, sizeof(double) * nc * kFactorNB));
  CREATE_CUDA(cudaMalloc(&c->d_alpha, sizeof(double) * nc));
  CREATE_CUDA(cudaMalloc(&c->d_w, sizeof(double) * nc));
  CREATE_CUDA(cudaMalloc(&c->d_small, sizeof(double) * (2 * (size_t)kMaxAppend * nc + 2 * (size_t)kMaxAppend * kMaxAppend)));
  CREATE_CUDA(cudaMalloc(&c->d_info, 32));   // Cholesky status + measured row 1-norms of L^-1 (factor.cu InfoRecord)
  CREATE_CUDA(cudaMemset(c->d_Linv, 0, sizeof(double) * nc * nc));
  CREATE_CUDA(cudaMemset(c->d_alpha, 0, sizeof(double) * nc));
  CREATE_CUDA(cudaMemset(c->d_xy, 0, sizeof(double) * 2 * nc));
  if (cfg->precision == SBO_FP32) {
    CREATE_CUDA(cudaMalloc(&c->d_planes, sizeof(uint16_t) * 2 * nc * nc));
    CREATE_CUDA(cudaMemset(c->d_planes, 0, sizeof(uint16_t) * 2 * nc * nc));
    CREATE_CUDA(cudaMalloc(&c->d_obs_f32, sizeof(float) * 4 * nc));
    CREATE_CUDA(cudaMemset(c->d_obs_f32, 0, sizeof(float) * 4 * nc));
    CREATE_CUDA(cudaMalloc(&c->d_chunk_bbox, sizeof(float) * 4 * (nc / 64 + 1)));
    CREATE_CUDA(cudaMalloc(&c->d_alpha_f32, sizeof(float) * nc));
    CREATE_CUDA(cudaMemset(c->d_alpha_f32, 0, sizeof(float) * nc));
  }
  CREATE_CUDA(cudaMalloc(&c->d_exec_cols, sizeof(unsigned long long)));   // executed-work counter of the last full refresh
  CREATE_CUDA(cudaMemset(c->d_exec_cols, 0, sizeof(unsigned long long)));
  CREATE_CUDA(cudaMalloc(&c->d_mu, es * mc));
  CREATE_CUDA(cudaMalloc(&c->d_s2, es * mc));
  CREATE_CUDA(cudaMalloc(&c->d_mask, mc + 16));
  CREATE_CUDA(cudaMalloc(&c->d_partials, sizeof(SetsResult) * (size_t)sets_grid_size(c)));
  CREATE_CUDA(cudaMalloc(&c->d_result, sizeof(SetsResult)));
  CREATE_CUDA(cudaMalloc(&c->d_ticket, sizeof(unsigned int)));
  CREATE_CUDA(cudaMemset(c->d_ticket, 0, sizeof(unsigned int)));
  CREATE_CUDA(cudaMalloc(&c->d_sel_out, sizeof(int32_t)));
  CREATE_CUDA(cudaMallocHost(&c->h_result, sizeof(SetsResult)));
  CREATE_CUDA(cudaMallocHost(&c->h_sel_out, sizeof(int32_t)));
#undef CREATE_CUDA

  if (cfg->precision == SBO_FP32 && cfg->kernel != SBO_KERNEL_SIMT) {
    const int32_t st = tc_init(c);
    if (st != SBO_OK && cfg->kernel == SBO_KERNEL_TCGEN05) {
      g_create_error = c->err;
      free_all(c);
      delete c;
      return st;
    }
  }
  *out = c;
  return SBO_OK;
}

int32_t sbo_destroy(sbo_ctx* ctx) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  if (ctx->nccl_comm) sbo_comm_destroy(ctx);
  free_all(ctx);
  delete ctx;
  return SBO_OK;
}

int32_t sbo_set_stream(sbo_ctx* ctx, void* cuda_stream) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  ctx->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  return SBO_OK;
}

int32_t sbo_synchronize(sbo_ctx* ctx) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SBO_OK;
}

int32_t sbo_set_grid(sbo_ctx* ctx, double x0, double y0, double rx, double ry, int32_t W, int32_t H, int32_t row0,
                     int32_t row1) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (!(std::isfinite(x0) && std::isfinite(y0) && std::isfinite(rx) && std::isfinite(ry)) || W < 1 || H < 1 || row0 < 0 ||
      row1 < row0 || row1 > H)
    return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_set_grid: need finite origin/pitch, W,H >= 1, 0 <= row0 <= row1 <= H");
  const int64_t m_local = (int64_t)(row1 - row0) * W;
  if (m_local > ctx->cfg.max_cells) return fail(ctx, SBO_ERR_CAPACITY, "sbo_set_grid: shard has more cells than max_cells");
  ctx->x0 = x0; ctx->y0 = y0; ctx->rx = rx; ctx->ry = ry;
  ctx->W = W; ctx->H = H; ctx->row0 = row0; ctx->row1 = row1;
  ctx->cell0 = (int64_t)row0 * W;
  ctx->m_local = m_local;
  cudaSetDevice(ctx->device);
  tc_eager_cancel(ctx);
  ++ctx->grid_epoch;
  ctx->grid_set = true;
  ctx->tc_skip_plan = -1;
  ctx->posterior_valid = false;
  ctx->source = PosteriorSource::kNone;
  ctx->mask_mirror = nullptr;   // another shard / layout: the next step copies the whole mask
  return SBO_OK;
}

int32_t sbo_num_observations(const sbo_ctx* ctx) { return ctx ? ctx->n : SBO_ERR_INVALID_ARG; }

int64_t sbo_launch_count(const sbo_ctx* ctx) { return ctx ? ctx->launches : 0; }

int32_t sbo_tc_plan(sbo_ctx* ctx, int32_t* cta_group, int32_t* skipping, double* kept_fraction) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  const bool skip = ctx->tc_ok && ctx->grid_set ? tc_plan_skip(ctx) : false;
  if (cta_group) *cta_group = ctx->tc_cg;
  if (skipping) *skipping = skip ? 1 : 0;
  if (kept_fraction) *kept_fraction = skip ? ctx->tc_kept_frac : 1.0;
  return SBO_OK;
}

static int* option_slot(sbo_ctx* c, int32_t option) {
  switch (option) {
    case SBO_OPT_TC_CTA_GROUP: return &c->tc_cg;
    case SBO_OPT_TC_PAIR: return &c->opt.tc_pair;
    case SBO_OPT_TC_EAGER: return &c->opt.tc_eager;
    case SBO_OPT_TC_INC: return &c->opt.tc_inc;
    case SBO_OPT_THIN: return &c->opt.thin;
    case SBO_OPT_THIN_2D: return &c->opt.thin_2d;
    case SBO_OPT_TC_GRADED: return &c->opt.tc_graded;
    case SBO_OPT_TC_STRIP: return &c->opt.tc_strip;
    case SBO_OPT_REFRESH_EVERY: return &c->opt.refresh_every;
    case SBO_OPT_DRIFT_CHECK_EVERY: return &c->opt.drift_check_every;
    case SBO_OPT_DRIFT_ROWS: return &c->opt.drift_rows;
    case SBO_OPT_DIST_FIT: return &c->opt.dist_fit;
    case SBO_OPT_TC_PHASES: return &c->opt.tc_phases;
    case SBO_OPT_MASK_DELTA: return &c->opt.mask_delta;
    default: return nullptr;
  }
}

int32_t sbo_set_option(sbo_ctx* ctx, int32_t option, int32_t value) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (option == SBO_OPT_TC_SKIP) {
    ctx->tc_skip = value != 0;
    ctx->tc_skip_plan = -1;
    return SBO_OK;
  }
  int* slot = option_slot(ctx, option);
  if (!slot) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_set_option: unknown option");
  if (value < 0 && option != SBO_OPT_REFRESH_EVERY && option != SBO_OPT_DRIFT_CHECK_EVERY) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_set_option: negative value");
  if (option == SBO_OPT_TC_CTA_GROUP) {
    if (value != 1 && value != 2) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_set_option: SBO_OPT_TC_CTA_GROUP is 1 or 2");
    SBO_CUDA(ctx, cudaSetDevice(ctx->device));
    tc_eager_cancel(ctx);
    cudaStreamSynchronize(ctx->stream);
    return ctx->tc_ok ? tc_set_cta_group(ctx, value) : SBO_OK;
  }
  if (option == SBO_OPT_TC_EAGER && value == 0) tc_eager_cancel(ctx);
  if (option == SBO_OPT_TC_GRADED) ctx->tc_skip_plan = -1;
  if (option == SBO_OPT_MASK_DELTA) {
    ctx->mask_mirror = nullptr;   // flags were not kept while the option was off: the next step copies the whole mask
    if (value && !ctx->d_mask_dirty) {
      const size_t nseg = ((size_t)ctx->cfg.max_cells >> kMaskSegShift) + 2;
      SBO_CUDA(ctx, cudaSetDevice(ctx->device));
      SBO_CUDA(ctx, cudaMalloc(&ctx->d_mask_dirty, nseg * sizeof(unsigned int)));
      SBO_CUDA(ctx, cudaMemset(ctx->d_mask_dirty, 0, nseg * sizeof(unsigned int)));
      SBO_CUDA(ctx, cudaMallocHost(&ctx->h_mask_dirty, nseg * sizeof(unsigned int)));
    }
    value = value != 0;
  }
  *slot = value;
  return SBO_OK;
}

int32_t sbo_get_option(const sbo_ctx* ctx, int32_t option, int32_t* value) {
  if (!ctx || !value) return SBO_ERR_INVALID_ARG;
  if (option == SBO_OPT_TC_SKIP) {
    *value = ctx->tc_skip ? 1 : 0;
    return SBO_OK;
  }
  const int* slot = option_slot(const_cast<sbo_ctx*>(ctx), option);
  if (!slot) return SBO_ERR_INVALID_ARG;
  *value = *slot;
  return SBO_OK;
}

int32_t sbo_tc_executed_flops(sbo_ctx* ctx, double* flops) {
  if (!ctx || !flops) return SBO_ERR_INVALID_ARG;
  *flops = 0.0;
  if (!ctx->d_exec_cols) return SBO_OK;
  unsigned long long cols = 0;
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  SBO_CUDA(ctx, cudaMemcpyAsync(&cols, ctx->d_exec_cols, sizeof(cols), cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (ctx->cfg.precision == SBO_FP64) *flops = 2.0 * 128.0 * 128.0 * 32.0 * (double)cols;   // DMMA kernel: (tile, row block, stage) visits of 128 cells x 128 rows x 32 observations
  else *flops = 2.0 * 256.0 * 16.0 * (double)cols;   // every counted MMA is M 256 (CTA pair) x N x K 16
  return SBO_OK;
}

int32_t sbo_fit_info(const sbo_ctx* ctx, float* ms, int32_t* shared_ranks) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (ms) *ms = ctx->fit_ms;
  if (shared_ranks) *shared_ranks = ctx->fit_dist_used;
  return SBO_OK;
}

int32_t sbo_mask_copy_info(const sbo_ctx* ctx, int64_t* bytes, int32_t* runs) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (bytes) *bytes = ctx->mask_copied_bytes;
  if (runs) *runs = ctx->mask_copied_runs;
  return SBO_OK;
}

int32_t sbo_tc_plan_detail(sbo_ctx* ctx, double out[6]) {
  if (!ctx || !out) return SBO_ERR_INVALID_ARG;
  return tc_plan_detail(ctx, out);
}

int32_t sbo_device_info(const sbo_ctx* ctx, int32_t* sm_count, int32_t* tcgen05_ok) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (sm_count) *sm_count = ctx->sm_count;
  if (tcgen05_ok) *tcgen05_ok = ctx->tc_ok ? 1 : 0;
  return SBO_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// observations
// ---------------------------------------------------------------------------------------------
extern "C" {

// Sort spatially, upload, factorise.  On failure the factor buffers hold garbage (the caller decides what to restore).
static int32_t fit_from(sbo_ctx* ctx, const double* xy, const double* y, int32_t n) {
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  tc_eager_cancel(ctx);
  // spatial order: bands of two length-scales in y, by x inside a band (see sbo_copy_order)
  std::vector<int32_t> order;
  spatial_order(ctx->cfg.length_scale, xy, n, order);
  std::vector<double> sxy(2 * (size_t)n), r((size_t)n);
  for (int i = 0; i < n; ++i) {
    sxy[2 * i] = xy[2 * (size_t)order[i]];
    sxy[2 * i + 1] = xy[2 * (size_t)order[i] + 1];
    r[i] = y[order[i]] - ctx->cfg.prior_mean;
  }
  if (n > 0) {
    SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_xy, sxy.data(), sizeof(double) * 2 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_r, r.data(), sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // sxy, r are locals
  }
  ctx->order.swap(order);
  ctx->n = n;
  ctx->posterior_valid = false;
  ctx->h_xy_sorted.swap(sxy);
  ctx->tc_skip_plan = -1;
  if (xy != ctx->h_xy_arrival.data()) {   // (the ingestion path refits from these very vectors)
    ctx->h_xy_arrival.assign(xy, xy + 2 * (size_t)n);
    ctx->h_y_arrival.assign(y, y + (size_t)n);
  }
  ctx->n_at_fit = n;
  const int32_t st = factor_full(ctx);
  if (st != SBO_OK) {
    tc_eager_cancel(ctx);
    return st;
  }
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SBO_OK;
}

int32_t sbo_set_observations(sbo_ctx* ctx, const double* xy, const double* y, int32_t n) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (n < 0 || (n > 0 && (!xy || !y))) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_set_observations: null input or n < 0");
  if (n > ctx->n_cap || n > ctx->cfg.max_obs) return fail(ctx, SBO_ERR_CAPACITY, "sbo_set_observations: n exceeds max_obs");
  if (!all_finite(xy, 2 * (size_t)n) || !all_finite(y, (size_t)n))
    return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_set_observations: NaN/Inf in the observations");
  // "ctx left in last good state" (SURVEY B4): the fit overwrites the factor buffers in place (L^-1 alone is 2 GiB at
  // n = 16 384, so there is no second copy); a failed fit is followed by a refit of the observations held before the
  // call -- a cost paid on the failure path only.  The stored posterior stays valid if it belonged to exactly that set.
  const int32_t prev_n = ctx->n;
  const std::vector<double> prev_xy(ctx->h_xy_arrival.begin(), ctx->h_xy_arrival.begin() + 2 * (size_t)prev_n);
  const std::vector<double> prev_y(ctx->h_y_arrival.begin(), ctx->h_y_arrival.begin() + (size_t)prev_n);
  const bool prev_post = ctx->posterior_valid && ctx->posterior_n == prev_n;
  const PosteriorSource prev_source = ctx->source;
  const int32_t st = fit_from(ctx, xy, y, n);
  if (st == SBO_OK) return SBO_OK;
  const std::string why = ctx->err;
  int32_t st2 = SBO_ERR_STATE;
  if (prev_n > 0) st2 = fit_from(ctx, prev_xy.data(), prev_y.data(), prev_n);
  if (st2 == SBO_OK) {
    ctx->posterior_valid = prev_post;
    ctx->source = prev_source;
    ctx->err = why + " (the previous " + std::to_string(prev_n) + " observations were restored)";
    return st;
  }
  // nothing to go back to (first fit, or the restore failed as well): an empty, consistent observation set
  ctx->n = 0;
  ctx->order.clear();
  ctx->h_xy_sorted.clear();
  ctx->h_xy_arrival.clear();
  ctx->h_y_arrival.clear();
  ctx->n_at_fit = 0;
  ctx->posterior_valid = false;
  cudaMemsetAsync(ctx->d_Linv, 0, sizeof(double) * (size_t)ctx->n_cap * ctx->n_cap, ctx->stream);
  cudaMemsetAsync(ctx->d_alpha, 0, sizeof(double) * (size_t)ctx->n_cap, ctx->stream);
  if (ctx->d_planes) cudaMemsetAsync(ctx->d_planes, 0, sizeof(uint16_t) * 2 * (size_t)ctx->n_cap * ctx->n_cap, ctx->stream);
  if (ctx->d_alpha_f32) cudaMemsetAsync(ctx->d_alpha_f32, 0, sizeof(float) * (size_t)ctx->n_cap, ctx->stream);
  cudaStreamSynchronize(ctx->stream);
  ctx->err = why;
  return st;
}

int32_t sbo_append_observations(sbo_ctx* ctx, const double* xy, const double* y, int32_t k) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (k < 0 || (k > 0 && (!xy || !y))) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_append_observations: null input or k < 0");
  if ((int64_t)ctx->n + k > ctx->cfg.max_obs) return fail(ctx, SBO_ERR_CAPACITY, "sbo_append_observations: n + k exceeds max_obs");
  if (!all_finite(xy, 2 * (size_t)k) || !all_finite(y, (size_t)k))
    return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_append_observations: NaN/Inf in the observations");
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  tc_eager_cancel(ctx);
  std::vector<double> r((size_t)k);
  for (int i = 0; i < k; ++i) r[i] = y[i] - ctx->cfg.prior_mean;
  for (int done = 0; done < k;) {
    const int kk = std::min(k - done, (int)kMaxAppend);
    const int n = ctx->n;
    SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_xy + 2 * (size_t)n, xy + 2 * (size_t)done, sizeof(double) * 2 * (size_t)kk,
                                  cudaMemcpyHostToDevice, ctx->stream));
    SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_r + n, r.data() + done, sizeof(double) * (size_t)kk, cudaMemcpyHostToDevice,
                                  ctx->stream));
    SBO_TRY(factor_append(ctx, kk));  // on failure n is unchanged: rows >= n of L^-1 are not part of the state
    for (int i = 0; i < kk; ++i) ctx->order.push_back(n + i);
    ctx->h_xy_sorted.insert(ctx->h_xy_sorted.end(), xy + 2 * (size_t)done, xy + 2 * (size_t)(done + kk));
    ctx->h_xy_arrival.insert(ctx->h_xy_arrival.end(), xy + 2 * (size_t)done, xy + 2 * (size_t)(done + kk));
    ctx->h_y_arrival.insert(ctx->h_y_arrival.end(), y + done, y + done + kk);
    ctx->n = n + kk;
    if (ctx->tc_skip_plan == 0) ctx->tc_skip_plan = -1;  // a decision to skip survives appends; a decision not to is re-examined
    done += kk;
  }
  return SBO_OK;   // factor_append left the stream synchronised
}

// N4 -- observation ingestion.  One record per SpatialMeasurement message as the publisher fills it
// (reference src/turtlesim_spatial_publisher.py:137-145).  The GP's noise level is a hyper-parameter (lpsc.yaml:37),
// so `uncertainty` is carried but not used; z and the stamp likewise.  The count returned is what the mapping node
// publishes as spatial_data_size and what triggers a step in the optimiser node when it changes (:552-566).
int32_t sbo_ingest_measurements(sbo_ctx* ctx, const sbo_measurement* recs, int32_t count, int32_t* spatial_data_size,
                                int32_t* refitted) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (refitted) *refitted = 0;
  if (spatial_data_size) *spatial_data_size = ctx->n;
  if (count < 0 || (count > 0 && !recs)) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_ingest_measurements: null input or count < 0");
  if (count == 0) return SBO_OK;
  if ((int64_t)ctx->n + count > ctx->cfg.max_obs) return fail(ctx, SBO_ERR_CAPACITY, "sbo_ingest_measurements: n + count exceeds max_obs");
  std::vector<double> xy(2 * (size_t)count), y((size_t)count);
  for (int i = 0; i < count; ++i) {
    xy[2 * i] = recs[i].x; xy[2 * i + 1] = recs[i].y; y[i] = recs[i].value;
  }
  if (!all_finite(xy.data(), xy.size()) || !all_finite(y.data(), y.size()))
    return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_ingest_measurements: NaN/Inf in a measurement");
  // Appended observations sit behind the spatially ordered ones: fine for a few, but the chunk boxes of the
  // tcgen05 kernel (and its rounding behaviour) want the spatial order back once the tail is long.  Refit from
  // all observations, re-sorted, when the tail would exceed a quarter of the last full fit (at least 64).
  const int64_t tail = (int64_t)ctx->n - ctx->n_at_fit + count;
  const bool refit = ctx->n == 0 || tail > std::max<int64_t>(64, ctx->n_at_fit / 4);
  int32_t st;
  if (refit) {
    std::vector<double> all_xy(ctx->h_xy_arrival.begin(), ctx->h_xy_arrival.begin() + 2 * (size_t)ctx->n);
    std::vector<double> all_y(ctx->h_y_arrival.begin(), ctx->h_y_arrival.begin() + (size_t)ctx->n);
    all_xy.insert(all_xy.end(), xy.begin(), xy.end());
    all_y.insert(all_y.end(), y.begin(), y.end());
    st = sbo_set_observations(ctx, all_xy.data(), all_y.data(), (int32_t)all_y.size());   // failure: previous set restored
    if (refitted && st == SBO_OK) *refitted = 1;
  } else {
    st = sbo_append_observations(ctx, xy.data(), y.data(), count);
  }
  if (spatial_data_size) *spatial_data_size = ctx->n;
  return st;
}

// N3 -- reference :650-706 in one call: the goal itself when it lies within the (eroded) safe polygon, else the
// frontier cell chosen by GetNextSubgoal (:508-549, on the GPU), its coordinates (the x_coords / y_coords the
// reference reads at :671-672 are the grid formula) run through polydist against the polygon's outer ring.
int32_t sbo_subgoal_point(sbo_ctx* ctx, const int32_t* frontier, int64_t F, double goal_x, double goal_y,
                          const double* rings_xy, const int32_t* ring_sizes, int32_t n_rings, double* subgoal_xy,
                          int32_t* cell, int32_t* kind, double* dist) {
  if (!ctx || !subgoal_xy || !kind) return SBO_ERR_INVALID_ARG;
  *kind = SBO_SUBGOAL_NONE;
  if (cell) *cell = -1;
  if (dist) *dist = 0.0;
  int32_t within = 0;
  const int32_t st = sbo_point_within_polygon(rings_xy, ring_sizes, n_rings, goal_x, goal_y, &within);
  if (st != SBO_OK) return fail(ctx, st, "sbo_subgoal_point: bad polygon arguments");
  if (within) {
    subgoal_xy[0] = goal_x; subgoal_xy[1] = goal_y;
    *kind = SBO_SUBGOAL_GOAL;
    return SBO_OK;
  }
  int32_t c = -1;
  SBO_TRY(sbo_select_subgoal(ctx, frontier, F, goal_x, goal_y, &c));
  if (c < 0) return SBO_OK;   // "No valid subgoal found" (:700-703)
  const int iy = c / ctx->W, ix = c - iy * ctx->W;
  const double raw_x = ctx->x0 + (double)ix * ctx->rx, raw_y = ctx->y0 + (double)iy * ctx->ry;
  double d = 0.0;
  SBO_TRY(sbo_polydist(rings_xy, n_rings > 0 ? ring_sizes[0] : 0, raw_x, raw_y, subgoal_xy, &d));
  if (cell) *cell = c;
  if (dist) *dist = d;
  *kind = SBO_SUBGOAL_PROJECTED;
  return SBO_OK;
}

int32_t sbo_copy_linv(sbo_ctx* ctx, double* host_dst) {
  if (!ctx || !host_dst) return SBO_ERR_INVALID_ARG;
  const size_t n = (size_t)ctx->n;
  if (n == 0) return SBO_OK;
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  SBO_CUDA(ctx, cudaMemcpy2DAsync(host_dst, n * sizeof(double), ctx->d_Linv, (size_t)ctx->n_cap * sizeof(double),
                                  n * sizeof(double), n, cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SBO_OK;
}

int32_t sbo_copy_alpha(sbo_ctx* ctx, double* host_dst) {
  if (!ctx || !host_dst) return SBO_ERR_INVALID_ARG;
  if (ctx->n == 0) return SBO_OK;
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  std::vector<double> tmp((size_t)ctx->n);
  SBO_CUDA(ctx, cudaMemcpyAsync(tmp.data(), ctx->d_alpha, sizeof(double) * (size_t)ctx->n, cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int i = 0; i < ctx->n; ++i) host_dst[ctx->order[i]] = tmp[i];  // back to the caller's order
  return SBO_OK;
}

int32_t sbo_copy_order(sbo_ctx* ctx, int32_t* host_dst) {
  if (!ctx || !host_dst) return SBO_ERR_INVALID_ARG;
  for (int i = 0; i < ctx->n; ++i) host_dst[i] = ctx->order[i];
  return SBO_OK;
}

// ---------------------------------------------------------------------------------------------
// the BO step
// ---------------------------------------------------------------------------------------------
// The step's mask on its way to the host.  Default: the whole shard.  SBO_OPT_MASK_DELTA: when `mask_host` is the buffer
// the previous copy went to (and the shard is the same), only the 64 KiB segments the sets pass flagged as changed since
// then -- contiguous flagged segments travel as one copy; the flags accumulate on the device over steps that copy nothing
// (sbo_step_device) and are cleared by the copy that consumes them.  One extra stream synchronisation (the flags).
static int32_t copy_mask_out(sbo_ctx* ctx, uint8_t* mask_host) {
  const size_t m = (size_t)ctx->m_local;
  ctx->mask_copied_bytes = 0;
  ctx->mask_copied_runs = 0;
  if (!mask_host || m == 0) return SBO_OK;
  if (!ctx->opt.mask_delta) {
    SBO_CUDA(ctx, cudaMemcpyAsync(mask_host, ctx->d_mask, m, cudaMemcpyDeviceToHost, ctx->stream));
    ctx->mask_copied_bytes = (int64_t)m;
    ctx->mask_copied_runs = 1;
    return SBO_OK;
  }
  const size_t seg = (size_t)1 << kMaskSegShift, nseg = (m + seg - 1) >> kMaskSegShift;
  const bool mirrored = ctx->mask_mirror == mask_host && ctx->mask_mirror_m == (int64_t)m;
  ctx->mask_mirror = nullptr;   // until this copy is known to have completed (step_common)
  if (!mirrored) {
    SBO_CUDA(ctx, cudaMemcpyAsync(mask_host, ctx->d_mask, m, cudaMemcpyDeviceToHost, ctx->stream));
    SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_mask_dirty, 0, nseg * sizeof(unsigned int), ctx->stream));
    ctx->mask_copied_bytes = (int64_t)m;
    ctx->mask_copied_runs = 1;
    return SBO_OK;
  }
  SBO_CUDA(ctx, cudaMemcpyAsync(ctx->h_mask_dirty, ctx->d_mask_dirty, nseg * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_mask_dirty, 0, nseg * sizeof(unsigned int), ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  size_t s = 0;
  while (s < nseg) {
    if (!ctx->h_mask_dirty[s]) { ++s; continue; }
    size_t e = s + 1;
    while (e < nseg && ctx->h_mask_dirty[e]) ++e;
    const size_t off = s << kMaskSegShift, len = std::min(m, e << kMaskSegShift) - off;
    SBO_CUDA(ctx, cudaMemcpyAsync(mask_host + off, ctx->d_mask + off, len, cudaMemcpyDeviceToHost, ctx->stream));
    ctx->mask_copied_bytes += (int64_t)len;
    ctx->mask_copied_runs++;
    s = e;
  }
  return SBO_OK;
}

static int32_t step_common(sbo_ctx* ctx, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out, bool incremental) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  int used = 0;
  const int32_t st = run_step(ctx, beta, f_min, incremental, &used);
  if (st != SBO_OK) {
    ctx->mask_mirror = nullptr;   // d_mask may or may not have been rewritten
    return st;
  }
  SBO_TRY(copy_mask_out(ctx, mask_host));
  SBO_TRY(finish_step(ctx, used, out));
  if (ctx->opt.mask_delta && mask_host && ctx->m_local > 0) {   // the copy has completed: mask_host mirrors d_mask
    ctx->mask_mirror = mask_host;
    ctx->mask_mirror_m = ctx->m_local;
  }
  return SBO_OK;
}

int32_t sbo_step(sbo_ctx* ctx, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out) {
  return step_common(ctx, beta, f_min, mask_host, out, false);
}

int32_t sbo_step_incremental(sbo_ctx* ctx, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out) {
  return step_common(ctx, beta, f_min, mask_host, out, true);
}

int32_t sbo_step_device(sbo_ctx* ctx, double beta, double f_min, sbo_step_out* out) {
  return step_common(ctx, beta, f_min, nullptr, out, false);
}

static int32_t ensure_array_staging(sbo_ctx* ctx, int64_t m) {
  if (ctx->d_mu64 && ctx->stage_cap >= m) return SBO_OK;
  for (double** p : {&ctx->d_mu64, &ctx->d_std64, &ctx->d_q0, &ctx->d_q1}) {
    if (*p) cudaFree(*p);
    *p = nullptr;
    SBO_CUDA(ctx, cudaMalloc(p, sizeof(double) * (size_t)(m > 0 ? m : 1)));
  }
  ctx->stage_cap = m;
  return SBO_OK;
}

int32_t sbo_compute_sets(sbo_ctx* ctx, const double* mu, const double* std_dev, int64_t m, double beta, double f_min,
                         uint8_t* mask_host, double* q0, double* q1, sbo_step_out* out) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (m < 0 || (m > 0 && (!mu || !std_dev))) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_compute_sets: null input or m < 0");
  if (m > ctx->cfg.max_cells) return fail(ctx, SBO_ERR_CAPACITY, "sbo_compute_sets: m exceeds max_cells");
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  SBO_TRY(ensure_array_staging(ctx, m));
  ctx->source = PosteriorSource::kHostArrays;
  ctx->source_m = m;
  ctx->last_beta = beta;
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
  if (m > 0) {
    SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_mu64, mu, sizeof(double) * (size_t)m, cudaMemcpyHostToDevice, ctx->stream));
    SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_std64, std_dev, sizeof(double) * (size_t)m, cudaMemcpyHostToDevice, ctx->stream));
  }
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
  ctx->mask_mirror = nullptr;   // d_mask is rewritten without changed-segment flags
  SBO_TRY(launch_sets_arrays(ctx, ctx->d_mu64, ctx->d_std64, m, beta, f_min, ctx->d_mask, q0 ? ctx->d_q0 : nullptr,
                             q1 ? ctx->d_q1 : nullptr));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));
  if (m > 0) {
    if (mask_host) SBO_CUDA(ctx, cudaMemcpyAsync(mask_host, ctx->d_mask, (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    if (q0) SBO_CUDA(ctx, cudaMemcpyAsync(q0, ctx->d_q0, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    if (q1) SBO_CUDA(ctx, cudaMemcpyAsync(q1, ctx->d_q1, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
  }
  return finish_step(ctx, 0, out);
}

int32_t sbo_compute_sets_device(sbo_ctx* ctx, const double* mu_dev, const double* std_dev, int64_t m, double beta,
                                double f_min, uint8_t* mask_dev, sbo_step_out* out) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (m < 0 || (m > 0 && (!mu_dev || !std_dev))) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_compute_sets_device: null input");
  if (!mask_dev && m > ctx->cfg.max_cells) return fail(ctx, SBO_ERR_CAPACITY, "sbo_compute_sets_device: m exceeds max_cells");
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
  if (!mask_dev) ctx->mask_mirror = nullptr;   // d_mask is rewritten without changed-segment flags
  SBO_TRY(launch_sets_arrays(ctx, mu_dev, std_dev, m, beta, f_min, mask_dev ? mask_dev : ctx->d_mask, nullptr, nullptr));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));
  const PosteriorSource keep = ctx->source;
  const int64_t keep_m = ctx->source_m;
  ctx->source = PosteriorSource::kHostArrays;  // so that n_cells reports m
  ctx->source_m = m;
  const int32_t st = finish_step(ctx, 0, out);
  ctx->source = keep;
  ctx->source_m = keep_m;
  return st;
}

// ---------------------------------------------------------------------------------------------
// frontier stage helpers (reference FindSafetyContourIndices :418-497, safe-cell gather :944-949)
// ---------------------------------------------------------------------------------------------
int32_t sbo_frontier_pixel_maps(sbo_ctx* ctx, int32_t* col_of_px, int32_t* row_of_py) {
  if (!ctx || !col_of_px || !row_of_py) return SBO_ERR_INVALID_ARG;
  if (!ctx->grid_set) return fail(ctx, SBO_ERR_STATE, "sbo_frontier_pixel_maps: grid not set");
  const int W = ctx->W, H = ctx->H;
  // :431-434 -- bounds of the cell coordinates truncated to int (coordinates ascend with the index on a regular grid)
  auto axis = [&](double o, double r, int dim, int32_t* map, const char* name) -> int32_t {
    const double c0 = o, c1 = o + (double)(dim - 1) * r;
    const int mn = (int)std::min(c0, c1), mx = (int)std::max(c0, c1);
    if (mx == mn)
      return fail(ctx, SBO_ERR_INVALID_ARG, std::string("sbo_frontier_pixel_maps: ") + name +
                                                " extent below one unit: the reference divides by zero here (:450-451)");
    for (int i = 0; i < dim; ++i) map[i] = -1;
    for (int i = 0; i < dim; ++i) {               // ascending cell order: the last writer of a pixel wins (:468-475)
      const double c = o + (double)i * r;         // same expression as the grid formula of sbo_set_grid
      const int px = (int)((c - mn) / (double)(mx - mn) * dim);  // :450-451
      if (px >= 0 && px < dim) map[px] = i;       // :453
    }
    return SBO_OK;
  };
  SBO_TRY(axis(ctx->x0, ctx->rx, W, col_of_px, "x"));
  SBO_TRY(axis(ctx->y0, ctx->ry, H, row_of_py, "y"));
  return SBO_OK;
}

int32_t sbo_frontier_image(sbo_ctx* ctx, uint8_t* img_host, int32_t* col_of_px, int32_t* row_of_py) {
  if (!ctx || !img_host || !col_of_px || !row_of_py) return SBO_ERR_INVALID_ARG;
  if (ctx->source != PosteriorSource::kStep)
    return fail(ctx, SBO_ERR_STATE, "sbo_frontier_image: no safe mask on the device yet (call sbo_step first)");
  SBO_TRY(sbo_frontier_pixel_maps(ctx, col_of_px, row_of_py));
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  return frontier_image(ctx, col_of_px, row_of_py, img_host);
}

int32_t sbo_compact_cells(sbo_ctx* ctx, int32_t mode, int32_t* idx_host, int64_t cap, int64_t* count) {
  if (!ctx || !count || cap < 0 || (cap > 0 && !idx_host)) return SBO_ERR_INVALID_ARG;
  if (mode != SBO_CELLS_SAFE && mode != SBO_CELLS_BORDER) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_compact_cells: bad mode");
  if (ctx->source != PosteriorSource::kStep)
    return fail(ctx, SBO_ERR_STATE, "sbo_compact_cells: no safe mask on the device yet (call sbo_step first)");
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  return compact_cells(ctx, mode, idx_host, cap, count);
}

}  // extern "C"
namespace sbo {
int32_t select_upload(sbo_ctx* ctx, const int32_t* frontier, int64_t F) {
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  if (F > ctx->sel_cap) {
    for (void** p : {(void**)&ctx->d_frontier, (void**)&ctx->d_sel_keys, (void**)&ctx->d_sel_width}) {
      if (*p) cudaFree(*p);
      *p = nullptr;
    }
    const int64_t cap = F + F / 2 + 1024;
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_frontier, sizeof(int32_t) * (size_t)cap));
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_sel_keys, sizeof(unsigned long long) * (size_t)cap));
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_sel_width, sizeof(double) * (size_t)cap));
    ctx->sel_cap = cap;
  }
  SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_frontier, frontier, sizeof(int32_t) * (size_t)F, cudaMemcpyHostToDevice, ctx->stream));
  return SBO_OK;
}
}  // namespace sbo
extern "C" {

int32_t sbo_select_subgoal(sbo_ctx* ctx, const int32_t* frontier, int64_t F, double goal_x, double goal_y,
                           int32_t* out_cell) {
  if (!ctx || !out_cell) return SBO_ERR_INVALID_ARG;
  *out_cell = -1;
  if (F < 0 || (F > 0 && !frontier)) return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_select_subgoal: null frontier or F < 0");
  if (F == 0) return SBO_OK;  // "No frontier points found" (:503-506)
  if (ctx->source == PosteriorSource::kNone)
    return fail(ctx, SBO_ERR_STATE, "sbo_select_subgoal: no posterior yet (call sbo_step or sbo_compute_sets first)");
  if (!ctx->grid_set) return fail(ctx, SBO_ERR_STATE, "sbo_select_subgoal: grid not set");
  if (F > 0xFFFFFFFFll) return fail(ctx, SBO_ERR_CAPACITY, "sbo_select_subgoal: frontier longer than 2^32");
  SBO_TRY(select_upload(ctx, frontier, F));
  SBO_TRY(launch_select(ctx, ctx->d_frontier, F, goal_x, goal_y));
  SBO_CUDA(ctx, cudaMemcpyAsync(ctx->h_sel_out, ctx->d_sel_out, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (*ctx->h_sel_out == -2)
    return fail(ctx, SBO_ERR_INVALID_ARG, "sbo_select_subgoal: frontier holds a cell outside this context's cells");
  *out_cell = *ctx->h_sel_out;
  return SBO_OK;
}

// ---------------------------------------------------------------------------------------------
// read-back
// ---------------------------------------------------------------------------------------------
static int32_t copy_posterior(sbo_ctx* ctx, const void* src, void* dst, bool to_f64) {
  if (!ctx || !dst) return SBO_ERR_INVALID_ARG;
  if (!ctx->posterior_valid) return fail(ctx, SBO_ERR_STATE, "sbo_copy_*: no posterior yet (call sbo_step first)");
  const size_t m = (size_t)ctx->m_local;
  if (m == 0) return SBO_OK;
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!to_f64 || ctx->cfg.precision == SBO_FP64) {
    SBO_CUDA(ctx, cudaMemcpyAsync(dst, src, elem_size(ctx) * m, cudaMemcpyDeviceToHost, ctx->stream));
    SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SBO_OK;
  }
  std::vector<float> tmp(m);
  SBO_CUDA(ctx, cudaMemcpyAsync(tmp.data(), src, sizeof(float) * m, cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  double* d = static_cast<double*>(dst);
  for (size_t i = 0; i < m; ++i) d[i] = (double)tmp[i];
  return SBO_OK;
}

int32_t sbo_copy_mu(sbo_ctx* ctx, void* host_dst) { return copy_posterior(ctx, ctx ? ctx->d_mu : nullptr, host_dst, false); }
int32_t sbo_copy_sigma2(sbo_ctx* ctx, void* host_dst) { return copy_posterior(ctx, ctx ? ctx->d_s2 : nullptr, host_dst, false); }
int32_t sbo_copy_mu_f64(sbo_ctx* ctx, double* host_dst) { return copy_posterior(ctx, ctx ? ctx->d_mu : nullptr, host_dst, true); }
int32_t sbo_copy_sigma2_f64(sbo_ctx* ctx, double* host_dst) { return copy_posterior(ctx, ctx ? ctx->d_s2 : nullptr, host_dst, true); }

int32_t sbo_copy_mask(sbo_ctx* ctx, uint8_t* host_dst) {
  if (!ctx || !host_dst) return SBO_ERR_INVALID_ARG;
  if (ctx->source == PosteriorSource::kNone) return fail(ctx, SBO_ERR_STATE, "sbo_copy_mask: no step yet");
  const size_t m = (size_t)(ctx->source == PosteriorSource::kHostArrays ? ctx->source_m : ctx->m_local);
  if (m == 0) return SBO_OK;
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  SBO_CUDA(ctx, cudaMemcpyAsync(host_dst, ctx->d_mask, m, cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SBO_OK;
}

int32_t sbo_merge_step_out(const sbo_step_out* parts, int32_t count, sbo_step_out* merged) {
  if (!parts || !merged || count < 1) return SBO_ERR_INVALID_ARG;
  sbo_step_out r = parts[0];
  auto host_better = [](const sbo_argmax& a, const sbo_argmax& b) {
    return a.idx >= 0 && (b.idx < 0 || a.value > b.value || (a.value == b.value && a.idx < b.idx));
  };
  for (int i = 1; i < count; ++i) {
    const sbo_step_out& o = parts[i];
    if (host_better(o.argmax_ucb, r.argmax_ucb)) r.argmax_ucb = o.argmax_ucb;
    if (host_better(o.argmax_width, r.argmax_width)) r.argmax_width = o.argmax_width;
    r.safe_count += o.safe_count;
    r.clamped_sigma2 += o.clamped_sigma2;
    r.low_sigma2 += o.low_sigma2;
    r.refreshed = r.refreshed || o.refreshed;
    r.drift = std::max(r.drift, o.drift);
    r.n_cells += o.n_cells;
    r.ms_posterior = std::max(r.ms_posterior, o.ms_posterior);
    r.ms_sets = std::max(r.ms_sets, o.ms_sets);
    r.ms_comm = std::max(r.ms_comm, o.ms_comm);
    r.ms_total = std::max(r.ms_total, o.ms_total);
  }
  *merged = r;
  return SBO_OK;
}

}  // extern "C"
