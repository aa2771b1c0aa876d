// common.cuh -- context, error plumbing and small device helpers shared by the
// kernels of libsbo.so (C-ABI in include/sbo.h).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/sbo.h"

namespace sbo {

constexpr int kFactorNB = 128;   // block size of the fp64 factorisation kernels
constexpr int kObsPad = 256;     // observation capacity is rounded up to this (tcgen05 N tile)
constexpr int kMaxAppend = 128;  // rows per rank-k append pass
constexpr int kThinMaxRows = 8;  // incremental steps with at most this many new rows take the thin SIMT kernel

// Result record of the CI/mask/argmax pass, written by the device.
struct SetsResult {
  double ucb_val;
  long long ucb_idx;
  double width_val;
  long long width_idx;
  long long safe_count;
  long long clamped;
  long long low_sigma2;   // cells with sigma^2 < 0.02 sigma_f^2
};

// sbo_set_option state (include/sbo.h sbo_option); the defaults are the product path
struct Options {
  int tc_pair = 1, tc_eager = 1, tc_inc = 1, thin = 1, thin_2d = 1, tc_graded = 1, tc_strip = 4;
  int refresh_every = -1;       // -1: by precision (256 fp32, never fp64)
  int drift_check_every = -1;   // -1: by precision (16 fp32, 256 fp64)
  int drift_rows = 2;           // sample blocks per check (a block = 1 grid row; 8 rows for the DMMA kernel's 16 x 8 cell patches)
  int tc_phases = 0;            // full refresh in this many launches over consecutive row-block pair ranges (0: by size)
  int dist_fit = 2;             // fit shared between the ranks / group members: 0 never, 1 always, 2 from 8192 observations on
  int mask_delta = 0;           // steps copy back only the mask segments that changed since the last copy into the same host buffer
};

constexpr int kMaskSegShift = 16;   // SBO_OPT_MASK_DELTA: one changed-flag per 2^16 cells (64 KiB of mask bytes)

struct FitPeers;   // factor.cu: rendezvous of the members of a same-process group during a shared fit

enum class PosteriorSource : int { kNone = 0, kStep = 1, kHostArrays = 2 };

}  // namespace sbo

struct sbo_ctx {
  sbo_config cfg{};
  int device = 0;
  int sm_count = 0;
  bool tc_ok = false;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[6]{};
  cudaStream_t aux_stream = nullptr;      // second chain of the factorisation (rows of L^-1)
  std::vector<cudaEvent_t> ev_factor;     // [0] fork/join, [1 + 2k] diag k done, [2 + 2k] panel k done
  std::string err;
  int64_t launches = 0;
  sbo::Options opt;
  bool attr_factor = false, attr_f64 = false;   // per-context (= per-device) cudaFuncSetAttribute done
  int inc_steps = 0;             // incremental steps since the last full posterior (refresh policy)
  int inc_refreshed = 0;         // the last sbo_step_incremental ran a full refresh
  float inc_drift = -1.f;        // result of the last drift check (-1: none)
  void* d_drift_mu = nullptr;    // drift check scratch: exact mu / sigma^2 of the sample rows (lazy)
  void* d_drift_s2 = nullptr;
  double* d_drift_out = nullptr; // [2] max deviations
  int64_t drift_cap = 0;
  bool drift_pending = false;   // a drift check was queued in this step: finish_step reads it
  bool force_refresh = false;   // the last drift check exceeded tol / 4: the next incremental step is a full refresh

  // grid
  bool grid_set = false;
  double x0 = 0, y0 = 0, rx = 0, ry = 0;
  int W = 0, H = 0, row0 = 0, row1 = 0;
  int64_t cell0 = 0, m_local = 0;

  // observations / factor state (all fp64, leading dimension n_cap)
  int n = 0;
  int n_prev = 0;  // n before the last append (for the incremental step)
  int n_cap = 0;
  std::vector<int32_t> order;  // order[pos] = caller index of the observation stored at pos
  double* d_xy = nullptr;     // [n_cap][2]
  double* d_r = nullptr;      // [n_cap] y - prior_mean
  double* d_A = nullptr;      // [n_cap][n_cap] K, then L in the lower triangle
  double* d_Linv = nullptr;   // [n_cap][n_cap] L^-1, zero above the diagonal and beyond n
  double* d_Dinv = nullptr;   // [n_cap/NB][NB][NB] inverses of the diagonal blocks of L
  double* d_alpha = nullptr;  // [n_cap]
  double* d_w = nullptr;      // [n_cap] work vector
  double* d_small = nullptr;  // append scratch: K21/L21/T [kMaxAppend][n_cap] x2 + S
  int* d_info = nullptr;      // Cholesky status (0 or 1+failing row)
  // tcgen05 operands (fp32 mode): fp16 hi/lo planes of 2^shift_b L^-1, stacked [2*n_cap][n_cap]
  uint16_t* d_planes = nullptr;
  int shift_a = 0, shift_b = 0;  // power-of-two scales of the K* and L^-1 planes (tc_plane_shifts)
  int tc_cg = 2;                 // CTAs per tcgen05.mma group (2 = CTA pair, 1 = single CTA)
  bool tc_skip = true;           // leave out K* chunks that are < 2^-44 sigma_f^2 for a whole cell tile
  std::vector<double> h_xy_arrival, h_y_arrival;   // every observation in arrival order (ingestion refits from these)
  int n_at_fit = 0;              // observations at the last full factorisation
  double tc_kept_frac = 1.0;     // planner's estimate: share of (row block, chunk) visits that survive the box test
  double tc_pass_frac = 1.0;     // ... and share of the dense walk's MMA passes that are issued (ring visits issue 1 of 3)
  double linv_r1 = 0.0;          // max row 1-norm of L^-1, measured by the fit / append (0: unknown)
  int tc_skip_plan = -1;         // cached decision of tc_plan_skip for this grid + observation set (-1: stale)
  std::vector<double> h_xy_sorted;  // host copy of d_xy (internal order), for the skip planner
  unsigned long long* d_exec_cols = nullptr;   // executed-work counter of the last full tcgen05 refresh (sbo_tc_executed_flops)
  float* d_chunk_bbox = nullptr; // [n_cap / 64] float4 scaled bounding boxes of the observation chunks
  void* d_chunk_bbox64 = nullptr; // [n_cap / 64] double4 boxes in the caller's coordinates (fp64 kernel, lazy)
  float* d_alpha_f32 = nullptr;  // [n_cap]
  float* d_obs_f32 = nullptr;    // [n_cap] float4: scaled observation coordinates as fp32 (hi, lo) pairs (posterior_tc.cu)
  void* tmap_planes = nullptr;   // host copy of the CUtensorMap (128 B), 256-row box
  bool planes_valid = false;
  // eager first pairs of the posterior, launched by factor_full on a side stream (posterior_tc.cu)
  cudaStream_t eager_stream = nullptr;
  cudaEvent_t ev_eager_done = nullptr;
  void* d_rs_part = nullptr;     // [rs_cap] float2 running row sums of the partial launch
  void* d_pplanes = nullptr;     // fp16 hi/lo planes of P, the row operand of the incremental GEMM (posterior_tc_inc.cuh), lazy
  size_t pplanes_cap = 0;        // bytes
  int64_t rs_cap = 0;
  bool eager_inflight = false;   // a partial result exists (or is being computed) for eager_n observations on grid eager_grid_epoch
  int eager_pairs = 0, eager_n = 0;
  uint64_t grid_epoch = 0, eager_grid_epoch = 0;

  // posterior of the last step (this shard)
  void* d_mu = nullptr;      // [max_cells] float or double
  void* d_s2 = nullptr;      // [max_cells]
  uint8_t* d_mask = nullptr; // [max_cells]
  // SBO_OPT_MASK_DELTA (api.cu: copy_mask_out): flags of the mask segments whose bytes changed since the host last received them
  unsigned int* d_mask_dirty = nullptr;   // [max_cells >> kMaskSegShift, + 2] (allocated when the option is switched on)
  unsigned int* h_mask_dirty = nullptr;   // pinned, same length
  const uint8_t* mask_mirror = nullptr;   // the host buffer that holds d_mask as of the last copy (null: none / unknown)
  int64_t mask_mirror_m = 0;              // m_local at that copy
  int64_t mask_copied_bytes = 0;          // mask bytes the last step copied to the host
  int32_t mask_copied_runs = 0;           // ... in this many contiguous runs (1: the whole mask)
  double* d_mu64 = nullptr;  // compute_sets staging (lazy) [max_cells]
  double* d_std64 = nullptr;
  double* d_q0 = nullptr;
  double* d_q1 = nullptr;
  int64_t stage_cap = 0;     // capacity (cells) of the four staging arrays above
  sbo::PosteriorSource source = sbo::PosteriorSource::kNone;
  int64_t source_m = 0;
  double last_beta = 0.0;
  bool posterior_valid = false;  // mu/s2 hold a full posterior for n_prev or n observations
  int posterior_n = 0;           // number of observations that posterior was computed with

  // reduction scratch
  sbo::SetsResult* d_partials = nullptr;  // [grid]
  sbo::SetsResult* d_result = nullptr;    // [1] (+ gathered [world])
  unsigned int* d_ticket = nullptr;
  sbo::SetsResult* h_result = nullptr;    // pinned
  int32_t* d_frontier = nullptr;          // select scratch (lazy)
  unsigned long long* d_sel_keys = nullptr;
  double* d_sel_width = nullptr;
  int64_t sel_cap = 0;
  int32_t* h_sel_out = nullptr;  // pinned
  int32_t* d_sel_out = nullptr;

  // frontier / compaction scratch (lazy; frontier.cu)
  uint8_t* d_fr_img = nullptr;
  int32_t* d_fr_lut = nullptr;   // [W + H] pixel -> cell tables
  long long fr_cap = 0;
  int32_t* d_cp_idx = nullptr;   // [m_local]
  unsigned int* d_cp_counts = nullptr;
  long long cp_cap = 0;

  // shared fit (factor_dist.cuh): block columns of the Cholesky / L^-1 work split over the ranks
  cudaStream_t fit_hi = nullptr, fit_lo = nullptr, fit_cm = nullptr;   // Cholesky chain, rows of L^-1, exchange
  double* d_fit_stage = nullptr;   // one panel message: header + D_k + L panel k (NCCL transport)
  double* d_fit_send = nullptr;    // my block columns of L^-1, packed (NCCL transport)
  double* d_fit_recv = nullptr;    // [world] slots of packed block columns
  size_t fit_stage_cap = 0, fit_send_cap = 0, fit_recv_cap = 0;   // doubles
  sbo::FitPeers* fit_peers = nullptr;   // same-process group (not owned); null: NCCL transport when nccl_comm is set
  int fit_rank = 0;
  int fit_dist_used = 0;           // the last full fit was shared between this many ranks (0: replicated)
  float fit_ms = 0.f;              // device time of the last full fit (events on the context's stream)
  cudaEvent_t ev_fit[2]{};

  // communicator
  void* nccl_comm = nullptr;
  int rank = 0, world = 1;
  sbo::SetsResult* d_gather = nullptr;  // [world]
  sbo::SetsResult* h_gather = nullptr;  // pinned [world]
};

namespace sbo {

int32_t fail(sbo_ctx* ctx, int32_t status, const std::string& msg);
int32_t cuda_fail(sbo_ctx* ctx, cudaError_t e, const char* what);

#define SBO_CUDA(ctx, call)                                        \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return sbo::cuda_fail(ctx, e__, #call); \
  } while (0)

#define SBO_TRY(call)                  \
  do {                                 \
    int32_t s__ = (call);              \
    if (s__ != SBO_OK) return s__;     \
  } while (0)

#define SBO_LAUNCH_CHECK(ctx, name)                                   \
  do {                                                                \
    (ctx)->launches++;                                                \
    cudaError_t e__ = cudaGetLastError();                             \
    if (e__ != cudaSuccess) return sbo::cuda_fail(ctx, e__, name);    \
  } while (0)

// ---- entry points of the individual translation units ----------------------

// sets.cu
int32_t launch_sets_step(sbo_ctx* ctx, double beta, double f_min);                       // mu/s2 -> mask, result
int32_t launch_sets_arrays(sbo_ctx* ctx, const double* d_mu, const double* d_std, int64_t m, double beta,
                           double f_min, uint8_t* d_mask, double* d_q0, double* d_q1);   // fp64 mu/std
int sets_grid_size(const sbo_ctx* ctx);

// select.cu
int32_t launch_select(sbo_ctx* ctx, const int32_t* d_frontier, int64_t F, double gx, double gy, int mode = 0);
int32_t select_upload(sbo_ctx* ctx, const int32_t* frontier, int64_t F);   // api.cu: scratch sized for F entries, frontier H2D

// factor.cu
int32_t factor_full(sbo_ctx* ctx);                 // d_xy, d_r, n -> d_A (L), d_Linv, d_alpha, planes
int32_t factor_append(sbo_ctx* ctx, int k);        // rows [n, n+k) already in d_xy / d_r
int32_t refresh_planes(sbo_ctx* ctx, int row_begin, int row_end, cudaStream_t stream = nullptr);   // nullptr: ctx->stream

// plan.cu (host)
void spatial_order(double length_scale, const double* xy, int n, std::vector<int32_t>& order);
double tc_xscale(const sbo_config& cfg);
void chunk_boxes(const double* xy_sorted, int n, double xs, std::vector<double>& box);
void tile_visit_cost(const std::vector<double>& box, int n, double xlo, double xhi, double y, double bits, double ring,
                     double* total, double* kept, double* passes);
double tc_skip_bits_of(const sbo_config& cfg, int n, double r1);
double tc_ring_bits_of(const sbo_config& cfg, int n, double r1, int shift_a);
double f64_skip_bits_of(const sbo_config& cfg, int n, double r1);

// posterior_simt.cu
int32_t launch_posterior_simt(sbo_ctx* ctx, int j_lo, int j_hi, bool accumulate);

// posterior_tc.cu
int32_t tc_init(sbo_ctx* ctx);                     // tensor map, attributes, plane scales; sets ctx->tc_ok
int32_t tc_set_cta_group(sbo_ctx* ctx, int cg);    // re-encodes the tensor map of the planes for 1 or 2 CTAs per MMA group
void tc_plane_shifts(const sbo_config& cfg, int* shift_a, int* shift_b);
int32_t launch_posterior_tc(sbo_ctx* ctx, int j_lo, int j_hi, bool accumulate);
void tc_destroy(sbo_ctx* ctx);
bool tc_plan_skip(sbo_ctx* ctx);                   // worth walking chunk masks on this grid / observation set?
int32_t tc_plan_detail(sbo_ctx* ctx, double out[6]);
bool tc_inc_applicable(sbo_ctx* ctx, int k);       // incremental step with k new rows through the GEMM kernel?
int32_t launch_posterior_tc_inc(sbo_ctx* ctx, int row_lo);
int tc_eager_pairs(sbo_ctx* ctx);                   // row-block pairs factor_full may launch early (0: none)
int32_t tc_eager_launch(sbo_ctx* ctx, int pairs, cudaEvent_t ready);
void tc_eager_cancel(sbo_ctx* ctx);                 // waits for an eager launch in flight and drops its partial result
int32_t tc_pack_observations(sbo_ctx* ctx, cudaStream_t stream = nullptr);       // d_xy -> d_obs_f32 (scaled fp32 hi/lo pairs)

// posterior_thin.cu
int32_t launch_posterior_f64(sbo_ctx* ctx, int j_lo, int j_hi, bool accumulate);   // fp64 mode, DMMA
int32_t f64_chunk_boxes(sbo_ctx* ctx, int max_chunks, const void** boxes, double* skip_r2);   // posterior_f64.cu: chunk boxes + reach of the fp64 block-sparse walk
int32_t launch_posterior_thin(sbo_ctx* ctx, int row_lo);   // incremental step, n - row_lo <= kThinMaxRows

// frontier.cu
int32_t frontier_image(sbo_ctx* ctx, const int32_t* col_of_px, const int32_t* row_of_py, uint8_t* img_host);
int32_t compact_cells(sbo_ctx* ctx, int mode, int32_t* idx_host, int64_t cap, int64_t* count_out);

// comm.cu
int32_t comm_gather_reduce(sbo_ctx* ctx);          // d_result -> d_result (global), over all ranks
int32_t comm_broadcast(sbo_ctx* ctx, void* buf, size_t bytes, int root, cudaStream_t stream);              // in place
int32_t comm_all_gather(sbo_ctx* ctx, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t stream);

// factor.cu: same-process members of a shared fit (group.cu creates one per group and attaches it to the members)
FitPeers* fit_peers_create(int world);
void fit_peers_destroy(FitPeers* p);
void fit_peers_attach(FitPeers* p, sbo_ctx* ctx, int rank);

// ---- device helpers ---------------------------------------------------------

#ifdef __CUDACC__
// Cell coordinate exactly as the oracle computes it: product and sum rounded
// separately (no FMA contraction).
__device__ __forceinline__ double cell_coord(double origin, double pitch, int i) {
  return __dadd_rn(origin, __dmul_rn((double)i, pitch));
}

// Total order of the argmax: larger value first, then lower index.
__device__ __forceinline__ bool better(double v, long long i, double bv, long long bi) {
  return i >= 0 && (bi < 0 || v > bv || (v == bv && i < bi));
}
#endif

}  // namespace sbo
