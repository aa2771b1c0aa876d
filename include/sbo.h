/*
 * sbo.h -- C-ABI of the B200-native safe-BO step library (libsbo.so).
 *
 * Drop-in boundary for the per-step hot path of
 * matthewyjiang/safe_bayesian_optimization's `safe_bayesian_optimization_node`
 * (file references below are into that repository,
 * src/safe_bayesian_optimization_node.cpp unless stated).  The reference has
 * no FFI/plugin interface of its own -- everything on the path is a private
 * member of `OptimizerNode` (:128) -- so each entry point names the member
 * function or the lines of the node it replaces.  INTEGRATION.md shows the
 * edit a maintainer makes in the node.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no C++/torch types.
 *   - every function returns an sbo_status (0 = OK, < 0 = error); nothing
 *     throws across the boundary.  sbo_last_error(ctx) gives the text; the
 *     caller logs and skips the step, as the node does today (:569-572,
 *     :618-621, :503-506).
 *   - one caller thread per context (the node runs a SingleThreadedExecutor,
 *     :1503).  Calls are synchronous unless the name ends in _async.
 *   - host pointers are borrowed for the duration of the call; outputs go to
 *     caller-owned buffers (pinned memory recommended, not required).
 *   - there is no CPU fallback: sbo_create fails without an sm_100 GPU.
 *
 * Grid convention: cell = iy * W + ix, x = x0 + ix * rx, y = y0 + iy * ry
 * (row-major, y outer -- the layout of `safety_image.at<uchar>(y, x)`, :454).
 */
#ifndef SBO_H
#define SBO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SBO_ABI_VERSION 2   /* 2: sbo_step_out gained low_sigma2 / refreshed / drift; options; row-band planner; device groups.
                               Additive since (same version): sbo_fit_info, SBO_OPT_DIST_FIT, SBO_OPT_TC_PHASES, sbo_compute_sets_noalloc,
                               SBO_OPT_MASK_DELTA, sbo_mask_copy_info */

typedef enum {
  SBO_OK = 0,
  SBO_ERR_INVALID_ARG = -1,   /* null pointer, bad size, NaN/Inf input          */
  SBO_ERR_NO_DEVICE = -2,     /* no CUDA device of compute capability 10.x      */
  SBO_ERR_CUDA = -3,          /* CUDA runtime/driver failure (text in last_error) */
  SBO_ERR_NOT_POSDEF = -4,    /* Cholesky met a non-positive pivot              */
  SBO_ERR_CAPACITY = -5,      /* n > max_obs or m > max_cells                   */
  SBO_ERR_STATE = -6,         /* grid / observations / posterior not set yet    */
  SBO_ERR_COMM = -7           /* NCCL failure                                   */
} sbo_status;

typedef enum {
  SBO_FP32 = 0,   /* mu, sigma^2 stored fp32; contraction on tcgen05: 3xFP16 split (hi.hi + hi.lo + lo.hi, kind::f16,
                     fp32 accumulators in TMEM) -- the fp32-grade split, 2^-22 per product.  Achieved accuracy (measured at
                     C2-C4, profiles/parity_*): |d mu| <= 3e-6 max|mu|, |d sigma^2| <= 2e-6 sigma_f^2.  Relative to sigma^2
                     ITSELF that is 1e-5 only where sigma^2 >= 0.2 sigma_f^2: next to observations (sigma^2 down to ~0.008
                     sigma_f^2) fp32 storage of |v|^2 ~ 0.99 sigma_f^2 alone is 6e-8 = 7e-6 sigma^2, and the measured figure
                     is 2.4e-4.  sbo_step_out.low_sigma2 counts the cells with sigma^2 < 0.02 sigma_f^2; use SBO_FP64 when
                     those need 1e-5 of sigma^2. */
  SBO_FP64 = 1    /* mu, sigma^2 stored fp64; contraction on the fp64 pipe (DMMA m8n8k4 / DFMA), 1e-10 */
} sbo_precision;

typedef enum {
  SBO_KERNEL_AUTO = 0,     /* tcgen05 path for SBO_FP32, the DMMA kernel for SBO_FP64 */
  SBO_KERNEL_SIMT = 1,     /* force the SIMT kernel (any precision)             */
  SBO_KERNEL_TCGEN05 = 2,  /* force the tcgen05 kernel (SBO_FP32 only)          */
  SBO_KERNEL_THIN = 3,     /* reported only: the O(m n k) FMA kernel of sbo_step_incremental after <= 8 new rows
                              (fp32 steps on large grids run as a tensor-core GEMM instead and report SBO_KERNEL_TCGEN05) */
  SBO_KERNEL_DMMA = 4      /* reported only: the fp64 mma.sync kernel that SBO_KERNEL_AUTO selects for SBO_FP64 */
} sbo_kernel_choice;

/* GP hyper-parameters of the external mapping node the library replaces
 * (config/lpsc.yaml:35-37) plus capacity limits.  The GP is zero-mean on
 * (y - prior_mean); sigma^2 is the latent variance unless
 * include_noise_in_sigma != 0 (then + noise_var). */
typedef struct {
  double sigma_f;                 /* lpsc.yaml:37 sigma_f                     */
  double length_scale;            /* lpsc.yaml:36 length_scale                */
  double noise_var;               /* lpsc.yaml:35 noise_level, as a variance  */
  double prior_mean;              /* constant mean                            */
  int32_t include_noise_in_sigma; /* 0 = latent variance                      */
  int32_t precision;              /* sbo_precision                            */
  int32_t kernel;                 /* sbo_kernel_choice                        */
  int32_t device;                 /* CUDA device ordinal                      */
  int32_t max_obs;                /* capacity in observations (n)             */
  int32_t reserved0;
  int64_t max_cells;              /* capacity in LOCAL cells (this shard)     */
} sbo_config;

typedef struct sbo_ctx sbo_ctx;

typedef struct {
  int64_t idx;    /* global cell index, -1 when the safe set is empty */
  double value;
} sbo_argmax;

/* Result of one BO step (caller-allocated POD). */
typedef struct {
  int64_t n_cells;            /* cells of this shard                                         */
  int64_t safe_count;         /* S_.count() (:426, :873); global when a communicator is set   */
  sbo_argmax argmax_ucb;      /* max of mu + beta*sigma over safe cells; ties -> lowest index */
  sbo_argmax argmax_width;    /* max of 2*beta*sigma (Q1 - Q0, :516) over safe cells          */
  int64_t clamped_sigma2;     /* cells whose sigma^2 came out < 0 and was clamped to 0        */
  int32_t n_obs;
  int32_t kernel_used;        /* sbo_kernel_choice actually run                               */
  float ms_posterior;         /* GPU time: posterior kernel(s)                                */
  float ms_sets;              /* GPU time: CI + mask + argmax kernel                          */
  float ms_comm;              /* GPU time: collective                                         */
  float ms_total;             /* GPU time: whole step on the stream                           */
  /* ---- ABI 2 ---- */
  int64_t low_sigma2;         /* cells with sigma^2 < 0.02 sigma_f^2 (where fp32 mode cannot hold 1e-5 of sigma^2 itself) */
  int32_t refreshed;          /* sbo_step_incremental: 1 = the refresh policy ran a full posterior instead          */
  float drift;                /* sbo_step_incremental: max |sigma^2 - exact| / sigma_f^2 and |mu - exact| / max(1,|mu|)
                                 over the re-evaluated sample rows of this step's drift check; -1 = no check this step */
} sbo_step_out;

/* ---- life cycle --------------------------------------------------------- */

int32_t sbo_abi_version(void);
/* replaces: OptimizerNode ctor parameter block (:48-72) + the service client
 * of the external GP node (:75-77). */
int32_t sbo_create(const sbo_config* cfg, sbo_ctx** out);
int32_t sbo_destroy(sbo_ctx* ctx);
const char* sbo_last_error(const sbo_ctx* ctx);   /* ctx may be NULL: last create error */
const char* sbo_status_string(int32_t status);

/* Use an existing CUDA stream (cudaStream_t as void*) instead of the
 * context's own; NULL restores the own stream. */
int32_t sbo_set_stream(sbo_ctx* ctx, void* cuda_stream);
int32_t sbo_synchronize(sbo_ctx* ctx);

/* Tuning / A-B switches.  The defaults are the product path; nothing here is read from the environment, so a
 * shipped libsbo.so computes the same thing whatever the caller's environment holds. */
typedef enum {
  SBO_OPT_TC_CTA_GROUP = 1,      /* 2 (default): CTA pairs, tcgen05 cta_group::2; 1: single CTAs                         */
  SBO_OPT_TC_SKIP = 2,           /* 1 (default): block-sparse walk of K* where the planner finds enough to skip (SBO_FP64:
                                    always, with its own cut-off: what is left out moves sigma^2 by < 1e-12 sigma_f^2); 0: dense */
  SBO_OPT_TC_PAIR = 3,           /* 1 (default): K*-stationary pair kernel for the full refresh; 0: row-block kernel      */
  SBO_OPT_TC_EAGER = 4,          /* 1 (default): first row-block pairs start inside the factorisation                    */
  SBO_OPT_TC_INC = 5,            /* 1 (default): incremental steps on large grids as a tensor-core GEMM                  */
  SBO_OPT_THIN = 6,              /* 1 (default): <= 8 new rows through the O(m n k) FMA kernel                            */
  SBO_OPT_THIN_2D = 7,           /* 1 (default): fp64 one-new-row step as a 128 x 128-cell tile on the fp64 tensor path (DMMA);
                                    2: the same tile as DFMA register tiles; 0: the column-strip kernel                     */
  SBO_OPT_TC_GRADED = 8,         /* 1 (default): distance-graded split depth -- (tile, chunk) visits whose K* entries are
                                    all < 2^-(b-10) sigma_f^2 issue hi.hi only (same error bound as the skipped chunks)   */
  SBO_OPT_TC_STRIP = 9,          /* tile pairs per strip of the strip-major tile order (default 4; 0 = raster order)       */
  SBO_OPT_REFRESH_EVERY = 10,    /* sbo_step_incremental runs a full refresh after this many incremental steps
                                    (default 256 in SBO_FP32, 0 = never in SBO_FP64)                                        */
  SBO_OPT_DRIFT_CHECK_EVERY = 11,/* every this many incremental steps, re-evaluate SBO_OPT_DRIFT_ROWS sample blocks exactly
                                    and refresh when the deviation exceeds tol / 4 (default: 16 in SBO_FP32, 256 in
                                    SBO_FP64, where a check costs several incremental steps; 0 = never)                     */
  SBO_OPT_DRIFT_ROWS = 12,       /* sample blocks per check, spread over the shard (default 2; a block is one grid row,
                                    eight consecutive rows in SBO_FP64: the DMMA kernel's tile is a 16 x 8 patch of cells)  */
  SBO_OPT_DIST_FIT = 13,         /* the full fit of a context that has a communicator (sbo_comm_init) or belongs to a group
                                    is SHARED between the ranks: block columns of the Cholesky and of L^-1 dealt in snake
                                    order, one panel broadcast per step, one all-gather of L^-1 at the end -- bit-identical
                                    to the replicated fit.  2 (default): from 8192 observations on; 1: always; 0: never.
                                    Every rank must hold the same value and call sbo_set_observations with the same data  */
  SBO_OPT_TC_PHASES = 14,        /* full refresh of the block-sparse walk as this many launches over consecutive row-block
                                    pair ranges (running row sums carried between them, bit-identical): keeps the rows of
                                    L^-1 in flight inside the L2.  0 (default): by size (1 below 8192 observations)         */
  SBO_OPT_MASK_DELTA = 15        /* 0 (default): every step copies its whole mask to `mask_host`.  1: a step given the SAME
                                    `mask_host` as the step before (same grid / band) copies only the 64 KiB segments of the
                                    mask in which a byte changed since that copy -- the CI / mask pass compares against the
                                    previous mask and raises one flag per segment.  The caller must leave the buffer as the
                                    library wrote it between the two calls (the reference rebuilds S_ and its image every
                                    step, :409, :444-456; this is its steady state without the 1 byte/cell transfer: after one
                                    new observation the mask changes near that observation only).  The first step, a step
                                    with another buffer, or one after sbo_set_grid / sbo_compute_sets copies everything.
                                    Steps without a host buffer (sbo_step_device) keep accumulating flags.  Costs one byte
                                    read per cell in the pass and one extra stream synchronisation per step.              */
} sbo_option;
int32_t sbo_set_option(sbo_ctx* ctx, int32_t option, int32_t value);
int32_t sbo_get_option(const sbo_ctx* ctx, int32_t option, int32_t* value);

/* ---- grid --------------------------------------------------------------- */

/* replaces: request->resolution (:580-581) and n_width_cells/n_height_cells/
 * x_coords/y_coords of the response (:612-613, :639-641).  [row0,row1) is the
 * band of grid rows this context owns (0,H for a single GPU); cells are
 * numbered globally. */
int32_t sbo_set_grid(sbo_ctx* ctx, double x0, double y0, double rx, double ry,
                     int32_t W, int32_t H, int32_t row0, int32_t row1);

/* ---- observations: the external GP node's fit ---------------------------- */

/* Full factorisation: K = k(X,X) + noise_var I = L L^T (blocked fp64
 * Cholesky), L^-1 (blocked triangular inverse), alpha = K^-1 (y - mean).
 * xy is n x 2 row-major.  replaces: mapping_package/terrain_mapping_node's
 * fit (launch/safe_bayesian_optimization.launch.py:111-117). */
int32_t sbo_set_observations(sbo_ctx* ctx, const double* xy, const double* y, int32_t n);
/* Rank-k append: L21 = K21 L11^-T, L22 = chol(K22 - L21 L21^T), new rows of
 * L^-1 = [-L22^-1 L21 L11^-1, L22^-1], alpha refreshed.  O(k n^2). */
int32_t sbo_append_observations(sbo_ctx* ctx, const double* xy, const double* y, int32_t k);
int32_t sbo_num_observations(const sbo_ctx* ctx);
/* Device time of the last full fit (ms, events on the context's stream) and how many ranks shared it (0: this
 * context factorised alone; see SBO_OPT_DIST_FIT).  replaces: nothing (the reference only logs the service
 * round trip, :589, :598-604). */
int32_t sbo_fit_info(const sbo_ctx* ctx, float* ms, int32_t* shared_ranks);
/* Internal ordering.  sbo_set_observations stores the observations sorted spatially (bands of 2
 * length-scales in y, by x inside a band): the GP is invariant under the permutation, the sorted
 * order makes the cross-covariance tiles of a cell tile block-sparse and keeps the tensor-core
 * accumulation short.  order[pos] = index in the caller's arrays of the observation stored at
 * `pos`; appended observations keep their arrival order behind the sorted ones. */
int32_t sbo_copy_order(sbo_ctx* ctx, int32_t* host_dst);
/* Read back the factor state for checking: alpha (n, in the CALLER's order) and L^-1 (n x n
 * row-major fp64, of the covariance in the INTERNAL order, see sbo_copy_order). */
int32_t sbo_copy_linv(sbo_ctx* ctx, double* host_dst);
int32_t sbo_copy_alpha(sbo_ctx* ctx, double* host_dst);

/* ---- the BO step --------------------------------------------------------- */

/* One BO step over this context's cells: GP posterior (mu, sigma^2), CI,
 * safe mask, argmax.  `mask_host` (n_cells bytes, 0/255, may be NULL)
 * receives the mask ready to wrap as cv::Mat(H, W, CV_8UC1, ptr).
 * replaces: request_terrain_map (:568-623) + the unpack loop (:639-644) +
 * ComputeSets (:399-416) + the rasterise loop of FindSafetyContourIndices
 * (:444-456). */
int32_t sbo_step(sbo_ctx* ctx, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out);
/* Mask bytes the last sbo_step / sbo_step_incremental of this context copied to the host, and in how many contiguous
 * runs (1 run of n_cells bytes unless SBO_OPT_MASK_DELTA is on; 0 / 0 when nothing changed or mask_host was NULL).
 * replaces: nothing (bookkeeping for the changed-segment copy). */
int32_t sbo_mask_copy_info(const sbo_ctx* ctx, int64_t* bytes, int32_t* runs);

/* Same step after a rank-k append, in O(m n k): only the k newest rows of
 * L^-1 are contracted (sigma^2 -= |rows_new k*|^2) and mu is recomputed with
 * the refreshed alpha.  Requires a previous sbo_step/sbo_step_incremental on
 * the same grid with the observations that preceded the append. */
int32_t sbo_step_incremental(sbo_ctx* ctx, double beta, double f_min, uint8_t* mask_host,
                             sbo_step_out* out);

/* ComputeSets on caller-supplied posterior arrays (fp64 host, length m): the
 * reference's own arithmetic and nothing else.  q0/q1 (may be NULL) receive
 * Q_.col(0)/Q_.col(1).  Bit-exact with the reference.
 * replaces: ComputeConfidenceIntervals (:411-416) + UpdateSafeSet (:409). */
int32_t sbo_compute_sets(sbo_ctx* ctx, const double* mu, const double* std_dev, int64_t m,
                         double beta, double f_min, uint8_t* mask_host, double* q0, double* q1,
                         sbo_step_out* out);

/* GetNextSubgoal's selection on a host-supplied frontier list (global cell
 * indices, as produced by cv::findContours on the mask, :461-492), using the
 * posterior of the last step (or the arrays of the last sbo_compute_sets):
 * distance to goal, sort by (distance, position), closest max(1, F/4),
 * strictly-largest CI width.  *out_cell = -1 when F == 0.
 * replaces: :508-549. */
int32_t sbo_select_subgoal(sbo_ctx* ctx, const int32_t* frontier, int64_t F, double goal_x,
                           double goal_y, int32_t* out_cell);

/* ---- the stages right after the mask -------------------- */

/* The pixel <-> cell maps of FindSafetyContourIndices on this grid, with the
 * reference's own arithmetic: bounds truncated to int (:431-434), pixel =
 * int((coord - min) / (max - min) * dim) (:450-451), out-of-image cells
 * dropped (:453), last writer of a pixel wins (:468-475).  col_of_px[W] /
 * row_of_py[H]: cell column / row stored at that pixel column / row, or -1.
 * These W + H integers replace the nested unordered_map over all m cells.
 * Error (INVALID_ARG) when an axis extent truncates to zero: the reference
 * divides by zero there. */
int32_t sbo_frontier_pixel_maps(sbo_ctx* ctx, int32_t* col_of_px, int32_t* row_of_py);
/* The CV_8UC1 image the reference hands to cv::findContours (:444-456), built
 * on the GPU from the mask of the last sbo_step by a gather through the maps
 * above; img_host is H*W bytes (0/255), the maps are returned as well.  A
 * contour point (x, y) maps back to cell row_of_py[y] * W + col_of_px[x].
 * Single-shard contexts only.  replaces: :431-456 and :468-475. */
int32_t sbo_frontier_image(sbo_ctx* ctx, uint8_t* img_host, int32_t* col_of_px, int32_t* row_of_py);

typedef enum {
  SBO_CELLS_SAFE = 0,     /* every safe cell (the gather in front of the alpha shape, :944-949)          */
  SBO_CELLS_BORDER = 1    /* safe cells with an unsafe cell or the grid frame in their 8-neighbourhood   */
} sbo_cell_set;
/* Global indices of the selected cells of this shard in ascending order
 * (stream compaction on the GPU).  *count = how many exist; at most `cap` are
 * written to idx_host (may be NULL when cap == 0: count only). */
int32_t sbo_compact_cells(sbo_ctx* ctx, int32_t mode, int32_t* idx_host, int64_t cap, int64_t* count);

/* Bulk read-back of the last step's posterior for this shard.  fp32 or fp64
 * according to cfg.precision (sbo_copy_*_f64 always converts to double). */
int32_t sbo_copy_mu(sbo_ctx* ctx, void* host_dst);
int32_t sbo_copy_sigma2(sbo_ctx* ctx, void* host_dst);
int32_t sbo_copy_mu_f64(sbo_ctx* ctx, double* host_dst);
int32_t sbo_copy_sigma2_f64(sbo_ctx* ctx, double* host_dst);
int32_t sbo_copy_mask(sbo_ctx* ctx, uint8_t* host_dst);

/* ---- N3: from the chosen cell to the published subgoal (host arithmetic) --- */

/* bg::within(goal_point, eroded_concave_polygon_) of reference src/safe_bayesian_optimization_node.cpp:657
 * (Boost.Geometry winding strategy: boundary points are not within; inside or on a hole is not within).
 * rings_xy = the rings' vertices back to back as (x, y) pairs, ring_sizes[0] = outer ring, the rest holes;
 * either orientation; an open ring gets its closing segment.  n_rings = 0: the empty polygon. */
int32_t sbo_point_within_polygon(const double* rings_xy, const int32_t* ring_sizes, int32_t n_rings,
                                 double px, double py, int32_t* within);
/* polydist(polygon, point) of reference src/libraries/polygeom_lib.cpp:400-474 as the node calls it (:677-691:
 * outer ring copied, bg::correct to closed + counter-clockwise).  As written there it returns the nearest
 * VERTEX (its rolled vertex list equals the list, :437), first minimum in corrected order; this restates the
 * same arithmetic.  nv = 0: dist 1e8, point (0, 0) (:419-423). */
int32_t sbo_polydist(const double* ring_xy, int32_t nv, double px, double py, double* proj_xy, double* dist);

typedef enum {
  SBO_SUBGOAL_NONE = 0,       /* goal outside the polygon and no frontier: "No valid subgoal found" (:700-703) */
  SBO_SUBGOAL_GOAL = 1,       /* goal within the polygon: the goal itself (:657-666) */
  SBO_SUBGOAL_PROJECTED = 2   /* frontier cell of GetNextSubgoal, projected with polydist (:667-699) */
} sbo_subgoal_kind;
/* Reference :650-706 in one call.  cell (may be NULL) = the frontier cell used, -1 otherwise; dist (may be NULL)
 * = polydist's distance. */
int32_t sbo_subgoal_point(sbo_ctx* ctx, const int32_t* frontier, int64_t F, double goal_x, double goal_y,
                          const double* rings_xy, const int32_t* ring_sizes, int32_t n_rings,
                          double* subgoal_xy, int32_t* cell, int32_t* kind, double* dist);

/* ---- N4: observation ingestion --------------------------------------------- */

/* One SpatialMeasurement as the publisher fills it (reference src/turtlesim_spatial_publisher.py:137-145):
 * position, value, uncertainty, stamp (unit / source_name strings are not needed by the GP). */
typedef struct {
  double x, y, z;
  double value;
  double uncertainty;        /* carried, not used: the noise level is a hyper-parameter (config/lpsc.yaml:37) */
  int32_t stamp_sec;
  uint32_t stamp_nanosec;
} sbo_measurement;
/* Feed `count` measurements: rank-k append, or a full re-sorted refit when there are no observations yet or the
 * appended tail would exceed max(64, n_at_last_fit / 4).  spatial_data_size (may be NULL) = the observation count
 * afterwards -- the number whose change triggers a step in the node (:552-566); refitted (may be NULL) = 1 when
 * the full fit ran.  On error nothing is ingested. */
int32_t sbo_ingest_measurements(sbo_ctx* ctx, const sbo_measurement* recs, int32_t count,
                                int32_t* spatial_data_size, int32_t* refitted);

/* ---- device-resident variants (bench / pipelines that keep data in HBM) -- */

/* Step with no host traffic: results stay on the device; `out` is filled
 * after a stream synchronise.  mask_dev may be NULL (internal buffer). */
int32_t sbo_step_device(sbo_ctx* ctx, double beta, double f_min, sbo_step_out* out);
/* CI+mask+argmax on device-resident fp64 mu/std (length m). */
int32_t sbo_compute_sets_device(sbo_ctx* ctx, const double* mu_dev, const double* std_dev,
                                int64_t m, double beta, double f_min, uint8_t* mask_dev,
                                sbo_step_out* out);

/* ---- allocation-free variant ------------------------------------------------
 *
 * sbo_compute_sets_noalloc has NO COUNTERPART IN THE REFERENCE: it was added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * It is sbo_compute_sets_device without a context: the same kernel (ComputeConfidenceIntervals + UpdateSafeSet + argmax
 * on device-resident fp64 mu / std, bit-identical results), launched on the CALLER's stream on the CURRENT device, with
 * the grid-reduction scratch in a fixed __device__ array inside libsbo.so (112.25 KiB per GPU that loads the library,
 * held until the process exits).  It allocates nothing -- no cudaMalloc / cudaFree, no allocating library routine -- and
 * does not synchronise: mask_dev (m bytes, 0/255) and *result_dev (a device pointer the caller owns) are complete when
 * `cuda_stream` reaches this point.  Limits and rules:
 *   - m <= 134 217 728 cells (2^27); beyond that SBO_ERR_CAPACITY (text via sbo_last_error(NULL));
 *   - every call on a device shares the one scratch array, which keeps the last call's contents (the call clears what
 *     must start at zero): calls that OVERLAP on the GPU -- two streams, two host threads -- race; serialise them;
 *   - the first call on each device looks the array's address up (cudaGetSymbolAddress): make it outside a capture. */
typedef struct {
  double ucb_value;     int64_t ucb_idx;      /* max of mu + beta std over safe cells (idx -1: no safe cell), ties -> lowest index */
  double width_value;   int64_t width_idx;    /* max of Q1 - Q0 over safe cells */
  int64_t safe_count;
  int64_t clamped;                            /* always 0 here (std inputs) */
  int64_t low_sigma2;                         /* always 0 here */
} sbo_sets_record;
int32_t sbo_compute_sets_noalloc(const double* mu_dev, const double* std_dev, int64_t m, double beta, double f_min,
                                 uint8_t* mask_dev, sbo_sets_record* result_dev, void* cuda_stream);

/* ---- multi-GPU ------------------------------------------------------------ */

/* One process per GPU.  Rank 0 calls sbo_comm_unique_id, the host transports
 * the 128 bytes (e.g. torch.distributed broadcast), every rank calls
 * sbo_comm_init.  Afterwards sbo_step* reduce argmax/safe_count over all ranks
 * with one NCCL all-gather of the per-rank (value, index) records. */
int32_t sbo_comm_unique_id(uint8_t id_out[128]);
int32_t sbo_comm_init(sbo_ctx* ctx, const uint8_t id[128], int32_t rank, int32_t world);
int32_t sbo_comm_destroy(sbo_ctx* ctx);
/* Bands of grid rows of equal estimated WORK for `parts` GPUs (SURVEY 8(e) G1 shards cells by contiguous row bands;
 * under the block-sparse walk of K* equal row counts are not equal work).  A pure host function of its arguments -- no
 * GPU, no context -- so every rank of a multi-process job derives the same cuts without a collective.  row_cuts
 * [parts + 1]: rank g owns rows [row_cuts[g], row_cuts[g + 1]).  max_rows > 0 caps the rows of one band (a shard's
 * max_cells / W).  cost_share (may be NULL) [parts]: estimated share of the work per band. */
int32_t sbo_plan_row_bands(const sbo_config* cfg, const double* xy, int32_t n, double x0, double y0, double rx, double ry,
                           int32_t W, int32_t H, int32_t parts, int32_t max_rows, int32_t* row_cuts, double* cost_share);
/* Host-side merge of per-shard results with the library's total order
 * (value desc, index asc); used when the host does its own transport. */
int32_t sbo_merge_step_out(const sbo_step_out* parts, int32_t count, sbo_step_out* merged);

/* ---- one context over several GPUs (single process, single caller thread) ------------------------------------------
 *
 * The node is one process with a SingleThreadedExecutor (:1500-1508); it cannot run one process per GPU.  A group owns
 * one sbo_ctx per device, cuts the grid into bands of rows of equal estimated work (sbo_plan_row_bands, re-planned when
 * the observations are replaced), drives the members from one host thread per device inside each call, writes every
 * band of the safe mask straight into the caller's ONE H x W image and folds the per-band records with the library's
 * total order.  cfg->device is ignored (devices[] decides; a device may be listed more than once); cfg->max_cells is
 * the capacity in cells of the WHOLE grid.  Semantics, status codes and "last good state" behaviour are those of the
 * single-context calls they mirror.  sbo_group_member gives the member contexts for everything else (options, plan,
 * read-back of a band). */
typedef struct sbo_group sbo_group;
int32_t sbo_group_create(const sbo_config* cfg, const int32_t* devices, int32_t n_devices, sbo_group** out);
int32_t sbo_group_destroy(sbo_group* g);
const char* sbo_group_last_error(const sbo_group* g);   /* g may be NULL: last create error */
int32_t sbo_group_size(const sbo_group* g);
sbo_ctx* sbo_group_member(sbo_group* g, int32_t i);
int32_t sbo_group_bands(const sbo_group* g, int32_t* row_cuts /* [size + 1] */);
int32_t sbo_group_set_option(sbo_group* g, int32_t option, int32_t value);
int32_t sbo_group_set_grid(sbo_group* g, double x0, double y0, double rx, double ry, int32_t W, int32_t H);
int32_t sbo_group_set_observations(sbo_group* g, const double* xy, const double* y, int32_t n);
int32_t sbo_group_append_observations(sbo_group* g, const double* xy, const double* y, int32_t k);
int32_t sbo_group_num_observations(const sbo_group* g);
/* mask_host: the whole H x W image (may be NULL); out: merged over the bands (timings = the slowest member). */
int32_t sbo_group_step(sbo_group* g, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out);
int32_t sbo_group_step_incremental(sbo_group* g, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out);
/* GetNextSubgoal's rule (:508-549) on a frontier that crosses the bands. */
int32_t sbo_group_select_subgoal(sbo_group* g, const int32_t* frontier, int64_t F, double goal_x, double goal_y,
                                 int32_t* out_cell);
int32_t sbo_group_copy_mu_f64(sbo_group* g, double* host_dst);       /* W * H values */
int32_t sbo_group_copy_sigma2_f64(sbo_group* g, double* host_dst);

/* ---- introspection -------------------------------------------------------- */

/* Number of kernel launches issued by this context since creation. */
int64_t sbo_launch_count(const sbo_ctx* ctx);
/* How the tcgen05 posterior kernel will run on the current grid and observations: CTAs per MMA group (2 = CTA
 * pair), whether it walks block-sparsity masks of K* (chunks of 64 observations farther than the kernel's reach
 * from a whole cell tile are left out: every skipped entry is < 2^-b sigma_f^2, b from the error budget, see sbo_tc_plan_detail), and the planner's estimate of the
 * share of (row block, chunk) visits that remain (1.0 when dense). */
int32_t sbo_tc_plan(sbo_ctx* ctx, int32_t* cta_group, int32_t* skipping, double* kept_fraction);
/* The same plan in full: [0] cut-off in bits (chunks below 2^-bits sigma_f^2 are skipped), [1] ring cut-off (visits below
 * 2^-ring issue hi.hi only; = [0] when graded split depth is off), [2] kept share of the (row block, chunk) visits,
 * [3] share of the dense walk's MMA passes that are issued (executed tensor flops / dense 3-product flops), [4] measured
 * max row 1-norm of L^-1 (0 = a-priori bound in use), [5] tiles per strip of the tile order (0 = raster). */
int32_t sbo_tc_plan_detail(sbo_ctx* ctx, double out[6]);
/* Tensor-core flops the last full tcgen05 refresh of this context EXECUTED (counted by the MMA issuers: every split
 * product, block-triangular padding included, skipped chunks excluded): the numerator of a pipe-utilisation figure, as
 * opposed to the credited m n (n + 1) of SURVEY M4.  In SBO_FP64 the same for the DMMA kernel (visited stages of 128 cells x
 * 128 rows x 32 observations).  0 when the last step ran another kernel. */
int32_t sbo_tc_executed_flops(sbo_ctx* ctx, double* flops);
/* Device properties the library sees: SM count, and 1 if tcgen05 path usable. */
int32_t sbo_device_info(const sbo_ctx* ctx, int32_t* sm_count, int32_t* tcgen05_ok);

#ifdef __cplusplus
}
#endif
#endif /* SBO_H */
