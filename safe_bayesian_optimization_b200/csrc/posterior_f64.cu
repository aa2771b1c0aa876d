// posterior_f64.cu -- GP posterior over the cells of this shard in fp64 mode (BASELINE config C5, and C1):
//   mu(x*)      = prior_mean + k*^T alpha
//   sigma^2(x*) = sigma_f^2 - | L^-1 k* |^2
// tcgen05 has no fp64 kind; the fp64 tensor path of sm_100a is mma.sync m8n8k4 (DMMA), which runs at the same
// 37 TFLOP/s as DFMA on this part (scripts/probe/fp64_peak.cu) but needs one instruction per 256 FMAs instead
// of eight, so the issue slots, the register file and the fp64 pipe stay free for the fragment loads and the
// K* generation.  (The register-tile DFMA kernel in posterior_simt.cu reaches 14.5 TFLOP/s credited; it stays as
// the cross-check, kernel = SBO_KERNEL_SIMT.)
//
// Tile = a patch of 16 columns x 8 rows of cells (128 cells) x 128 rows j of L^-1, 32 observations per stage,
// 512 threads = 16 warps as 4 (cells) x 4 (rows j), warp tile 32 x 32 = 4 x 4 DMMA tiles, 32 fp64 accumulators
// per thread.  The patch shape is what makes K* cheap: the RBF kernel on a grid is separable,
//   k*(cell(ix, iy), x_k) = ex[ix][k] * (sigma_f^2 ey[iy][k]),
// so a stage needs 16 x 32 + 8 x 32 exponentials for its 4096 entries instead of 4096.  Software pipeline, one
// __syncthreads per stage: while the DMMAs of stage s run, the L^-1 slab of stage s+1 arrives by cp.async (rows of
// L^-1 are K-contiguous: a straight copy, zero-filled above the diagonal and outside [j_lo, j_hi)), the K* slab of
// stage s+1 is formed from the factor tables of stage s+1, and the factor tables of stage s+2 are computed.
// Shared-memory rows are padded to 36 doubles: the (row = lane / 4, k = lane % 4) fragment reads of a half-warp
// then cover all 32 banks once.  Row block jb walks k < 128 (jb + 1) only (L^-1 is lower triangular); when it is
// finished the accumulators are squared and row-summed in registers, V itself never leaves them.
#include <algorithm>

#include "common.cuh"

namespace sbo {
namespace {

constexpr int FK_ = 32;
constexpr int FT = 512;            // threads: 16 warps as 4 (cells) x 4 (rows j), warp tile 32 x 32
constexpr int MI = 4, NI = 4;      // DMMA tiles per warp tile
constexpr int PC = 16, PR = 8;     // patch of cells: columns x rows
constexpr int FM = PC * PR;        // 128 cells per tile
constexpr int NT = FK_ * (PC + PR + 1);   // factor-table entries of a stage: x factors, y factors, alpha
constexpr int TPT = (NT + FT - 1) / FT;   // ... per thread
constexpr int FN = 128;            // rows of L^-1 per row block
constexpr int FK = FK_;            // observations per stage
constexpr int LD = FK + 4;          // padded stage row (doubles): 2 LD = 8 (mod 32) banks

struct F64Params {
  const double* xy;     // [n][2]
  const double* Linv;   // [n_cap][ld]
  const double* alpha;  // [n]
  int n, ld;
  int j_lo, j_hi;       // rows of L^-1 to contract
  int accumulate;       // 1: sigma^2 -= |rows|^2 on top of the stored value, mu refreshed
  double x0, y0, rx, ry;
  int W, rows_local, row0;
  double sf2, inv2l2, prior_mean;
  const double4* bbox;            // [ceil(n / 64)] (xmin, xmax, ymin, ymax) of the 64-observation chunks; null: dense walk
  double skip_r2;                 // chunks farther than this (squared distance to the patch of cells) are left out
  unsigned long long* exec_stages;   // executed (tile, row block, 32-observation stage) visits of the launch
};

struct F64Smem {
  double A[2][FM * LD];      // K* stage, A[cell][k]
  double B[2][FN * LD];      // L^-1 stage, B[j][k]
  double ex[3][FK][PC];      // x factors of a stage
  double ey[3][FK][PR];      // sigma_f^2 * y factors; 0 beyond n and below the last grid row
  double al[3][FK];          // alpha; 0 beyond n
  double red[4][FM];         // row sums of the four warp columns
  double mu_part[4][FM];
  uint32_t cmask[FT / 32];   // chunks (64 observations) within reach of this tile's patch: bit c of word c / 32
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src, int src_bytes) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1])
               : "d"(a), "d"(b));
}

// Bounding box of every 64-observation chunk of the (spatially ordered) observations, in the caller's coordinates.
__global__ void chunk_bbox_f64_kernel(const double* __restrict__ xy, int n, int nchunks, double4* __restrict__ bbox) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (c >= nchunks) return;
  double xmin = INFINITY, xmax = -INFINITY, ymin = INFINITY, ymax = -INFINITY;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int k = c * 64 + h * 32 + lane;
    if (k < n) {
      const double x = xy[2 * (size_t)k], y = xy[2 * (size_t)k + 1];
      xmin = fmin(xmin, x); xmax = fmax(xmax, x);
      ymin = fmin(ymin, y); ymax = fmax(ymax, y);
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    xmin = fmin(xmin, __shfl_xor_sync(0xffffffffu, xmin, d)); xmax = fmax(xmax, __shfl_xor_sync(0xffffffffu, xmax, d));
    ymin = fmin(ymin, __shfl_xor_sync(0xffffffffu, ymin, d)); ymax = fmax(ymax, __shfl_xor_sync(0xffffffffu, ymax, d));
  }
  if (lane == 0) bbox[c] = make_double4(xmin, xmax, ymin, ymax);
}

// (row block, first observation) of a stage; row block jb walks the kept stages k0 = 0, 32, .. < kd(jb)
struct Cursor {
  int jb, k0;
};

static_assert(FT / FM == 4 && FT / 32 == (FM / 32) * (FN / 32) && (FN * FK / 2) % FT == 0 && (2 * LD) % 32 == 8, "thread maps");

__global__ void __launch_bounds__(FT, 1)
posterior_f64_kernel(F64Params p, double* __restrict__ mu_out, double* __restrict__ s2_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  F64Smem& sm = *reinterpret_cast<F64Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp >> 2, wn = warp & 3;      // warp tile: cells [32 wm, +32) x rows j [32 wn, +32)
  const int g = lane >> 2, t = lane & 3;        // DMMA fragment coordinates
  const int tiles_x = (p.W + PC - 1) / PC, tiles_y = (p.rows_local + PR - 1) / PR;
  const long long ntiles = (long long)tiles_x * tiles_y;
  const int n_pad = (p.n + FK - 1) / FK * FK;
  const int jb_lo = p.j_lo / FN, jb_hi = (p.j_hi + FN - 1) / FN;   // row blocks [jb_lo, jb_hi)
  auto kd = [&](int jb) { return min(n_pad, (jb + 1) * FN); };
  // Block-sparse walk (the fp64 counterpart of posterior_tc.cu's chunk masks): a stage is visited only if its
  // 64-observation chunk has a point within skip_r2 of the tile's patch of cells, i.e. can hold a K* entry above the
  // cut-off (plan.cu: f64_skip_bits_of -- what is left out moves sigma^2 by < 1e-12 sigma_f^2).  Row blocks without a
  // kept stage are not visited at all.  All threads walk the same mask: uniform control flow.
  const bool masked = p.bbox != nullptr;
  auto kept = [&](int k0) { const int c = k0 >> 6; return !masked || ((sm.cmask[c >> 5] >> (c & 31)) & 1u) != 0u; };
  auto seek = [&](Cursor& c) {   // first kept stage at or after c (c.jb == jb_hi: none)
    while (c.jb < jb_hi) {
      const int end = kd(c.jb);
      while (c.k0 < end && !kept(c.k0)) c.k0 += FK;
      if (c.k0 < end) return;
      c.jb += 1; c.k0 = 0;
    }
  };
  auto advance = [&](Cursor& c) {
    c.k0 += FK;
    seek(c);
  };
  unsigned long long stages_done = 0;

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int col0 = (int)(tile % tiles_x) * PC, r0 = (int)(tile / tiles_x) * PR;
    if (masked) {   // chunk mask of this tile: thread c tests chunk c (FT = 512 chunks = 32 768 observations at most)
      bool on = tid < (n_pad + 63) / 64;
      if (on) {
        const double xa = cell_coord(p.x0, p.rx, col0), xb = cell_coord(p.x0, p.rx, min(col0 + PC, p.W) - 1);
        const double ya = cell_coord(p.y0, p.ry, p.row0 + r0), yb = cell_coord(p.y0, p.ry, p.row0 + min(r0 + PR, p.rows_local) - 1);
        const double4 b = p.bbox[tid];
        const double dx = fmax(fmax(b.x - fmax(xa, xb), fmin(xa, xb) - b.y), 0.0);
        const double dy = fmax(fmax(b.z - fmax(ya, yb), fmin(ya, yb) - b.w), 0.0);
        on = dx * dx + dy * dy <= p.skip_r2;
      }
      const uint32_t bits = __ballot_sync(0xffffffffu, on);
      if (lane == 0) sm.cmask[warp] = bits;
      __syncthreads();
    }

    // ---- per-stage producers -------------------------------------------------------------------------
    // factor tables of the stage that starts at observation k0: the global loads are issued by table_load (early
    // in an iteration), the exponentials and the stores come after the DMMAs (table_store)
    struct TabIn { double v[TPT]; };
    auto table_load = [&](int k0) -> TabIn {
      TabIn o;
#pragma unroll
      for (int u = 0; u < TPT; ++u) {
        const int e = tid + u * FT;
        o.v[u] = 0.0;
        if (e < FK * PC) {
          const int k = k0 + (e >> 4);
          if (k < p.n) o.v[u] = p.xy[2 * (size_t)k];
        } else if (e < FK * PC + FK * PR) {
          const int k = k0 + (e - FK * PC) / PR;
          if (k < p.n) o.v[u] = p.xy[2 * (size_t)k + 1];
        } else if (e < NT) {
          const int k = k0 + (e - FK * PC - FK * PR);
          if (k < p.n) o.v[u] = p.alpha[k];
        }
      }
      return o;
    };
    auto table_store = [&](int k0, int tb, const TabIn& o) {
#pragma unroll
      for (int u = 0; u < TPT; ++u) {
        const int e = tid + u * FT;
        if (e < FK * PC) {
          const int kk = e >> 4, pc = e & 15;
          const double dx = cell_coord(p.x0, p.rx, col0 + pc) - o.v[u];
          sm.ex[tb][kk][pc] = k0 + kk < p.n ? exp(-dx * dx * p.inv2l2) : 0.0;
        } else if (e < FK * PC + FK * PR) {
          const int kk = (e - FK * PC) / PR, pr = (e - FK * PC) % PR;
          const double dy = cell_coord(p.y0, p.ry, p.row0 + r0 + pr) - o.v[u];
          sm.ey[tb][kk][pr] = (k0 + kk < p.n && r0 + pr < p.rows_local) ? p.sf2 * exp(-dy * dy * p.inv2l2) : 0.0;
        } else if (e < NT) {
          sm.al[tb][e - FK * PC - FK * PR] = o.v[u];
        }
      }
    };
    double mu_acc = 0.0;
    constexpr int KPT = FK * FM / FT;   // K* entries per thread and stage
    auto make_a = [&](int tb, int buf, bool with_mu) {   // K* stage from its tables: thread = (cell, KPT consecutive k)
      const int c = tid & (FM - 1), h = tid / FM, pc = c & 15, pr = c >> 4;   // a warp: 32 cells, one k group
      double v[KPT];
#pragma unroll
      for (int i = 0; i < KPT; ++i) v[i] = sm.ex[tb][h * KPT + i][pc] * sm.ey[tb][h * KPT + i][pr];
      if (with_mu) {
#pragma unroll
        for (int i = 0; i < KPT; ++i) mu_acc = fma(v[i], sm.al[tb][h * KPT + i], mu_acc);
      }
      double2* dst = reinterpret_cast<double2*>(&sm.A[buf][c * LD + h * KPT]);
#pragma unroll
      for (int i = 0; i < KPT / 2; ++i) dst[i] = make_double2(v[2 * i], v[2 * i + 1]);
    };
    auto load_b = [&](Cursor cur, int buf) {   // 128 rows x 16 k of L^-1: 1024 granules of two doubles
#pragma unroll
      for (int q = 0; q < FN * FK / 2 / FT; ++q) {
        const int gi = q * FT + tid, r = gi / (FK / 2), kq = (gi % (FK / 2)) * 2;
        const int j = cur.jb * FN + r, k = cur.k0 + kq;
        const bool live = j >= p.j_lo && j < p.j_hi && k <= j;
        const double* src = p.Linv + (size_t)(live ? j : 0) * p.ld + (live ? k : 0);
        cp_async16(&sm.B[buf][r * LD + kq], src, live ? 16 : 0);
      }
    };

    double rs[MI];
#pragma unroll
    for (int i = 0; i < MI; ++i) rs[i] = 0.0;

    if (jb_hi > jb_lo) {
      double acc[MI][NI][2];
#pragma unroll
      for (int mi = 0; mi < MI; ++mi)
#pragma unroll
        for (int ni = 0; ni < NI; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

      // ---- prologue: tables of stages 0 and 1, table inputs of stage 2, K* and L^-1 of stage 0 ----
      Cursor cur{jb_lo, 0};                       // stage being contracted
      seek(cur);
      Cursor nxt = cur; if (cur.jb < jb_hi) advance(nxt);   // its successors (jb == jb_hi: beyond the end)
      Cursor nx2 = nxt; if (nxt.jb < jb_hi) advance(nx2);
      Cursor nx3 = nx2; if (nx2.jb < jb_hi) advance(nx3);
      TabIn tab_cur = table_load(nx2.jb < jb_hi ? nx2.k0 : 0);   // inputs of the tables stored during iteration s = 0
      if (cur.jb < jb_hi) {                       // (a tile out of reach of every observation walks nothing)
        table_store(cur.k0, 0, table_load(cur.k0));
        if (nxt.jb < jb_hi) table_store(nxt.k0, 1, table_load(nxt.k0));
        load_b(cur, 0);
        __syncthreads();
        make_a(0, 0, cur.jb == jb_hi - 1);
        cp_async_wait_all();
        __syncthreads();
      }

      int s = 0;   // stage counter: A/B buffer s & 1, table buffer s % 3
      while (cur.jb < jb_hi) {
        const bool have_next = nxt.jb < jb_hi, have_nx2 = nx2.jb < jb_hi, have_nx3 = nx3.jb < jb_hi;
        // ---- contract stage s: FK / 4 k-steps of 4 x 4 DMMA tiles.  The staging work for the following stages
        // (L^-1 slab of s+1, K* slab of s+1, tables of s+2, table inputs of s+3) sits BETWEEN the k-steps, at
        // positions that differ for the four warps of a scheduler (they differ in wm): while one of them stages,
        // the other three keep the fp64 pipe busy, and after the barrier every warp starts with DMMAs. ----
        const double* As = &sm.A[s & 1][(wm * 32 + g) * LD + t];
        const double* Bs = &sm.B[s & 1][(wn * 32 + g) * LD + t];
        TabIn tab_new = tab_cur;
#pragma unroll
        for (int k4 = 0; k4 < FK / 4; ++k4) {
          double a[MI], b[NI];
#pragma unroll
          for (int mi = 0; mi < MI; ++mi) a[mi] = As[mi * 8 * LD + k4 * 4];
#pragma unroll
          for (int ni = 0; ni < NI; ++ni) b[ni] = Bs[ni * 8 * LD + k4 * 4];
#pragma unroll
          for (int mi = 0; mi < MI; ++mi)
#pragma unroll
            for (int ni = 0; ni < NI; ++ni) dmma(acc[mi][ni], a[mi], b[ni]);
          const int slot = (k4 - wm * (FK / 16)) & (FK / 4 - 1);
          if (slot == 0 && have_next) {
            load_b(nxt, (s + 1) & 1);
            if (have_nx3) tab_new = table_load(nx3.k0);
          }
          if (slot == FK / 16 && have_next) make_a((s + 1) % 3, (s + 1) & 1, nxt.jb == jb_hi - 1);
          if (slot == FK / 8 && have_nx2) table_store(nx2.k0, (s + 2) % 3, tab_cur);
        }
        tab_cur = tab_new;
        // ---- end of a row block: squared row sums of this 128 x 128 block of V ----
        ++stages_done;
        if (nxt.jb != cur.jb) {   // the last kept stage of this row block
#pragma unroll
          for (int mi = 0; mi < MI; ++mi) {
            double q = 0.0;
#pragma unroll
            for (int ni = 0; ni < NI; ++ni) {
              q = fma(acc[mi][ni][0], acc[mi][ni][0], q);
              q = fma(acc[mi][ni][1], acc[mi][ni][1], q);
              acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
            }
            rs[mi] += q;
          }
        }
        cp_async_wait_all();
        __syncthreads();
        cur = nxt; nxt = nx2; nx2 = nx3;
        if (nx3.jb < jb_hi) advance(nx3);
        ++s;
      }
    }

    // ---- finish the tile: reduce over the lanes and warps that share a cell, write mu / sigma^2 ----
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) {
      rs[mi] += __shfl_xor_sync(0xffffffffu, rs[mi], 1);
      rs[mi] += __shfl_xor_sync(0xffffffffu, rs[mi], 2);
      if (t == 0) sm.red[wn][wm * 32 + mi * 8 + g] = rs[mi];
    }
    sm.mu_part[tid / FM][tid & (FM - 1)] = mu_acc;
    __syncthreads();
    if (tid < FM) {
      const int pc = tid & 15, pr = tid >> 4;
      const int col = col0 + pc, lrow = r0 + pr;
      if (col < p.W && lrow < p.rows_local) {
        const long long lc = (long long)lrow * p.W + col;
        const double sq = (sm.red[0][tid] + sm.red[1][tid]) + (sm.red[2][tid] + sm.red[3][tid]);
        const double prev = p.accumulate ? s2_out[lc] : p.sf2;
        s2_out[lc] = prev - sq;
        mu_out[lc] = p.prior_mean + ((sm.mu_part[0][tid] + sm.mu_part[1][tid]) + (sm.mu_part[2][tid] + sm.mu_part[3][tid]));
      }
    }
    __syncthreads();   // red / mu_part / stage buffers / the chunk mask are reused by the next tile
  }
  if (p.exec_stages && tid == 0 && stages_done) atomicAdd(p.exec_stages, stages_done);
}

}  // namespace

// Block-sparse walk of the fp64 kernels: boxes of the 64-observation chunks (the observations are spatially ordered in
// every precision; appended ones sit behind in arrival order, their chunks simply have large boxes) and the squared reach
// beyond which K* < 2^-bits sigma_f^2 (bits: plan.cu f64_skip_bits_of, from the measured row norm of L^-1).  *boxes = null:
// dense walk (switched off, few observations, or more chunks than the caller's mask holds).
int32_t f64_chunk_boxes(sbo_ctx* ctx, int max_chunks, const void** boxes, double* skip_r2) {
  *boxes = nullptr;
  *skip_r2 = 0.0;
  const int nchunks = (ctx->n + 63) / 64;
  if (!ctx->tc_skip || ctx->n < 512 || nchunks > max_chunks) return SBO_OK;
  if (!ctx->d_chunk_bbox64) SBO_CUDA(ctx, cudaMalloc(&ctx->d_chunk_bbox64, sizeof(double4) * (size_t)(ctx->n_cap / 64 + 1)));
  chunk_bbox_f64_kernel<<<(nchunks * 32 + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_xy, ctx->n, nchunks,
                                                                               static_cast<double4*>(ctx->d_chunk_bbox64));
  SBO_LAUNCH_CHECK(ctx, "chunk_bbox_f64_kernel");
  *boxes = ctx->d_chunk_bbox64;
  const double bits = f64_skip_bits_of(ctx->cfg, ctx->n, ctx->linv_r1);
  *skip_r2 = (bits + 0.5) * 0.6931471805599453 * 2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale;
  return SBO_OK;
}

int32_t launch_posterior_f64(sbo_ctx* ctx, int j_lo, int j_hi, bool accumulate) {
  if (ctx->cfg.precision != SBO_FP64) return fail(ctx, SBO_ERR_STATE, "posterior_f64: context is not in fp64 mode");
  if (ctx->m_local == 0) return SBO_OK;
  if (accumulate && j_hi <= j_lo) return SBO_OK;   // no new rows: the stored posterior stands
  F64Params p;
  p.xy = ctx->d_xy; p.Linv = ctx->d_Linv; p.alpha = ctx->d_alpha;
  p.n = ctx->n; p.ld = ctx->n_cap;
  p.j_lo = j_hi > j_lo ? j_lo : 0; p.j_hi = j_hi > j_lo ? j_hi : 0;
  p.accumulate = accumulate ? 1 : 0;
  p.x0 = ctx->x0; p.y0 = ctx->y0; p.rx = ctx->rx; p.ry = ctx->ry;
  p.W = ctx->W; p.rows_local = ctx->row1 - ctx->row0; p.row0 = ctx->row0;
  p.sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  p.inv2l2 = 1.0 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale);
  p.prior_mean = ctx->cfg.prior_mean;
  if (p.ld % 2 != 0) return fail(ctx, SBO_ERR_STATE, "posterior_f64: odd leading dimension of L^-1");
  p.exec_stages = ctx->d_exec_cols;
  const void* boxes = nullptr;
  SBO_TRY(f64_chunk_boxes(ctx, FT, &boxes, &p.skip_r2));
  p.bbox = static_cast<const double4*>(boxes);
  if (ctx->d_exec_cols && !accumulate) SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_exec_cols, 0, sizeof(unsigned long long), ctx->stream));
  if (!ctx->attr_f64) {   // the attribute is per device: kept per context, not in a process-wide flag
    SBO_CUDA(ctx, cudaFuncSetAttribute(posterior_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(F64Smem)));
    ctx->attr_f64 = true;
  }
  const long long ntiles = (long long)((p.W + PC - 1) / PC) * ((p.rows_local + PR - 1) / PR);
  const unsigned grid = (unsigned)std::min<long long>(ntiles, ctx->sm_count);
  posterior_f64_kernel<<<grid, FT, sizeof(F64Smem), ctx->stream>>>(p, static_cast<double*>(ctx->d_mu), static_cast<double*>(ctx->d_s2));
  SBO_LAUNCH_CHECK(ctx, "posterior_f64_kernel");
  return SBO_OK;
}

}  // namespace sbo
