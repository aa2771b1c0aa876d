// posterior_thin.cu -- the incremental BO step after a small rank-k append (k <= 8), in O(m n k).
//
// After sbo_append_observations the factor gains k rows:  L^-1 = [[L11^-1, 0], [R, L22^-1]].  For every cell
//   sigma^2_new = sigma^2_old - sum_r u_r^2,   u_r = rows_new[r] . k*      (sequential conditioning)
//   mu_new      = mu_old + sum_r w_r u_r,      w = L^-1 (y - prior_mean)   (alpha_new = [alpha_old; 0] + rows_new^T w_new,
//                                                                           factor.cu append_alpha_kernel, so k* . alpha_new
//                                                                           = k* . alpha_old + sum_r w_r (rows_new[r] . k*))
// so the k projections u_r are all a cell needs: one pass over the n observations with k outputs (not 1 + k: the
// mean needs no contraction of its own).  This is what the ROS node's steady state looks
// like (one measurement per second, src/turtlesim_spatial_publisher.py:43) and what BASELINE configs C3/C5 time.
// The contraction has N = k (<= 8) columns: far too thin for a tensor-core tile (the 256-row block kernel of
// posterior_tc.cu spends 256/k times the necessary MMAs on it), and its cost is the K* entries themselves --
// one ex2 (fp32) or exp (fp64) per cell and observation if done naively.  The RBF kernel on a grid is separable,
//   k*(cell(ix, iy), obs j) = sigma_f^2 ex(ix, j) ey(iy, j),
// so a thread that owns C columns x R rows of cells needs ex only once per (column, observation) and the row
// factor arrives pre-multiplied:  the CTA stages, per observation, the tile-relative x and the R k products
// ey(r, j) Linv[new row q][j]; the inner loop is then one FMA per (cell, observation, new row)
// and 1/R exponentials per (cell, observation), with all staged values read as broadcast 16-byte vectors.
//
// Numerics: fp32 mode accumulates in fp32 over 64 observations and in fp64 across (an fp32 running sum over
// n = 16 384 terms would cost ~1e-5); coordinates are tile-relative fp32 (hi, lo) differences as in the
// tcgen05 kernel.  fp64 mode is plain fp64.
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "common.cuh"

namespace sbo {
namespace {

constexpr int kThinThreads = 256;
constexpr int kSub = 64;      // fp32 accumulation span

struct ThinParams {
  const float4* obs_f32;   // fp32 mode: scaled (x_hi, x_lo, y_hi, y_lo)
  const double* xy;        // fp64 mode
  const double* Linv;      // [n_cap][ld] fp64
  const double* w;         // [n] fp64: L^-1 (y - prior_mean); entries row_lo .. row_lo + k_new - 1 are read
  int n, ld;
  int row_lo, k_new;       // new rows [row_lo, row_lo + k_new)
  int W, rows_local, row0;
  double x0, y0, rx, ry;
  double xscale;           // fp32 mode: sqrt(log2(e) / 2 l^2)
  double sf2, inv2l2, prior_mean;
  const double4* bbox;     // fp64 2-D tile kernel: boxes of the 64-observation chunks (null: every chunk is staged)
  double skip_r2;          // ... and the squared reach beyond which a chunk is left out for a tile
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

// Tile = (kThinThreads * C) columns x R grid rows; thread = C columns x R rows.  NP products per observation:
// slot r KR + q is ey_r Linv[new row q].
template <typename T, int KR, int C, int R>
struct ThinCfg {
  static constexpr int kNP = R * KR;
  static constexpr int kChunk = sizeof(T) == 4 ? 256 : 128;   // observations staged per pass (<= 48 KB static smem)
  static_assert(kNP % 4 == 0, "products are read as 4-vectors");
};

template <typename T, int KR, int C, int R>
struct __align__(16) ThinSmem {
  T x[ThinCfg<T, KR, C, R>::kChunk];                               // tile-relative x (fp32 mode: scaled)
  T p[ThinCfg<T, KR, C, R>::kChunk][ThinCfg<T, KR, C, R>::kNP];    // zero beyond n and beyond the last grid row
};

// x factor of a K* entry from the tile-relative difference
__device__ __forceinline__ float xfactor(float dx, double) { return ex2f(-dx * dx); }
__device__ __forceinline__ double xfactor(double dx, double inv2l2) { return exp(-dx * dx * inv2l2); }

template <typename T, int KR, int C, int R>
__global__ void __launch_bounds__(kThinThreads)
posterior_thin_kernel(ThinParams p, T* __restrict__ mu_out, T* __restrict__ s2_out) {
  using Cfg = ThinCfg<T, KR, C, R>;
  constexpr int NP = Cfg::kNP, kChunk = Cfg::kChunk;
  constexpr bool kF32 = sizeof(T) == 4;
  __shared__ ThinSmem<T, KR, C, R> sm;
  const int tid = threadIdx.x;
  const int span = kThinThreads * C;
  const int tiles_per_row = (p.W + span - 1) / span;
  const int row_groups = (p.rows_local + R - 1) / R;
  const long long ntiles = (long long)row_groups * tiles_per_row;

  double wq[KR];
#pragma unroll
  for (int q = 0; q < KR; ++q) wq[q] = q < p.k_new ? p.w[p.row_lo + q] : 0.0;

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int r0 = (int)(tile / tiles_per_row) * R, col0 = (int)(tile % tiles_per_row) * span;
    // tile origin; fp32 mode works in scaled units with (hi, lo) pairs, fp64 mode in plain coordinates
    const double ox = cell_coord(p.x0, p.rx, col0);
    const double oxs = ox * p.xscale;
    const float oxh = (float)oxs, oxl = (float)(oxs - (double)oxh);
    T cx[C];
#pragma unroll
    for (int q = 0; q < C; ++q) {
      const double off = (double)(q * kThinThreads + tid) * p.rx;
      cx[q] = kF32 ? (T)(off * p.xscale) : (T)off;
    }
    double acc[C][NP];   // fp32 mode: sums across 64-observation spans; fp64 mode: the accumulators themselves
#pragma unroll
    for (int q = 0; q < C; ++q)
#pragma unroll
      for (int t = 0; t < NP; ++t) acc[q][t] = 0.0;

    for (int k0 = 0; k0 < p.n; k0 += kChunk) {
      // ---- stage observations k0 .. k0 + kChunk - 1 ----
      if (tid < kChunk) {
        const int j = k0 + tid;
        const bool live = j < p.n;
        T tx = (T)0, tl[KR];
        float4 of = make_float4(0.f, 0.f, 0.f, 0.f);
        double2 od = make_double2(0.0, 0.0);
#pragma unroll
        for (int q = 0; q < KR; ++q) tl[q] = (T)0;
        if (live) {
          if (kF32) {
            of = p.obs_f32[j];
            tx = (T)((of.x - oxh) + (of.y - oxl));
          } else {
            od = *reinterpret_cast<const double2*>(p.xy + 2 * (size_t)j);
            tx = (T)(od.x - ox);
          }
#pragma unroll
          for (int q = 0; q < KR; ++q)
            if (q < p.k_new) tl[q] = (T)p.Linv[(size_t)(p.row_lo + q) * p.ld + j];
        }
        sm.x[tid] = tx;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          T e = (T)0;
          if (live && r0 + r < p.rows_local) {
            const double oy = cell_coord(p.y0, p.ry, p.row0 + r0 + r);
            if (kF32) {
              const double oys = oy * p.xscale;
              const float oyh = (float)oys, oyl = (float)(oys - (double)oyh);
              const float yr = (of.z - oyh) + (of.w - oyl);
              e = (T)(ex2f(-yr * yr) * (float)p.sf2);
            } else {
              const double dy = od.y - oy;
              e = (T)(p.sf2 * exp(-dy * dy * p.inv2l2));
            }
          }
#pragma unroll
          for (int q = 0; q < KR; ++q) sm.p[tid][r * KR + q] = e * tl[q];
        }
      }
      __syncthreads();
      const int cnt = min(kChunk, p.n - k0);
      for (int s0 = 0; s0 < cnt; s0 += kSub) {
        T f[C][NP];
        if (kF32) {
#pragma unroll
          for (int q = 0; q < C; ++q)
#pragma unroll
            for (int t = 0; t < NP; ++t) f[q][t] = (T)0;
        } else {
#pragma unroll
          for (int q = 0; q < C; ++q)
#pragma unroll
            for (int t = 0; t < NP; ++t) f[q][t] = (T)acc[q][t];
        }
#pragma unroll 1
        for (int j4 = s0; j4 < s0 + kSub; j4 += 4) {  // the chunk is zero-padded: no bound check inside
          T xs[4];
          load4(&sm.x[j4], xs);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            T ex[C];
#pragma unroll
            for (int q = 0; q < C; ++q) ex[q] = xfactor(xs[i] - cx[q], p.inv2l2);
#pragma unroll
            for (int t0 = 0; t0 < NP; t0 += 4) {
              T pv[4];
              load4(&sm.p[j4 + i][t0], pv);
#pragma unroll
              for (int t = 0; t < 4; ++t)
#pragma unroll
                for (int q = 0; q < C; ++q) f[q][t0 + t] = fma(ex[q], pv[t], f[q][t0 + t]);
            }
          }
        }
#pragma unroll
        for (int q = 0; q < C; ++q)
#pragma unroll
          for (int t = 0; t < NP; ++t) acc[q][t] = kF32 ? acc[q][t] + (double)f[q][t] : (double)f[q][t];
      }
      __syncthreads();
    }

#pragma unroll
    for (int q = 0; q < C; ++q) {
      const int col = col0 + q * kThinThreads + tid;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        if (col < p.W && r0 + r < p.rows_local) {
          const long long lc = (long long)(r0 + r) * p.W + col;
          double red = 0.0, dmu = 0.0;
#pragma unroll
          for (int t = 0; t < KR; ++t) {
            const double u = acc[q][r * KR + t];
            red = fma(u, u, red);
            dmu = fma(wq[t], u, dmu);
          }
          s2_out[lc] = (T)((double)s2_out[lc] - red);
          mu_out[lc] = (T)((double)mu_out[lc] + dmu);
        }
      }
    }
  }
}

// ---- fp64, one new row, large grids: two-dimensional register tile ------------------------------------------------
// The kernel above reads R staged products per observation and thread as broadcast LDS.128 (two doubles each) for R
// DFMAs: 16 LDS.128 per 32 DFMA at R = 32, and a warp-wide broadcast of 16 bytes still costs the shared-memory pipe
// its four 8-lane phases -- it ran at 26 % of the DFMA rate, bound by those reads.  Here a thread owns 8 x 8 cells
// (columns 2 tx + 32 i + {0,1}, rows 2 ty + 32 i + {0,1} of a 128 x 128-cell CTA tile), the CTA stages BOTH factors
// per observation -- a[c] = ex(col c, j) and b[r] = sigma_f^2 ey(row r, j) Linv[new row][j] -- and the inner loop is
// 8 LDS.128 for 64 DFMA: u[r][c] += b[r] a[c].  One exp per 64 cell-observations instead of one per 32.
constexpr int kT2 = 128;      // cells per CTA tile edge
constexpr int kT2Ld = kT2 + 8; // row stride of the staged factors (doubles): 16-byte aligned rows, staggered banks
constexpr int kT2Chunk = 32;  // observations staged per pass: 2 x 32 x 136 doubles = 68 KB

// Staging by recurrence: along a grid line with pitch s h the RBF factor obeys
//   e(i+1) = e(i) g(i),  g(i) = exp(-(2 (X_i - x_j) s h + s^2 h^2) / 2 l^2),  g(i+1) = g(i) Q,  Q = exp(-s^2 h^2 / l^2),
// so a thread that owns observation jj and the 16 cells sg, sg + 8, ..., sg + 120 of the tile edge pays two exp and
// thirty multiplications instead of sixteen exp (the fp64 exp is ~45 DFMA-equivalents; with one exp per staged value the
// staging cost 70 % of the inner loop).  Sixteen steps: relative error ~16 ulp, far inside the fp64 tolerance.  The
// host only selects this kernel when a tile edge spans <= 4 length-scales: an anchor that underflows then sits >= 33
// length-scales from every cell of its segment, whose true factors are < e^-560.
__global__ void __launch_bounds__(kThinThreads, 1)
posterior_thin2d_f64_kernel(ThinParams p, double* __restrict__ mu_out, double* __restrict__ s2_out) {
  extern __shared__ __align__(16) double sm2[];
  double* a_s = sm2;                        // [kT2Chunk][kT2Ld]
  double* b_s = sm2 + kT2Chunk * kT2Ld;     // [kT2Chunk][kT2Ld]
  __shared__ uint32_t cmask[kThinThreads / 32];   // 256 chunks = 16 384 observations
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int sg = tid & 7, jj = tid >> 3;    // staging role: observation jj of the chunk, cells sg + 8 i of the tile edge
  const int tiles_x = (p.W + kT2 - 1) / kT2, tiles_y = (p.rows_local + kT2 - 1) / kT2;
  const long long ntiles = (long long)tiles_x * tiles_y;
  const double w = p.w[p.row_lo];
  const double* lrow = p.Linv + (size_t)p.row_lo * p.ld;
  const double sx = 8.0 * p.rx, sy = 8.0 * p.ry;
  const double qx = exp(-2.0 * sx * sx * p.inv2l2), qy = exp(-2.0 * sy * sy * p.inv2l2);

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int r0 = (int)(tile / tiles_x) * kT2, col0 = (int)(tile % tiles_x) * kT2;
    if (p.bbox) {   // chunks within reach of this 128 x 128 patch (same box test and cut-off as posterior_f64.cu)
      __syncthreads();   // the previous tile's reads of the mask are over
      bool on = tid < (p.n + 63) / 64;
      if (on) {
        const double xa = cell_coord(p.x0, p.rx, col0), xb = cell_coord(p.x0, p.rx, min(col0 + kT2, p.W) - 1);
        const double ya = cell_coord(p.y0, p.ry, p.row0 + r0), yb = cell_coord(p.y0, p.ry, p.row0 + min(r0 + kT2, p.rows_local) - 1);
        const double4 b = p.bbox[tid];
        const double dx = fmax(fmax(b.x - fmax(xa, xb), fmin(xa, xb) - b.y), 0.0);
        const double dy = fmax(fmax(b.z - fmax(ya, yb), fmin(ya, yb) - b.w), 0.0);
        on = dx * dx + dy * dy <= p.skip_r2;
      }
      const uint32_t bits = __ballot_sync(0xffffffffu, on);
      if ((tid & 31) == 0) cmask[tid >> 5] = bits;
      __syncthreads();
    }
    const double cx0 = cell_coord(p.x0, p.rx, col0 + sg), cy0 = cell_coord(p.y0, p.ry, p.row0 + r0 + sg);
    double acc[8][8];
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[r][c] = 0.0;

    for (int k0 = 0; k0 < p.n; k0 += kT2Chunk) {
      if (p.bbox && !((cmask[k0 >> 11] >> ((k0 >> 6) & 31)) & 1u)) continue;   // out of reach: K* < 2^-bits sigma_f^2 for the whole patch
      {
        const int j = k0 + jj;
        double ox = 0.0, oy = 0.0, ol = 0.0;  // observations beyond n: Linv = 0, so b = 0
        if (j < p.n) {
          const double2 od = *reinterpret_cast<const double2*>(p.xy + 2 * (size_t)j);
          ox = od.x; oy = od.y; ol = p.sf2 * lrow[j];
        }
        const double dx = cx0 - ox, dy = cy0 - oy;
        double ea = exp(-dx * dx * p.inv2l2), ga = exp(-(2.0 * dx * sx + sx * sx) * p.inv2l2);
        double eb = exp(-dy * dy * p.inv2l2) * ol, gb = exp(-(2.0 * dy * sy + sy * sy) * p.inv2l2);
        double* da = a_s + jj * kT2Ld + sg;
        double* db = b_s + jj * kT2Ld + sg;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          da[8 * i] = ea;
          db[8 * i] = eb;
          ea *= ga; ga *= qx;
          eb *= gb; gb *= qy;
        }
      }
      __syncthreads();
#pragma unroll 2
      for (int q = 0; q < kT2Chunk; ++q) {
        double a[8], b[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const double2 av = *reinterpret_cast<const double2*>(&a_s[q * kT2Ld + 2 * tx + 32 * i]);
          const double2 bv = *reinterpret_cast<const double2*>(&b_s[q * kT2Ld + 2 * ty + 32 * i]);
          a[2 * i] = av.x; a[2 * i + 1] = av.y;
          b[2 * i] = bv.x; b[2 * i + 1] = bv.y;
        }
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
          for (int c = 0; c < 8; ++c) acc[r][c] = fma(b[r], a[c], acc[r][c]);
      }
      __syncthreads();
    }

#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int row = r0 + 2 * ty + 32 * (r >> 1) + (r & 1);
      if (row >= p.rows_local) continue;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int col = col0 + 2 * tx + 32 * (c >> 1) + (c & 1);
        if (col < p.W) {
          const long long lc = (long long)row * p.W + col;
          const double u = acc[r][c];
          s2_out[lc] = s2_out[lc] - u * u;
          mu_out[lc] = fma(w, u, mu_out[lc]);
        }
      }
    }
  }
}

// ---- the same tile on the fp64 tensor path (mma.sync m8n8k4): what the fp64 one-new-row step runs by default ----------
// u[r][c] = sum_j b[j][r] a[j][c] is a GEMM over the observations (M = 128 grid rows, N = 128 grid columns, K = the 64
// observations of a chunk per stage).  The register tile above issues 64 DFMA + 8 LDS.128 per observation and thread and
// holds 57 % of the fp64 pipe; one DMMA is 256 FMAs behind two 8-byte fragment loads (posterior_f64.cu holds 77-79 % with
// the same instruction).  512 threads = 16 warps as 4 (rows) x 4 (columns), warp tile 32 x 32 = 4 x 4 DMMA tiles, 32
// accumulators per thread.  Staged rows are 132 doubles long: the (k = lane % 4, cell = lane / 4) fragment reads of a
// half-warp then cover all 32 banks once.
// Staging costs three exp per (observation, factor) instead of two per 16 staged values: with d0 = X_0 - x_j at the first
// cell of the tile edge and pitch h,
//   e(c) = exp(-d0^2 / 2l^2) * g^c * T[c],      g = exp(-2 d0 h / 2l^2),  T[c] = exp(-c^2 h^2 / 2l^2)   (c = 0 .. 15),
//   e(c + 16 (i + 1)) = e(c + 16 i) * G_i,      G_0 = exp(-(32 d0 h + 256 h^2) / 2l^2) * R[c],  R[c] = exp(-32 c h^2 / 2l^2),
//   G_(i+1) = G_i * exp(-512 h^2 / 2l^2);
// 128 threads compute (e(0), g, g^2, g^4, g^8, G_0 without R) per (observation, factor) into shared memory, then every
// thread forms its anchors with <= 6 multiplications (g^c from the bits of c; T, R from two 16-entry tables) and runs
// eight recurrence steps: 3e-14 relative within the reach of the walk (tests/test_identities.py; the exponentials themselves carry |x| ulp).  An anchor that underflows sits >= 34
// length-scales from every cell of the edge (the host requires 128 h <= 4 l); its ratios are zeroed with it so that a huge
// ratio cannot meet the zero.
constexpr int kD2Threads = 512;
constexpr int kD2Ld = kT2 + 4;
constexpr int kD2Chunk = 64;

__device__ __forceinline__ void dmma8x8x4(double (&c)[2], double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(kD2Threads, 1)
posterior_thin2d_dmma_kernel(ThinParams p, double* __restrict__ mu_out, double* __restrict__ s2_out) {
  extern __shared__ __align__(16) double sm2[];
  double* a_s = sm2;                         // [kD2Chunk][kD2Ld]: a_s[j][c] = ex(column c, observation j)
  double* b_s = sm2 + kD2Chunk * kD2Ld;      // [kD2Chunk][kD2Ld]: b_s[j][r] = sigma_f^2 ey(row r, j) Linv[new row][j]
  __shared__ __align__(16) double anc[2][kD2Chunk][8];   // per (factor, observation): e(0), g, g^2, g^4, g^8, G_0
  __shared__ double tab_t[2][16], tab_r[2][16];
  __shared__ uint32_t cmask[kD2Threads / 32];            // 512 chunks = 32 768 observations
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3, wm = warp >> 2, wn = warp & 3;
  const int sg = tid & 15, jq = tid >> 4;   // staging role: cells sg + 16 i of the edge, observations jq and jq + 32 of the chunk
  const int tiles_x = (p.W + kT2 - 1) / kT2, tiles_y = (p.rows_local + kT2 - 1) / kT2;
  const long long ntiles = (long long)tiles_x * tiles_y;
  const double w = p.w[p.row_lo];
  const double* lrow = p.Linv + (size_t)p.row_lo * p.ld;
  const double q16[2] = {exp(-512.0 * p.rx * p.rx * p.inv2l2), exp(-512.0 * p.ry * p.ry * p.inv2l2)};
  if (tid < 32) {
    const int f = tid >> 4, c = tid & 15;
    const double h = f ? p.ry : p.rx;
    tab_t[f][c] = exp(-(double)(c * c) * h * h * p.inv2l2);
    tab_r[f][c] = exp(-32.0 * (double)c * h * h * p.inv2l2);
  }
  __syncthreads();
  const double tt[2] = {tab_t[0][sg], tab_t[1][sg]}, tr[2] = {tab_r[0][sg], tab_r[1][sg]};

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int r0 = (int)(tile / tiles_x) * kT2, col0 = (int)(tile % tiles_x) * kT2;
    if (p.bbox) {   // chunks within reach of this 128 x 128 patch (same box test and cut-off as posterior_f64.cu)
      __syncthreads();   // the previous tile's reads of the mask are over
      bool on = tid < (p.n + 63) / 64;
      if (on) {
        const double xa = cell_coord(p.x0, p.rx, col0), xb = cell_coord(p.x0, p.rx, min(col0 + kT2, p.W) - 1);
        const double ya = cell_coord(p.y0, p.ry, p.row0 + r0), yb = cell_coord(p.y0, p.ry, p.row0 + min(r0 + kT2, p.rows_local) - 1);
        const double4 b = p.bbox[tid];
        const double dx = fmax(fmax(b.x - fmax(xa, xb), fmin(xa, xb) - b.y), 0.0);
        const double dy = fmax(fmax(b.z - fmax(ya, yb), fmin(ya, yb) - b.w), 0.0);
        on = dx * dx + dy * dy <= p.skip_r2;
      }
      const uint32_t bits = __ballot_sync(0xffffffffu, on);
      if (lane == 0) cmask[warp] = bits;
      __syncthreads();
    }
    const double base[2] = {cell_coord(p.x0, p.rx, col0), cell_coord(p.y0, p.ry, p.row0 + r0)};
    double acc[4][4][2];
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
      for (int ni = 0; ni < 4; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

    for (int k0 = 0; k0 < p.n; k0 += kD2Chunk) {
      if (p.bbox && !((cmask[k0 >> 11] >> ((k0 >> 6) & 31)) & 1u)) continue;   // out of reach: K* < 2^-bits sigma_f^2 for the whole patch
      if (tid < 2 * kD2Chunk) {   // anchors of the chunk: one thread per (factor, observation)
        const int f = tid >> 6, q = tid & (kD2Chunk - 1), j = k0 + q;
        const double h = f ? p.ry : p.rx;
        double d0 = 0.0, scale = 0.0;   // observations beyond n: b = 0 (Linv = 0); a = the harmless exp(0) table
        if (j < p.n) {
          d0 = base[f] - p.xy[2 * (size_t)j + f];
          scale = f ? p.sf2 * lrow[j] : 1.0;
        } else if (!f) {
          scale = 1.0;
        }
        const double e0 = exp(-d0 * d0 * p.inv2l2) * scale;
        double g1 = exp(-2.0 * d0 * h * p.inv2l2), g16 = exp(-(32.0 * d0 * h + 256.0 * h * h) * p.inv2l2);
        if (e0 == 0.0) g1 = g16 = 0.0;
        const double g2 = g1 * g1, g4 = g2 * g2, g8 = g4 * g4;
        double2* dst = reinterpret_cast<double2*>(anc[f][q]);
        dst[0] = make_double2(e0, g1);
        dst[1] = make_double2(g2, g4);
        dst[2] = make_double2(g8, g16);
      }
      __syncthreads();
#pragma unroll
      for (int rep = 0; rep < 2; ++rep) {
        const int q = jq + 32 * rep;
#pragma unroll
        for (int f = 0; f < 2; ++f) {
          const double2* an = reinterpret_cast<const double2*>(anc[f][q]);
          const double2 v0 = an[0], v1 = an[1], v2 = an[2];
          double e = v0.x * tt[f];
          if (sg & 1) e *= v0.y;
          if (sg & 2) e *= v1.x;
          if (sg & 4) e *= v1.y;
          if (sg & 8) e *= v2.x;
          double gg = v2.y * tr[f];
          double* dst = (f ? b_s : a_s) + q * kD2Ld + sg;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            dst[16 * i] = e;
            e *= gg;
            gg *= q16[f];
          }
        }
      }
      __syncthreads();
      const double* As = b_s + t * kD2Ld + wm * 32 + g;   // A[m = grid row][k = observation]
      const double* Bs = a_s + t * kD2Ld + wn * 32 + g;   // B[k = observation][n = grid column]
#pragma unroll 4
      for (int k4 = 0; k4 < kD2Chunk / 4; ++k4) {
        double a[4], b[4];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi) a[mi] = As[k4 * 4 * kD2Ld + mi * 8];
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) b[ni] = Bs[k4 * 4 * kD2Ld + ni * 8];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi)
#pragma unroll
          for (int ni = 0; ni < 4; ++ni) dmma8x8x4(acc[mi][ni], a[mi], b[ni]);
      }
      __syncthreads();
    }

    // accumulator (mi, ni) of a lane = u at row wm 32 + mi 8 + g, columns wn 32 + ni 8 + 2 t + {0, 1}.  The stored values
    // of a row of four tiles are fetched before the first is rewritten.
    const bool vec = (p.W & 1) == 0;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi) {
      const int row = r0 + wm * 32 + mi * 8 + g;
      if (row >= p.rows_local) continue;
      double s2v[4][2], muv[4][2];
#pragma unroll
      for (int ni = 0; ni < 4; ++ni) {
        const int col = col0 + wn * 32 + ni * 8 + 2 * t;
        const long long lc = (long long)row * p.W + col;
        if (vec && col + 1 < p.W) {
          const double2 sv = *reinterpret_cast<const double2*>(s2_out + lc), mv = *reinterpret_cast<const double2*>(mu_out + lc);
          s2v[ni][0] = sv.x; s2v[ni][1] = sv.y; muv[ni][0] = mv.x; muv[ni][1] = mv.y;
        } else {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            s2v[ni][e] = col + e < p.W ? s2_out[lc + e] : 0.0;
            muv[ni][e] = col + e < p.W ? mu_out[lc + e] : 0.0;
          }
        }
      }
#pragma unroll
      for (int ni = 0; ni < 4; ++ni) {
        const int col = col0 + wn * 32 + ni * 8 + 2 * t;
        const long long lc = (long long)row * p.W + col;
        const double u0 = acc[mi][ni][0], u1 = acc[mi][ni][1];
        const double n0 = s2v[ni][0] - u0 * u0, n1 = s2v[ni][1] - u1 * u1;
        const double m0 = fma(w, u0, muv[ni][0]), m1 = fma(w, u1, muv[ni][1]);
        if (vec && col + 1 < p.W) {
          *reinterpret_cast<double2*>(s2_out + lc) = make_double2(n0, n1);
          *reinterpret_cast<double2*>(mu_out + lc) = make_double2(m0, m1);
        } else {
          if (col < p.W) { s2_out[lc] = n0; mu_out[lc] = m0; }
          if (col + 1 < p.W) { s2_out[lc + 1] = n1; mu_out[lc + 1] = m1; }
        }
      }
    }
  }
}

// (An fp32 form of this tile -- 64 FFMA per 4 LDS.128, MUFU staging, fp64 flush every 64 observations -- measured 12.3e12
// against the strip kernel's 11.7e12 cell-observations/s at n = 16 384: 255 registers, one CTA per SM, no room to prefetch the
// next observation's factors.  Not kept.)

template <typename T, int KR, int C, int R>
void launch_thin(sbo_ctx* ctx, const ThinParams& p) {
  const int span = kThinThreads * C;
  const long long ntiles = (long long)((p.rows_local + R - 1) / R) * ((p.W + span - 1) / span);
  int per_sm = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, posterior_thin_kernel<T, KR, C, R>, kThinThreads, 0);
  const unsigned grid = (unsigned)std::min<long long>(ntiles, (long long)ctx->sm_count * std::max(per_sm, 1));
  posterior_thin_kernel<T, KR, C, R><<<grid, kThinThreads, 0, ctx->stream>>>(p, static_cast<T*>(ctx->d_mu), static_cast<T*>(ctx->d_s2));
}

}  // namespace

// Incremental posterior for new rows [row_lo, n) of L^-1, n - row_lo <= 8.
int32_t launch_posterior_thin(sbo_ctx* ctx, int row_lo) {
  if (ctx->m_local == 0) return SBO_OK;
  const int k_new = ctx->n - row_lo;
  if (k_new < 0 || k_new > kThinMaxRows) return fail(ctx, SBO_ERR_STATE, "posterior_thin: more than 8 new rows");
  ThinParams p;
  p.bbox = nullptr;
  p.skip_r2 = 0.0;
  p.obs_f32 = reinterpret_cast<const float4*>(ctx->d_obs_f32);
  p.xy = ctx->d_xy; p.Linv = ctx->d_Linv; p.w = ctx->d_w;
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
      case 1: launch_thin<float, 1, 2, 16>(ctx, p); break;
      case 2: launch_thin<float, 2, 2, 8>(ctx, p); break;
      case 4: launch_thin<float, 4, 1, 8>(ctx, p); break;
      default: launch_thin<float, 8, 1, 4>(ctx, p); break;
    }
  } else {
    const long long tiles2d = (long long)((p.W + kT2 - 1) / kT2) * ((p.rows_local + kT2 - 1) / kT2);
    const bool span_ok = kT2 * std::max(std::fabs(p.rx), std::fabs(p.ry)) <= 4.0 * ctx->cfg.length_scale;  // see the staging recurrence
    p.bbox = nullptr;
    p.skip_r2 = 0.0;
    if (kr == 1 && tiles2d >= ctx->sm_count && span_ok && ctx->opt.thin_2d) {   // enough 128 x 128 tiles for every SM
      const void* boxes = nullptr;
      const unsigned grid = (unsigned)std::min<long long>(tiles2d, ctx->sm_count);
      // (the attributes are per device, so not cached in a process-wide flag: a process may hold contexts on several GPUs)
      if (ctx->opt.thin_2d == 2) {   // the DFMA register tile, kept for A/B
        SBO_TRY(f64_chunk_boxes(ctx, kThinThreads, &boxes, &p.skip_r2));
        p.bbox = static_cast<const double4*>(boxes);
        const size_t smem = (size_t)2 * kT2Chunk * kT2Ld * sizeof(double);
        const cudaError_t ea = cudaFuncSetAttribute(posterior_thin2d_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (ea != cudaSuccess) return cuda_fail(ctx, ea, "posterior_thin2d_f64_kernel: cudaFuncSetAttribute(max dynamic smem)");
        posterior_thin2d_f64_kernel<<<grid, kThinThreads, smem, ctx->stream>>>(p, static_cast<double*>(ctx->d_mu), static_cast<double*>(ctx->d_s2));
        SBO_LAUNCH_CHECK(ctx, "posterior_thin2d_f64_kernel");
        return SBO_OK;
      }
      SBO_TRY(f64_chunk_boxes(ctx, kD2Threads, &boxes, &p.skip_r2));
      p.bbox = static_cast<const double4*>(boxes);
      const size_t smem = (size_t)2 * kD2Chunk * kD2Ld * sizeof(double);
      const cudaError_t ea = cudaFuncSetAttribute(posterior_thin2d_dmma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (ea != cudaSuccess) return cuda_fail(ctx, ea, "posterior_thin2d_dmma_kernel: cudaFuncSetAttribute(max dynamic smem)");
      posterior_thin2d_dmma_kernel<<<grid, kD2Threads, smem, ctx->stream>>>(p, static_cast<double*>(ctx->d_mu), static_cast<double*>(ctx->d_s2));
      SBO_LAUNCH_CHECK(ctx, "posterior_thin2d_dmma_kernel");
      return SBO_OK;
    }
    switch (kr) {
      case 1: launch_thin<double, 1, 1, 32>(ctx, p); break;   // one exp per 32 cells: the fp64 exp is ~14 DFMA-equivalents
      case 2: launch_thin<double, 2, 1, 16>(ctx, p); break;
      case 4: launch_thin<double, 4, 1, 8>(ctx, p); break;
      default: launch_thin<double, 8, 1, 4>(ctx, p); break;
    }
  }
  SBO_LAUNCH_CHECK(ctx, "posterior_thin_kernel");
  return SBO_OK;
}

}  // namespace sbo
