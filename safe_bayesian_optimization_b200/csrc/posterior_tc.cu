// posterior_tc.cu -- the tcgen05/TMEM GP-posterior kernel (fp32 mode).
//
//   V = K* L^-T   (m x n) . (n x n, lower triangular),   sigma^2 = sigma_f^2 - rowsum(V o V)
//   mu = prior_mean + K* alpha
//
// K* (RBF cross-covariance between every cell and every observation) is never materialised:
// generator warps compute 128 x 64 tiles of it on the fly, split every entry into two fp16 planes
// (hi = rn_f16(x), lo = rn_f16(x - hi), x pre-scaled by a power of two so that neither plane leaves
// the fp16 normal range where it matters) and store them straight into the 128B-swizzled K-major
// shared-memory layout the UMMA descriptors expect.  The matching tiles of the pre-split, pre-scaled
// hi/lo planes of L^-1 arrive by TMA.  One elected thread issues
//   D += A_lo.B_hi ; D += A_hi.B_lo ; D += A_hi.B_hi        (tcgen05.mma kind::f16, N256 K16)
// so the fp32 accumulator in TMEM carries ~2^-22 relative products ("3xFP16": the same 11+11
// significand bits as a 3xTF32 split, at twice the tensor rate and half the operand bytes).
// Epilogue warps pull finished 128 x 256 accumulators out of TMEM with tcgen05.ld, square and
// row-sum them in fp64 and keep only one number per cell: V never reaches shared or global memory.
// Two accumulator stages (2 x 256 TMEM columns) let the epilogue of row block jb overlap the MMAs
// of jb+1.  L^-1 is lower triangular, so row block jb only walks k < 256 (jb+1).
//
// CG = 2 (default): the two CTAs of a cluster (one TPC) run tcgen05.mma.cta_group::2 -- each CTA
// owns 128 cells (its own A tiles, its own accumulators) and HALF of every B tile (128 of the 256
// rows of L^-1), which the tensor cores of both SMs read.  That halves the shared-memory operand
// reads and the L2->SM traffic of L^-1 per SM, the two things that capped the 1-CTA kernel
// (profiles/ncu_posterior_tc_r01.md).  CG = 1 is the single-CTA form of the same code.
//
// Warp roles (768 threads, 1 CTA/SM, persistent over cell tiles):
//   0      TMA producer (converged warp, elect.sync)     1  MMA issuer (same, leader CTA)     2  TMEM alloc/dealloc
//   4-7    epilogue (TMEM lane quarter = warp % 4)
//   8-23   K* generators in two groups of 8 on alternate visits: warp w -> cells 32 (w & 3) .. +31, observations
//          32 (((w - 8) >> 2) & 1) .. +31 of the chunk
// The full refresh over >= 2 row blocks runs posterior_tc_pair.cuh (one K* generation per pair of row blocks), large-grid
// incremental steps posterior_tc_inc.cuh (a GEMM through the separable RBF); this kernel serves the rest: the
// incremental rank-k step on the rows that hold the new observations, n <= 256, cta_group::1.
// (profiles/ncu_posterior_tc_r01d.md has the measurements behind this organisation.)
#include <cuda.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace sbo {
namespace {

// -DSBO_TC_CLOCKS=1 compiles the per-role cycle counters in (SBO_TC_CLK=1 prints them); off by default: they cost registers
#ifndef SBO_TC_CLOCKS
#define SBO_TC_CLOCKS 0
#endif
// SBO_TC_DBG bit flags switch parts of the kernel off for timing experiments (results are then wrong by design):
// they only exist in -DSBO_TC_CLOCKS=1 builds.
#define TC_DBG(p, bit) (SBO_TC_CLOCKS && ((p).dbg & (bit)))

constexpr int BM = 128;          // cells per CTA tile (UMMA M per CTA)
constexpr int BN = 256;          // rows of L^-1 per accumulator (UMMA N)
constexpr int BK = 64;           // k per stage: 64 fp16 = one 128-byte swizzle row
constexpr int UK = 16;           // k per tcgen05.mma kind::f16
constexpr int kThreadsTc = 768;
constexpr int kGenWarps = 16;
constexpr uint32_t kABytes = BM * BK * 2;   // 16 KB per plane
constexpr uint32_t kTmemCols = 512;
constexpr int kMaskWords = 16;         // chunk masks cover 512 chunks = 32 768 observations; beyond that: dense

template <int CG> struct TcCfg {
  static constexpr uint32_t kBRows = BN / CG;                    // rows of a B tile held by one CTA
  static constexpr uint32_t kBBytes = kBRows * BK * 2;           // 32 KB (CG=1) / 16 KB (CG=2) per plane
  static constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytes;  // 96 KB / 64 KB
  static constexpr int kStages = CG == 1 ? 2 : 3;
};

struct __align__(16) TcShared {
  unsigned long long full[4];      // leader CTA's copy is the one in use
  unsigned long long empty[4];     // local (tcgen05.commit multicasts to both CTAs)
  unsigned long long tmem_full[2]; // local
  unsigned long long tmem_empty[2];// leader CTA's copy is the one in use
  uint32_t tmem_base;
  uint32_t pad[3];
  float4 obs_x[kGenWarps][8];      // per generator warp: scaled x - tile origin of its 32 observations
  float4 obs_y[kGenWarps][8];      //   y factor of K* (incl. sigma_f^2 and the A scale)
  float4 obs_a[kGenWarps][8];      //   alpha
  float mu_part[2][3][BM];         // partial mu of k-quarters 1..3, double-buffered by tile parity
  uint32_t cmask[kThreadsTc / 32][2 * kMaskWords];  // per warp: [0, 16) bit c set = chunk c is contracted for the current tile (pair); [16, 32) = with all three split products
};

struct TcParams {
  const float4* obs;       // [n_cap] scaled observation coordinates (x_hi, x_lo, y_hi, y_lo), see pack_obs_kernel
  const float* alpha;      // [n_cap] fp32
  float* mu;               // [m_local]
  float* s2;               // [m_local]
  int n, n_cap;
  int j_lo, j_hi;          // rows of L^-1 that count (incremental step: only the new rows)
  int accumulate;
  int W, rows_local, row0; // this shard: grid rows [row0, row0 + rows_local)
  double x0, y0, rx, ry;
  double xscale;           // sqrt(log2(e) / (2 l^2)): coordinates are pre-scaled so that K* = ex2(-dx^2) ey
  float a_scale;           // sigma_f^2 2^shift_a
  float mu_unscale;        // 2^-shift_a
  double v_unscale2;       // 2^(-2 (shift_a + shift_b))
  float v_unscale;         // 2^-(shift_a + shift_b) (debug dump only)
  float sf2, prior_mean;
  float* dbg_v;            // debug: V of tile 0 ([128][n_cap]) when non-null
  const float4* bbox;      // [n_cap / 64] scaled bounding box (xmin, xmax, ymin, ymax) of every 64-observation chunk
  int skip;                // 1: chunks whose K* entries are all below 2^-skip_bits sigma_f^2 for a tile are not contracted
  float skip_bits;         // tc_skip_bits(): chosen so that everything skipped moves sigma^2 by < 1e-7 sigma_f^2
  float ring_bits;         // graded split depth: visits whose K* entries are all < 2^-ring_bits sigma_f^2 issue hi.hi only (>= skip_bits: none)
  int strip_tiles;         // strip-major tile order: tiles per strip (even; 0 = raster), see TileMap
  int pi_lo, pi_hi;        // pair kernel: row-block pairs [pi_lo, pi_hi) of this launch
  int rs_mode;             // pair kernel: 0 whole job; bit 0: partial (store the running row sums in rs_part, no sigma^2 / mu); bit 1: resume from rs_part (3: a middle phase)
  float2* rs_part;         // [m_local] double-word fp32 running sums of a partial launch
  unsigned long long* exec_cols;   // pair kernel: += sum of N over the issued MMAs (M 256, K 16 each); executed flops = 2 * 256 * 16 * that
  long long* dbg_clk;      // SBO_TC_CLK=1: per-role cycle counters of cluster 0 (timing forensics)
  int dbg;                 // SBO_TC_DBG (timing experiments only, results are wrong): 1 generators skip the K* math, 2 epilogue skips the TMEM reads, 4 one MMA pass instead of three
};

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of `local` (a shared::cta address) in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
// arrive on a barrier addressed in the shared::cluster window (own or peer CTA).  Default .release.cta /
// .acquire.cta semantics on purpose: cluster-scope acquire makes ptxas emit CCTL.IVALL (an L1 flush) after
// every wait; the data handed over lives in shared memory and is ordered by fence.proxy.async.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_arrive_local(unsigned long long* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// TMA: 2-D box into this CTA's shared memory; completion bytes are signalled on `bar_cluster_addr`
// (a shared::cluster address: the leader CTA's barrier when CG = 2).
template <int CG>
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint32_t bar_cluster_addr, void* dst, int c0, int c1) {
  if constexpr (CG == 1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
  } else {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
  }
}

template <int CG> __device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  if constexpr (CG == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
}
template <int CG> __device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  if constexpr (CG == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
  else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}

// D[tmem] (+)= A[smem] . B[smem]^T, kind::f16 (fp16 operands, fp32 accumulate), issued by one thread.
template <int CG>
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  if constexpr (CG == 1) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// mbarrier arrive (in every CTA of the group) when every tcgen05.mma issued so far by this thread has completed.
template <int CG> __device__ __forceinline__ void umma_commit(unsigned long long* bar) {
  if constexpr (CG == 1) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
  } else {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"((uint16_t)3)
                 : "memory");
  }
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory descriptor: K-major operand, 128-byte swizzle, rows 128 B apart, 8-row groups
// 1024 B apart (SBO), descriptor version 1 (sm_100).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor: D=f32, A=B=f16 (format 0), both K-major, N=256, M=128 per CTA.
template <int CG> __device__ __forceinline__ uint32_t idesc_f16(uint32_t n_cols) {
  return (1u << 4) | ((n_cols >> 3) << 17) | ((uint32_t)((BM * CG) >> 4) << 24);
}

// One lane of a CONVERGED warp (the lowest).  The single-thread tcgen05 / TMA instructions are issued under this
// predicate from warp-uniform code: with operands in uniform registers ptxas emits them directly; inside an
// `if (lane == 0)` region it wraps every one of them in an ELECT / R2UR.BROADCAST / BRA.U.ANY waterfall loop
// (~20 instructions per MMA: the issuer, not the tensor pipe, then paces the kernel).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P1;\n\telect.sync _|P1, 0xffffffff;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t h2_bits(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// number of k-chunks of row block jb
__device__ __forceinline__ int chunks_of(int jb, int n_pad) { return min(n_pad, (jb + 1) * BN) / BK; }

// ---- block sparsity of K* ------------------------------------------------------------------------
// The observations are stored in spatial order (sbo_set_observations), so a chunk of 64 consecutive ones
// covers a small box, and K*[cell, k] = sigma_f^2 2^-(d^2) in the scaled coordinates.  A chunk whose box is
// farther than sqrt(skip_bits) from every cell of the tile (pair) contributes less than 2^-skip_bits sigma_f^2 per
// entry -- below the fp16 split's own resolution (2^-22) squared -- and is left out by every role alike:
// each warp builds the same bit mask for itself (no cross-warp hand-over) and walks its set bits.
__device__ __forceinline__ float box_dist2(const float4 b, float xlo, float xhi, float y) {
  const float dx = fmaxf(fmaxf(b.x - xhi, xlo - b.y), 0.f), dy = fmaxf(fmaxf(b.z - y, y - b.w), 0.f);
  return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}

// Tile order.  Raster order makes the co-resident CTA pairs (74 of them = 148 tiles) cover a one-row strip across the WHOLE
// width of the grid: together they touch every observation chunk of the ~4 bands of observations near that row, for every
// row block of L^-1 -- ~94 MB of fp16 planes per sweep at n = 16 384, more than the L2 holds once the pairs drift apart
// (measured round 1: 50 % of the TMA bytes came from DRAM, 2.4 TB/s).  Strip-major order walks a strip of `strip` tiles
// (a few length-scales wide) down all the rows of the shard before the next strip, so the co-resident pairs cover a compact
// patch and share most of their chunks.  Needs an even number of tiles per row (a tile pair never straddles rows).
struct TileMap {
  int tiles_per_row, rows, strip;   // strip: tiles per strip (even) or 0 = raster
  __device__ __forceinline__ void locate(long long tile, int& lrow, int& col0) const {
    if (strip == 0) {
      lrow = (int)(tile / tiles_per_row);
      col0 = (int)(tile % tiles_per_row) * BM;
    } else {
      const long long per_strip = (long long)strip * rows;
      const int nstrips = (tiles_per_row + strip - 1) / strip;
      const int s = (int)min(tile / per_strip, (long long)(nstrips - 1));
      const long long rem = tile - (long long)s * per_strip;
      const int w = min(strip, tiles_per_row - s * strip);
      lrow = (int)(rem / w);
      col0 = (s * strip + (int)(rem % w)) * BM;
    }
  }
};

// __noinline__ on purpose: every role of both CTAs must derive bit-identical masks (they walk them in lock step
// through the mbarrier pipelines), so there is exactly one compiled copy of this arithmetic.
struct MaskGeom {  // by value: taking the address of the kernel parameter block would move it to local memory
  double x0, y0, rx, ry, xscale;
  const float4* bbox;
  int row0, skip;
  float skip_bits, ring_bits;
  TileMap tm;
};

// mask[0 .. 16): chunks that are contracted; mask[16 .. 32): those of them that need all three split products (the rest
// -- every K* entry of the chunk below 2^-ring_bits sigma_f^2 for both tiles -- run hi.hi only: graded split depth).
template <int CG>
__device__ __noinline__ void build_chunk_mask(uint32_t* mask, const MaskGeom p, long long gt, long long ntiles, int nchunks) {
  const int lane = threadIdx.x & 31;
  float xlo[CG], xhi[CG], yy[CG];
#pragma unroll
  for (int t = 0; t < CG; ++t) {
    const long long tile = min(gt * CG + t, ntiles - 1);
    int lrow, col0;
    p.tm.locate(tile, lrow, col0);
    const double xs = cell_coord(p.x0, p.rx, col0) * p.xscale, ys = cell_coord(p.y0, p.ry, p.row0 + lrow) * p.xscale;
    const double w = (double)(BM - 1) * p.rx * p.xscale;
    xlo[t] = (float)fmin(xs, xs + w) - 1e-3f;   // margin: the box test runs in fp32
    xhi[t] = (float)fmax(xs, xs + w) + 1e-3f;
    yy[t] = (float)ys;
  }
  __syncwarp();  // the previous tile's walks of this warp's mask are over
#pragma unroll 1
  for (int w = 0; w < kMaskWords; ++w) {
    const int c = w * 32 + lane;
    bool on = false, full = false;
    if (c < nchunks) {
      on = full = true;
      if (p.skip) {
        const float4 b = p.bbox[c];
        on = full = false;
#pragma unroll
        for (int t = 0; t < CG; ++t) {
          const float d2 = box_dist2(b, xlo[t], xhi[t], yy[t]);
          on = on || d2 <= p.skip_bits + 0.5f;
          full = full || d2 <= p.ring_bits + 0.5f;
        }
      }
    }
    const uint32_t bits = __ballot_sync(0xffffffffu, on), fbits = __ballot_sync(0xffffffffu, full);
    if (lane == 0) {
      mask[w] = bits;
      mask[kMaskWords + w] = fbits;
    }
  }
  __syncwarp();
}
__device__ __forceinline__ bool chunk_full(const uint32_t* mask, int kc) { return (mask[kMaskWords + (kc >> 5)] >> (kc & 31)) & 1u; }

// Walk over the set bits of a chunk mask in ascending order: one shared-memory load per 32 chunks, a pop
// (bits &= bits - 1) per chunk -- this sits on the per-chunk path of every role.  kSkip = false (the host found
// next to nothing to skip for this grid and observation set, tc_plan_skip): plain counting, no mask at all.
template <bool kSkip> struct ChunkIter {
  const uint32_t* m;
  int w;
  uint32_t bits;
  __device__ __forceinline__ void reset(const uint32_t* mask) {
    if (kSkip) { m = mask; w = 0; bits = mask[0]; } else { w = 0; }
  }
  __device__ __forceinline__ int next() {  // next contracted chunk, or a large number when none is left
    if (!kSkip) return w++;
    while (!bits) {
      if (++w >= kMaskWords) return 1 << 30;
      bits = m[w];
    }
    const int c = w * 32 + __ffs(bits) - 1;
    bits &= bits - 1;
    return c;
  }
};

// The (row block, chunk) visits of one cell tile in pipeline order -- every role walks this same sequence.
template <bool kSkip> struct TileWalk {
  ChunkIter<kSkip> ci;
  const uint32_t* mask;
  int jb, jb_hi, n_pad, nch;
  __device__ __forceinline__ void start(const uint32_t* m, int jb_lo_, int jb_hi_, int n_pad_) {
    mask = m; jb = jb_lo_; jb_hi = jb_hi_; n_pad = n_pad_;
    nch = jb < jb_hi ? chunks_of(jb, n_pad) : 0;
    ci.reset(mask);
  }
  // chunk of the next visit (its row block is `jb`), or -1 when the tile is done
  __device__ __forceinline__ int next() {
    while (jb < jb_hi) {
      const int kc = ci.next();
      if (kc < nch) return kc;
      if (++jb < jb_hi) {
        nch = chunks_of(jb, n_pad);
        ci.reset(mask);
      }
    }
    return -1;
  }
};

// One thread's share of a K* chunk: cell r of the tile x 16 observations (2 groups of 8 = 2 16-byte
// swizzle units of the row), written as fp16 hi / lo planes.  Returns sum_k K*[r,k] alpha[k] (scaled).
//   K*[r,k] 2^shift_a = ex2(-(xs_k - cxs_r)^2) ey_k,   xs, cxs = x sqrt(log2(e) / 2 l^2),
//   ey_k = sigma_f^2 2^shift_a ex2(-(y_k - y_row)^2 log2(e) / 2 l^2)
// (the scale multiplies instead of shifting the exponent argument: the argument keeps its small magnitude,
// so the MUFU input carries full fp32 resolution where K* is large).
template <bool kMu>
__device__ __forceinline__ float gen_chunk(const float4* ox4, const float4* oy4, const float4* oa4, float cxs,
                                           uint32_t (&hi)[8], uint32_t (&lo)[8]) {
  float mu = 0.f;
#pragma unroll
  for (int qd = 0; qd < 2; ++qd) {
    const float4 xa = ox4[2 * qd], xb = ox4[2 * qd + 1];  // warp-wide broadcast loads
    const float4 ya = oy4[2 * qd], yb = oy4[2 * qd + 1];
    const float xs[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
    const float ey[8] = {ya.x, ya.y, ya.z, ya.w, yb.x, yb.y, yb.z, yb.w};
    float kv[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float dx = xs[e] - cxs;
      // MUFU ex2 is accurate enough here: swapping it for a 1-ulp polynomial changed neither mu nor
      // sigma^2 errors (profiles/accuracy_r01.md); the error budget is set by the accumulator.
      kv[e] = ex2_approx(-dx * dx) * ey[e];
    }
    if (kMu) {
      const float4 aa = oa4[2 * qd], ab = oa4[2 * qd + 1];
      const float as[8] = {aa.x, aa.y, aa.z, aa.w, ab.x, ab.y, ab.z, ab.w};
#pragma unroll
      for (int e = 0; e < 8; ++e) mu = fmaf(kv[e], as[e], mu);
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __half2 h = __floats2half2_rn(kv[2 * e], kv[2 * e + 1]);
      const float2 hf = __half22float2(h);
      hi[4 * qd + e] = h2_bits(h);
      lo[4 * qd + e] = h2_bits(__floats2half2_rn(kv[2 * e] - hf.x, kv[2 * e + 1] - hf.y));
    }
  }
  return mu;
}

// fp32 double-word accumulation (2Sum): (hi, lo) += x without an fp64 instruction on the generators' path
__device__ __forceinline__ void two_sum_acc(float& hi, float& lo, float x) {
  const float t = hi + x;
  const float z = t - hi;
  lo += (hi - (t - z)) + (x - z);
  hi = t;
}

// ---- packed fp32 pairs (sm_100: add/mul/fma.f32x2 -> FADD2 / FMUL2 / FFMA2, two results per issue slot) ----------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float a, float b) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

// One thread's share of HALF a K* chunk-half: cell r x 16 observations (two 16-byte swizzle units per plane), the
// arithmetic of gen_chunk on packed fp32 pairs: per pair of entries  FADD2 (dx), FMUL2 (dx^2), 2 MUFU.EX2, FMUL2 (ey),
// [FFMA2 (mu)], F2FP (hi), 2 HADD2.F32 (hi back to fp32), FADD2 (residual), F2FP (lo) = 5 issue slots per entry
// instead of 7.  ncx2 = (-cxs, -cxs).  mu2 accumulates sum_k K*[r,k] alpha[k] in two fp32 lanes (even / odd k).
template <bool kMu>
__device__ __forceinline__ void gen16(const float4* ox4, const float4* oy4, const float4* oa4, f32x2 ncx2, uint32_t (&hi)[8],
                                      uint32_t (&lo)[8], f32x2& mu2) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 xa = ox4[q], ya = oy4[q];  // warp-wide broadcast loads
    float4 aa = make_float4(0.f, 0.f, 0.f, 0.f);
    if (kMu) aa = oa4[q];
    const f32x2 xs[2] = {pack2(xa.x, xa.y), pack2(xa.z, xa.w)};
    const f32x2 ey[2] = {pack2(ya.x, ya.y), pack2(ya.z, ya.w)};
    const f32x2 al[2] = {pack2(aa.x, aa.y), pack2(aa.z, aa.w)};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const f32x2 d = add2(xs[h], ncx2);
      float t0, t1;
      unpack2(mul2(d, d), t0, t1);
      // MUFU ex2 is accurate enough here (profiles/accuracy_r01.md); the error budget is set by the accumulator.
      const f32x2 kv = mul2(pack2(ex2_approx(-t0), ex2_approx(-t1)), ey[h]);
      if (kMu) mu2 = fma2(kv, al[h], mu2);
      float k0, k1;
      unpack2(kv, k0, k1);
      const __half2 hh = __floats2half2_rn(k0, k1);
      const float2 hf = __half22float2(hh);
      float r0, r1;
      unpack2(add2(kv, pack2(-hf.x, -hf.y)), r0, r1);
      hi[2 * q + h] = h2_bits(hh);
      lo[2 * q + h] = h2_bits(__floats2half2_rn(r0, r1));
    }
  }
}

// Observation coordinates for the generators: scaled by sqrt(log2(e) / 2 l^2) and split into fp32 (hi, lo)
// pairs, so that a tile-relative coordinate is (x_hi - o_hi) + (x_lo - o_lo) in three fp32 adds -- exact to
// ~1e-7 of its own magnitude wherever the grid sits -- instead of fp64 arithmetic once per chunk.
__global__ void pack_obs_kernel(const double* __restrict__ xy, int n, int n_cap, double xscale, float4* __restrict__ out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_cap) return;
  float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
  if (k < n) {
    const double x = xy[2 * (size_t)k] * xscale, y = xy[2 * (size_t)k + 1] * xscale;
    o.x = (float)x;
    o.y = (float)(x - (double)o.x);
    o.z = (float)y;
    o.w = (float)(y - (double)o.z);
  }
  out[k] = o;
}

// Bounding box (scaled coordinates) of every 64-observation chunk; an empty chunk gets an inverted box at infinity.
__global__ void chunk_bbox_kernel(const float4* __restrict__ obs, int n, int nchunks, float4* __restrict__ bbox) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (c >= nchunks) return;
  float xmin = INFINITY, xmax = -INFINITY, ymin = INFINITY, ymax = -INFINITY;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int k = c * BK + h * 32 + lane;
    if (k < n) {
      const float4 o = obs[k];
      const float x = o.x + o.y, y = o.z + o.w;
      xmin = fminf(xmin, x); xmax = fmaxf(xmax, x);
      ymin = fminf(ymin, y); ymax = fmaxf(ymax, y);
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    xmin = fminf(xmin, __shfl_xor_sync(0xffffffffu, xmin, d)); xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, d));
    ymin = fminf(ymin, __shfl_xor_sync(0xffffffffu, ymin, d)); ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, d));
  }
  if (lane == 0) bbox[c] = make_float4(xmin, xmax, ymin, ymax);
}

// ---- the kernel ----------------------------------------------------------------------------------
template <int CG, bool kSkip>
__global__ void __launch_bounds__(kThreadsTc, 1)
posterior_tc_kernel(const __grid_constant__ CUtensorMap tmap, TcParams p) {
  using Cfg = TcCfg<CG>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  // 128B-swizzled tiles need 1024-byte alignment; pad by pointer arithmetic so the compiler keeps
  // the shared address space (STS/LDS, not generic stores).  The pad is identical in both CTAs of a
  // pair, which the cta_group::2 descriptors rely on (same offsets in both shared memories).
  uint8_t* tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  TcShared* sh = reinterpret_cast<TcShared*>(tiles + (size_t)kStages * Cfg::kStageBytes);
  // stage s: [A_hi 16K][A_lo 16K][B_hi][B_lo]
  auto a_hi = [&](int s) { return tiles + (size_t)s * Cfg::kStageBytes; };
  auto a_lo = [&](int s) { return tiles + (size_t)s * Cfg::kStageBytes + kABytes; };
  auto b_hi = [&](int s) { return tiles + (size_t)s * Cfg::kStageBytes + 2 * kABytes; };
  auto b_lo = [&](int s) { return tiles + (size_t)s * Cfg::kStageBytes + 2 * kABytes + Cfg::kBBytes; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = CG == 1 ? 0u : cluster_ctarank();
  const int tiles_per_row = (p.W + BM - 1) / BM;
  const long long ntiles = (long long)p.rows_local * tiles_per_row;
  const long long group = blockIdx.x / CG, ngroups = gridDim.x / CG;
  const long long ngtiles = (ntiles + CG - 1) / CG;  // tiles of CTA groups
  const int n_pad = (p.n + BK - 1) / BK * BK;
  const int jb_lo = p.j_lo / BN, jb_hi = (p.j_hi + BN - 1) / BN;
  const int nchunks = n_pad / BK;
  uint32_t* cmask = sh->cmask[warp];
  const TileMap tmap_tiles = {tiles_per_row, p.rows_local, 0};   // raster order (incremental steps: every tile walks the same few row blocks)
  const MaskGeom geom = {p.x0, p.y0, p.rx, p.ry, p.xscale, p.bbox, p.row0, p.skip, p.skip_bits, p.skip_bits, tmap_tiles};

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&sh->full[s], 1 + (kGenWarps / 2) * CG);  // leader's TMA producer + the 8 generator warps per CTA that own the visit
      mbar_init(&sh->empty[s], 1);          // tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sh->tmem_full[s], 1);        // tcgen05.commit
      mbar_init(&sh->tmem_empty[s], 4 * CG);  // every epilogue warp of the group
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<CG>(&sh->tmem_base, kTmemCols);
  tc_fence_before();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;
  const long long ck_begin = SBO_TC_CLOCKS ? clock64() : 0;

  if (warp == 0) {
    // =============================== TMA producer (converged warp, one elected lane issues) ===============================
    uint32_t it = 0;
    for (long long gt = group; gt < ngtiles; gt += ngroups) {
      if (kSkip) build_chunk_mask<CG>(cmask, geom, gt, ntiles, nchunks);
      for (int jb = jb_lo; jb < jb_hi; ++jb) {
        const int nch = chunks_of(jb, n_pad);
        bool fresh = true;
        ChunkIter<kSkip> ci;
        ci.reset(cmask);
        for (int kc = ci.next(); kc < nch; kc = ci.next(), ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          // chunk c = kc - 4 jb >= 0 of the diagonal block only meets rows >= 64 c of this row block (L^-1 is
          // lower triangular): the tile starts there and the MMA narrows to N = 256 - 64 c (see the issuer).
          // The first contracted chunk of a row block runs at full N: it initialises every accumulator column.
          // Only rows j_lo .. j_hi of L^-1 count (the incremental step: the new rows; a full refresh: rows < n), so the
          // tile and the MMA are also cut to the 64-row groups [c_min, c_end) of the row block that hold them.
          const int c_min = min(max(p.j_lo - jb * BN, 0) / BK, BN / BK - 1);
          const int c_end = max(min((p.j_hi - jb * BN + BK - 1) / BK, BN / BK), c_min + 1);
          const int c = min(max(fresh ? 0 : kc - jb * (BN / BK), c_min), c_end - 1);
          fresh = false;
          const int row = jb * BN + c * BK + (int)rank * ((c_end - c) * (BK / CG));
          mbar_wait(&sh->empty[s], ph ^ 1u);
          const uint32_t bar = CG == 1 ? smem_u32(&sh->full[s]) : mapa(smem_u32(&sh->full[s]), 0);
          if (elect_one()) {
            if TC_DBG(p, 64) {
              if (rank == 0) mbar_arrive_local(&sh->full[s]);
            } else {
              if (rank == 0) mbar_arrive_expect_tx(&sh->full[s], 2 * Cfg::kBBytes * CG);
              tma_load_2d<CG>(&tmap, bar, b_hi(s), kc * BK, row);
              tma_load_2d<CG>(&tmap, bar, b_lo(s), kc * BK, p.n_cap + row);
            }
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer (leader CTA; converged warp, one elected lane issues) ===============================
    uint32_t it = 0, acc_it = 0;
    long long ck_full = 0, ck_issue = 0, ck_acc = 0;
    const bool clk_on = SBO_TC_CLOCKS && p.dbg_clk != nullptr && blockIdx.x == 0;
    for (long long gt = group; rank == 0 && gt < ngtiles; gt += ngroups) {
      if (kSkip) build_chunk_mask<CG>(cmask, geom, gt, ntiles, nchunks);
      for (int jb = jb_lo; jb < jb_hi; ++jb) {
        const int nch = chunks_of(jb, n_pad);
        ChunkIter<kSkip> ci;
        ci.reset(cmask);
        int kc = ci.next();
        if (kc >= nch) continue;  // no chunk of this row block matters for the tile: V = 0, nothing to do
        const uint32_t as = acc_it & 1u, aph = (acc_it >> 1) & 1u;
        ++acc_it;
        const long long cka = clk_on ? clock64() : 0;
        mbar_wait(&sh->tmem_empty[as], aph ^ 1u);  // epilogue has drained this accumulator
        if (clk_on) ck_acc += clock64() - cka;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        bool fresh = true;
        for (; kc < nch; kc = ci.next(), ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          const long long ckf = clk_on ? clock64() : 0;
          mbar_wait(&sh->full[s], ph);  // TMA bytes landed and generators stored their planes
          const long long ckg = clk_on ? clock64() : 0;
          tc_fence_after();
          const uint64_t dah = umma_desc_sw128(smem_u32(a_hi(s))), dal = umma_desc_sw128(smem_u32(a_lo(s)));
          const uint64_t dbh = umma_desc_sw128(smem_u32(b_hi(s))), dbl = umma_desc_sw128(smem_u32(b_lo(s)));
          // diagonal block, chunk c: columns < 64 c of the accumulator get no contribution (zeros of L^-1)
          // ... and groups outside [c_min, c_end) hold no row in j_lo .. j_hi (see the producer)
          const int c_min = min(max(p.j_lo - jb * BN, 0) / BK, BN / BK - 1);
          const int c_end = max(min((p.j_hi - jb * BN + BK - 1) / BK, BN / BK), c_min + 1);
          const uint32_t c = (uint32_t)min(max(fresh ? 0 : kc - jb * (BN / BK), c_min), c_end - 1);
          const uint32_t idesc = idesc_f16<CG>((c_end - c) * BK), d_col = d_tmem + c * BK;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < BK / UK; ++ks) {
              const uint64_t adv = (uint64_t)((ks * UK * 2) >> 4);  // +32 bytes along K inside the swizzle row
              umma_f16<CG>(d_col, dal + adv, dbh + adv, idesc, (fresh && ks == 0) ? 0u : 1u);
              if TC_DBG(p, 4) continue;
              umma_f16<CG>(d_col, dah + adv, dbl + adv, idesc, 1u);
              umma_f16<CG>(d_col, dah + adv, dbh + adv, idesc, 1u);
            }
            umma_commit<CG>(&sh->empty[s]);  // stage reusable once these MMAs have read it
          }
          __syncwarp();
          fresh = false;
          if (clk_on) { ck_full += ckg - ckf; ck_issue += clock64() - ckg; }
        }
        if (elect_one()) umma_commit<CG>(&sh->tmem_full[as]);  // accumulator complete
        __syncwarp();
      }
    }
    if (clk_on && lane == 0) { p.dbg_clk[40] = ck_full; p.dbg_clk[41] = ck_issue; p.dbg_clk[42] = ck_acc; p.dbg_clk[43] = it; }
  } else if (warp >= 4 && warp < 8) {
    // =============================== epilogue ===============================
    const int q = warp & 3;
    const int r = q * 32 + lane;  // TMEM lane == row of the tile == cell
    uint32_t acc_it = 0;
    for (long long gt = group; gt < ngtiles; gt += ngroups) {
      const long long tile = gt * CG + rank;
      if (kSkip) build_chunk_mask<CG>(cmask, geom, gt, ntiles, nchunks);
      float rs_hi = 0.f, rs_lo = 0.f;  // sum of squares of this cell's row of V (scaled), double-word fp32
      ChunkIter<kSkip> ci0;
      ci0.reset(cmask);
      const int first_chunk = ci0.next();
      for (int jb = jb_lo; jb < jb_hi; ++jb) {
        if (first_chunk >= chunks_of(jb, n_pad)) continue;  // row block skipped by the issuer: V = 0
        const uint32_t as = acc_it & 1u, aph = (acc_it >> 1) & 1u;
        ++acc_it;
        mbar_wait(&sh->tmem_full[as], aph);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
        const int j0 = jb * BN;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          const int jc = j0 + c * 32;
          if (jc >= p.j_hi || jc + 32 <= p.j_lo) continue;  // columns outside the contracted rows (warp-uniform)
          if TC_DBG(p, 2) continue;
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_ld_wait();
          if (p.dbg_v && tile == 0) {
#pragma unroll
            for (int i = 0; i < 32; ++i) p.dbg_v[(size_t)r * p.n_cap + jc + i] = __uint_as_float(v[i]) * p.v_unscale;
          }
          if (jc < p.j_lo || jc + 32 > p.j_hi) {  // partial group (incremental step): mask the columns outside
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (jc + i < p.j_lo || jc + i >= p.j_hi) v[i] = 0u;
          }
          // squares are exact inside the FMA; four 8-term fp32 partial sums (relative error <= ~1e-7 each,
          // uncorrelated), then a double-word fp32 running sum (2Sum).  No fp64 here: with F2F/DADD the four
          // epilogue warps spent 3/4 of their samples in math-pipe throttle behind the generators' MUFU traffic
          // and the short row blocks of small n were paced by the epilogue (profiles/ncu_posterior_tc_r01b.md).
          float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float x0 = __uint_as_float(v[i]), x1 = __uint_as_float(v[i + 1]);
            const float x2 = __uint_as_float(v[i + 2]), x3 = __uint_as_float(v[i + 3]);
            s0 = fmaf(x0, x0, s0);
            s1 = fmaf(x1, x1, s1);
            s2 = fmaf(x2, x2, s2);
            s3 = fmaf(x3, x3, s3);
          }
          two_sum_acc(rs_hi, rs_lo, (s0 + s1) + (s2 + s3));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CG == 1) mbar_arrive_local(&sh->tmem_empty[as]);
          else mbar_arrive_cluster(mapa(smem_u32(&sh->tmem_empty[as]), 0));
        }
      }
      if (tile < ntiles) {
        const int lrow = (int)(tile / tiles_per_row), col = (int)(tile % tiles_per_row) * BM + r;
        if (col < p.W) {
          const long long lc = (long long)lrow * p.W + col;
          const double prev = p.accumulate ? (double)p.s2[lc] : (double)p.sf2;
          p.s2[lc] = (float)(prev - ((double)rs_hi + (double)rs_lo) * p.v_unscale2);
        }
      }
    }
  } else if (warp >= 8) {
    // =============================== K* generators ===============================
    // Two groups of 8 warps take alternate visits (grp = visit parity): a warp has two chunk times for its share,
    // and the per-chunk fixed cost (observation staging, fence, barrier traffic) is paid once per 32 observations.
    const int gw = warp - 8;          // 0..15
    const int grp = gw >> 3;          // parity of the visits this warp generates
    const int oh = (gw >> 2) & 1;     // which 32 of the chunk's 64 observations
    const int pq = grp * 2 + oh;      // slot of this warp's partial mu
    const int r = (gw & 3) * 32 + lane;
    float* ox = reinterpret_cast<float*>(sh->obs_x[gw]);
    float* oy = reinterpret_cast<float*>(sh->obs_y[gw]);
    float* oa = reinterpret_cast<float*>(sh->obs_a[gw]);
    const uint32_t row_off = (uint32_t)r * 128u;
    const uint32_t swz = (uint32_t)(r & 7);
    const float cxs = (float)((double)r * p.rx * p.xscale);  // scaled offset of this thread's cell from the tile origin
    const f32x2 ncx2 = pack2(-cxs, -cxs);
    uint32_t it = 0;
    uint32_t tile_par = 0;
    // observation data of this warp's NEXT visit (two visits ahead) is fetched while the current one is generated;
    // across tiles the guess is "the next tile starts like this one": checked at use (c_next != kc -> reload)
    float4 o_next = make_float4(0.f, 0.f, 0.f, 0.f);
    float a_next = 0.f;
    int c_next = -1;
    long long ck_math = 0, ck_wait = 0, ck_post = 0, ck_iters = 0, ck_stage = 0, ck_sync = 0, ck_h2 = 0, ck_fence = 0;
    const bool clk_on = SBO_TC_CLOCKS && p.dbg_clk != nullptr && blockIdx.x < 2 && (gw == 0 || gw == 15) && lane == 0;
    for (long long gt = group; gt < ngtiles; gt += ngroups, tile_par ^= 1u) {
      const long long tile = min(gt * CG + (long long)rank, ntiles - 1);  // an odd last tile: the peer recomputes it, no store
      const bool live = gt * CG + rank < ntiles;
      const int lrow = (int)(tile / tiles_per_row), col0 = (int)(tile % tiles_per_row) * BM;
      // scaled tile origin as an fp32 (hi, lo) pair
      const double ox0 = cell_coord(p.x0, p.rx, col0) * p.xscale, oy0 = cell_coord(p.y0, p.ry, p.row0 + lrow) * p.xscale;
      const float oxh = (float)ox0, oxl = (float)(ox0 - (double)oxh);
      const float oyh = (float)oy0, oyl = (float)(oy0 - (double)oyh);
      float mu_hi = 0.f, mu_lo = 0.f;  // fp32 inside a 32-observation half-chunk, double-word fp32 across chunks
      if (kSkip) build_chunk_mask<CG>(cmask, geom, gt, ntiles, nchunks);
      TileWalk<kSkip> w, scout;
      w.start(cmask, jb_lo, jb_hi, n_pad);
      scout = w;
      const int first0 = scout.next(), first1 = scout.next();  // visits 0 and 1 of this tile
      // visit parity is counted per tile (vi), not along the pipeline (it): which warps sum which observations into mu
      // must not depend on the tiles this CTA happened to process before (row-band shards agree bit for bit)
      uint32_t vi = 0;
      for (int kc = w.next(); kc >= 0; kc = w.next(), ++it, ++vi) {
        int kn = scout.next();  // the chunk two visits ahead
        if (kn < 0) kn = grp == 0 ? first0 : first1;  // next tile: this group's first visit there
        if ((vi & 1u) != (uint32_t)grp) continue;
        const bool last = (w.jb == jb_hi - 1);
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1u;
        const long long ck0 = clk_on ? clock64() : 0;
        // per-observation terms of this half-chunk, one observation per lane
        if (!TC_DBG(p, 16)) {
          const int k = kc * BK + oh * 32 + lane;
          if (c_next != kc) {  // warp-uniform
            o_next = p.obs[k];
            a_next = p.alpha[k];
          }
          const float yr = (o_next.z - oyh) + (o_next.w - oyl);
          ox[lane] = (o_next.x - oxh) + (o_next.y - oxl);  // (the previous visit's readers passed the __syncwarp below)
          oy[lane] = k < p.n ? ex2_approx(-yr * yr) * p.a_scale : 0.f;
          oa[lane] = a_next;
          c_next = kn >= 0 ? kn : kc;
          const int kp = c_next * BK + oh * 32 + lane;  // < n_cap: the buffers are padded
          o_next = p.obs[kp];
          a_next = p.alpha[kp];
        }
        const long long ckA = clk_on ? clock64() : 0;
        __syncwarp();
        const long long ckB = clk_on ? clock64() : 0;
        uint8_t* pa_hi = a_hi(s) + row_off;
        uint8_t* pa_lo = a_lo(s) + row_off;
        f32x2 mu2 = pack2(0.f, 0.f);
        long long ck1 = 0, ck2 = 0;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t hi[8], lo[8];
          if TC_DBG(p, 1) {
#pragma unroll
            for (int e = 0; e < 8; ++e) { hi[e] = 0x3c003c00u; lo[e] = 0u; }
          } else if (last) gen16<true>(sh->obs_x[gw] + 4 * half, sh->obs_y[gw] + 4 * half, sh->obs_a[gw] + 4 * half, ncx2, hi, lo, mu2);
          else gen16<false>(sh->obs_x[gw] + 4 * half, sh->obs_y[gw] + 4 * half, sh->obs_a[gw] + 4 * half, ncx2, hi, lo, mu2);
          if (half == 0) {
            ck1 = clk_on ? clock64() : 0;
            mbar_wait(&sh->empty[s], ph ^ 1u);  // MMAs that read this stage have completed
            ck2 = clk_on ? clock64() : 0;
          }
          if (!TC_DBG(p, 8)) {
            const uint32_t u0 = (uint32_t)(oh * 4 + half * 2);
            const uint32_t off0 = ((u0 ^ swz) << 4), off1 = (((u0 + 1u) ^ swz) << 4);
            *reinterpret_cast<uint4*>(pa_hi + off0) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(pa_lo + off0) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            *reinterpret_cast<uint4*>(pa_hi + off1) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
            *reinterpret_cast<uint4*>(pa_lo + off1) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
          }
        }
        if (last) {
          float m0, m1;
          unpack2(mu2, m0, m1);
          two_sum_acc(mu_hi, mu_lo, m0 + m1);
        }
        const long long ckD = clk_on ? clock64() : 0;
        if (!TC_DBG(p, 32)) fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
        __syncwarp();
        const long long ckE = clk_on ? clock64() : 0;
        if (lane == 0) {
          if constexpr (CG == 1) mbar_arrive_local(&sh->full[s]);
          else mbar_arrive_cluster(mapa(smem_u32(&sh->full[s]), 0));
        }
        if (clk_on) {
          const long long ck3 = clock64();
          ck_math += ck1 - ck0; ck_wait += ck2 - ck1; ck_post += ck3 - ck2; ++ck_iters;
          ck_stage += ckA - ck0; ck_sync += ckB - ckA; ck_h2 += ckD - ck2; ck_fence += ckE - ckD;
        }
      }
      const double mu_acc = (double)mu_hi + (double)mu_lo;
      // combine the four partial mu (2 groups x 2 observation halves) and write it out
      if (pq > 0) sh->mu_part[tile_par][pq - 1][r] = (float)mu_acc;
      named_bar_sync(1, kGenWarps * 32);
      if (pq == 0 && live) {
        const int col = col0 + r;
        if (col < p.W) {
          const double part = (double)sh->mu_part[tile_par][0][r] + (double)sh->mu_part[tile_par][1][r] +
                              (double)sh->mu_part[tile_par][2][r];
          p.mu[(long long)lrow * p.W + col] = (float)((double)p.prior_mean + (mu_acc + part) * (double)p.mu_unscale);
        }
      }
    }
    if (clk_on) {
      long long* o = p.dbg_clk + 8 * (blockIdx.x * 2 + (gw == 15));
      o[0] = ck_math; o[1] = ck_wait; o[2] = ck_post; o[3] = ck_iters; o[4] = ck_stage; o[5] = ck_sync; o[6] = ck_h2; o[7] = ck_fence;
    }
  }

  tc_fence_before();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  if (SBO_TC_CLOCKS && p.dbg_clk && blockIdx.x == 0 && threadIdx.x == 0) p.dbg_clk[63] = clock64() - ck_begin;
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<CG>(tmem_base, kTmemCols);
  }
}

#include "posterior_tc_pair.cuh"
#include "posterior_tc_inc.cuh"

}  // namespace

// ---- host side -------------------------------------------------------------------------------------
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn g_encode_tiled = nullptr;   // cuTensorMapEncodeTiled, resolved by tc_init

template <int CG> constexpr size_t tc_smem_bytes() {
  return 1024 /*align slack*/ + (size_t)TcCfg<CG>::kStages * TcCfg<CG>::kStageBytes + sizeof(TcShared);
}

}  // namespace

// Power-of-two scales of the fp16 operand planes.  A = K* <= sigma_f^2; every entry of L^-1 is bounded by
// ||L^-1||_2 = lambda_min(K)^-1/2 <= noise_var^-1/2 (K = k(X,X) + noise_var I), so both bounds hold for every
// observation set and survive appends.  Scaled maxima stay <= 2^15 < 65504.
void tc_plane_shifts(const sbo_config& cfg, int* shift_a, int* shift_b) {
  const double sf2 = cfg.sigma_f * cfg.sigma_f;
  const double bmax = 1.0 / std::sqrt(std::max(cfg.noise_var, 1e-30));
  *shift_a = std::max(-60, std::min(60, (int)std::floor(15.0 - std::log2(sf2))));
  *shift_b = std::max(-60, std::min(60, (int)std::floor(15.0 - std::log2(bmax))));
}

// (Re-)encode the tensor map of the fp16 planes: the TMA box is this CTA's share of a 256-row B tile.
int32_t tc_set_cta_group(sbo_ctx* ctx, int cg) {
  if (!g_encode_tiled) return fail(ctx, SBO_ERR_STATE, "tc_set_cta_group: tcgen05 path not initialised");
  CUtensorMap* map = ctx->tmap_planes ? static_cast<CUtensorMap*>(ctx->tmap_planes) : new CUtensorMap;
  const cuuint64_t nc = (cuuint64_t)ctx->n_cap;
  const cuuint64_t gdim[2] = {nc, 2 * nc};                 // inner: k (columns of L^-1), outer: hi rows then lo rows
  const cuuint64_t gstride[1] = {nc * sizeof(__half)};     // bytes between rows
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)(BN / cg)};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = g_encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, ctx->d_planes, gdim, gstride, box, estr,
                                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    if (!ctx->tmap_planes) delete map;
    return fail(ctx, SBO_ERR_CUDA, "tc_set_cta_group: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  }
  ctx->tmap_planes = map;
  ctx->tc_cg = cg;
  return SBO_OK;
}

int32_t tc_init(sbo_ctx* ctx) {
  ctx->tc_ok = false;
  if (!(ctx->cfg.noise_var > 0.0))
    return fail(ctx, SBO_ERR_INVALID_ARG, "tc_init: the tcgen05 kernel needs noise_var > 0 (it bounds the fp16 planes of L^-1)");
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
    return fail(ctx, SBO_ERR_CUDA, "tc_init: cuTensorMapEncodeTiled not available from the driver");
  ctx->tc_skip = true;   // sbo_set_option(SBO_OPT_TC_SKIP, 0): contract every chunk (dense), for A/B measurements
  tc_plane_shifts(ctx->cfg, &ctx->shift_a, &ctx->shift_b);
  g_encode_tiled = reinterpret_cast<EncodeTiledFn>(fn);
  SBO_TRY(tc_set_cta_group(ctx, 2));
  // (per device: sbo_create has made ctx->device current)
  e = cudaFuncSetAttribute(posterior_tc_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<1>());
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(posterior_tc_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<1>());
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(posterior_tc_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<2>());
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(posterior_tc_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<2>());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(posterior_tc_inc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<2>());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(posterior_tc_inc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<2>());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(posterior_tc_inc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<2>());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(posterior_tc_inc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<2>());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(posterior_tc_inc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_smem_bytes<2>());
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(posterior_tc_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_pair_smem_bytes());
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(posterior_tc_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_pair_smem_bytes());
  if (e != cudaSuccess) {
    delete static_cast<CUtensorMap*>(ctx->tmap_planes);
    ctx->tmap_planes = nullptr;
    return cuda_fail(ctx, e, "tc_init: cudaFuncSetAttribute(max dynamic smem)");
  }
  ctx->tc_ok = true;
  return SBO_OK;
}

void tc_destroy(sbo_ctx* ctx) {
  if (ctx->eager_stream) {
    cudaStreamSynchronize(ctx->eager_stream);
    cudaStreamDestroy(ctx->eager_stream);
    cudaEventDestroy(ctx->ev_eager_done);
    ctx->eager_stream = nullptr;
  }
  if (ctx->d_pplanes) cudaFree(ctx->d_pplanes);
  ctx->d_pplanes = nullptr;
  ctx->pplanes_cap = 0;
  if (ctx->d_rs_part) cudaFree(ctx->d_rs_part);
  ctx->d_rs_part = nullptr;
  ctx->rs_cap = 0;
  ctx->eager_inflight = false;
  if (ctx->tmap_planes) delete static_cast<CUtensorMap*>(ctx->tmap_planes);
  ctx->tmap_planes = nullptr;
  ctx->tc_ok = false;
}

static float* g_dbg_v = nullptr;

int32_t tc_pack_observations(sbo_ctx* ctx, cudaStream_t stream) {
  if (!stream) stream = ctx->stream;
  if (!ctx->d_obs_f32) return fail(ctx, SBO_ERR_STATE, "packed observation buffer missing (fp32 mode only)");
  const double xscale = std::sqrt(1.4426950408889634 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale));
  pack_obs_kernel<<<(ctx->n_cap + 255) / 256, 256, 0, stream>>>(ctx->d_xy, ctx->n, ctx->n_cap, xscale,
                                                                     reinterpret_cast<float4*>(ctx->d_obs_f32));
  SBO_LAUNCH_CHECK(ctx, "pack_obs_kernel");
  const int nchunks = ctx->n_cap / BK;
  chunk_bbox_kernel<<<(nchunks * 32 + 255) / 256, 256, 0, stream>>>(reinterpret_cast<const float4*>(ctx->d_obs_f32), ctx->n,
                                                                         nchunks, reinterpret_cast<float4*>(ctx->d_chunk_bbox));
  SBO_LAUNCH_CHECK(ctx, "chunk_bbox_kernel");
  return SBO_OK;
}

// Cut-offs of the block-sparse walk (plan.cu has the derivation).  The measured row 1-norm of L^-1 replaces its a-priori
// bound where the walk is always planned after the fit: beyond kEagerMaxObs observations (below that the eager first
// pairs start inside the fit, before the norm exists, and must walk the same masks as the launch that resumes them).
constexpr int kEagerMaxObs = 2048;
static double tc_r1(const sbo_ctx* ctx) { return ctx->n > kEagerMaxObs ? ctx->linv_r1 : 0.0; }
float tc_skip_bits(const sbo_ctx* ctx) { return (float)tc_skip_bits_of(ctx->cfg, ctx->n, tc_r1(ctx)); }
float tc_ring_bits(const sbo_ctx* ctx) {
  if (!ctx->opt.tc_graded) return tc_skip_bits(ctx);
  return (float)tc_ring_bits_of(ctx->cfg, ctx->n, tc_r1(ctx), ctx->shift_a);
}

// Is there enough to skip on this grid with these observations to pay for the mask walk (~3 % at C2, where
// almost every chunk matters everywhere)?  Host-side estimate on a sample of cell tiles: the share of the dense walk's
// MMA passes that the box test keeps, with the same boxes and cut-offs as the kernel.  Cached until the grid or the
// observations change.
// The mask walk costs 3-5 % where nothing can be left out; it pays once the walk issues < 90 % of the dense walk's MMA passes
// (C2: 92 % of the visits survive but 13 % of the passes go, the far ring running hi.hi only).
constexpr double kSkipWorthIt = 0.90;

bool tc_plan_skip(sbo_ctx* ctx) {
  if (!ctx->tc_skip) return false;
  if (ctx->tc_skip_plan >= 0) return ctx->tc_skip_plan != 0;
  const int n = ctx->n, nchunks = (n + BK - 1) / BK;
  ctx->tc_skip_plan = 0;
  ctx->tc_kept_frac = 1.0;
  ctx->tc_pass_frac = 1.0;
  if (nchunks < 8 || (int)ctx->h_xy_sorted.size() < 2 * n) return false;
  const double xs = tc_xscale(ctx->cfg);
  std::vector<double> box;
  chunk_boxes(ctx->h_xy_sorted.data(), n, xs, box);
  const int rows = ctx->row1 - ctx->row0, tiles_per_row = (ctx->W + BM - 1) / BM;
  const double bits = tc_skip_bits(ctx), ring = tc_ring_bits(ctx);
  const int sr = std::min(rows, 24), sc = std::min(tiles_per_row, 24);
  double kept = 0.0, total = 0.0, passes = 0.0;
  for (int a = 0; a < sr; ++a)
    for (int b = 0; b < sc; ++b) {
      const int row = ctx->row0 + (int)((long long)a * rows / sr), col0 = (int)((long long)b * tiles_per_row / sc) * BM;
      const double tx0 = (ctx->x0 + col0 * ctx->rx) * xs, tx1 = tx0 + (BM - 1) * ctx->rx * xs;
      double t, k, p;
      tile_visit_cost(box, n, std::min(tx0, tx1), std::max(tx0, tx1), (ctx->y0 + row * ctx->ry) * xs, bits, ring, &t, &k, &p);
      total += t; kept += k; passes += p;
    }
  ctx->tc_kept_frac = total > 0.0 ? kept / total : 1.0;
  ctx->tc_pass_frac = total > 0.0 ? passes / (3.0 * total) : 1.0;
  ctx->tc_skip_plan = (total > 0.0 && passes < kSkipWorthIt * 3.0 * total) ? 1 : 0;   // MMA passes, not visits: far-ring visits issue 1 of 3
  return ctx->tc_skip_plan != 0;
}

int32_t tc_plan_detail(sbo_ctx* ctx, double out[6]) {
  const bool skip = ctx->tc_ok && ctx->grid_set ? tc_plan_skip(ctx) : false;
  out[0] = tc_skip_bits(ctx);
  out[1] = skip ? tc_ring_bits(ctx) : out[0];
  out[2] = skip ? ctx->tc_kept_frac : 1.0;
  out[3] = skip ? ctx->tc_pass_frac : 1.0;
  out[4] = tc_r1(ctx);
  const int tpr = (ctx->W + BM - 1) / BM, strip = 2 * ctx->opt.tc_strip;
  out[5] = (strip > 0 && tpr % 2 == 0 && tpr > strip) ? strip : 0;
  return SBO_OK;
}

static void tc_fill_params(sbo_ctx* ctx, TcParams& p, int j_lo, int j_hi, bool accumulate) {
  p.obs = reinterpret_cast<const float4*>(ctx->d_obs_f32);
  p.alpha = ctx->d_alpha_f32;
  p.mu = static_cast<float*>(ctx->d_mu);
  p.s2 = static_cast<float*>(ctx->d_s2);
  p.n = ctx->n;
  p.n_cap = ctx->n_cap;
  p.j_lo = j_hi > j_lo ? j_lo : 0;
  p.j_hi = j_hi > j_lo ? j_hi : 0;
  p.accumulate = accumulate ? 1 : 0;
  p.W = ctx->W;
  p.rows_local = ctx->row1 - ctx->row0;
  p.row0 = ctx->row0;
  p.x0 = ctx->x0; p.y0 = ctx->y0; p.rx = ctx->rx; p.ry = ctx->ry;
  const double sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  p.xscale = std::sqrt(1.4426950408889634 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale));
  p.a_scale = (float)std::ldexp(sf2, ctx->shift_a);
  p.mu_unscale = (float)std::ldexp(1.0, -ctx->shift_a);
  p.v_unscale2 = std::ldexp(1.0, -2 * (ctx->shift_a + ctx->shift_b));
  p.v_unscale = (float)std::ldexp(1.0, -(ctx->shift_a + ctx->shift_b));
  p.sf2 = (float)sf2;
  p.prior_mean = (float)ctx->cfg.prior_mean;
  p.dbg_v = g_dbg_v;
  p.bbox = reinterpret_cast<const float4*>(ctx->d_chunk_bbox);
  p.skip = tc_plan_skip(ctx) ? 1 : 0;
  p.skip_bits = tc_skip_bits(ctx);
  p.ring_bits = p.skip ? tc_ring_bits(ctx) : p.skip_bits;
  // strip-major tile order (pair kernel): needs whole tile pairs per row and per strip
  const int tpr = (ctx->W + BM - 1) / BM;
  const int strip = 2 * ctx->opt.tc_strip;
  p.strip_tiles = (strip > 0 && tpr % 2 == 0 && tpr > strip) ? strip : 0;
  p.pi_lo = 0;
  p.pi_hi = ((ctx->n + BN - 1) / BN + 1) / 2;
  p.rs_mode = 0;
  p.rs_part = nullptr;
  p.exec_cols = ctx->d_exec_cols;
  p.dbg = 0;
  p.dbg_clk = nullptr;
}

// Number of phases of the full refresh (see launch_posterior_tc).  SBO_OPT_TC_PHASES: 0 = by size, k >= 1 = exactly k.
// By size: the kept part of the planes a patch of concurrently processed tiles walks, ~1.4 x kept_fraction x n^2 / 2 entries
// of 4 bytes (hi + lo), in pieces of at most ~24 MB; nothing to gain while the whole of it sits in the L2 (n < 8192).
static int tc_phase_count(sbo_ctx* ctx, int npairs) {
  int f = ctx->opt.tc_phases;
  if (f <= 0) {
    if (ctx->n < 8192) return 1;
    const double bytes = 1.4 * ctx->tc_kept_frac * 0.5 * (double)ctx->n * ctx->n * 4.0;
    f = (int)std::ceil(bytes / (24.0 * 1048576.0));
  }
  return std::max(1, std::min(f, npairs));
}

static cudaError_t tc_launch_pair(sbo_ctx* ctx, const TcParams& p, cudaStream_t stream, long long clusters) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * clusters));
  cfg.blockDim = dim3(kThreadsTc);
  cfg.dynamicSmemBytes = tc_pair_smem_bytes();
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const CUtensorMap& tmap = *static_cast<CUtensorMap*>(ctx->tmap_planes);
  return p.skip ? cudaLaunchKernelEx(&cfg, posterior_tc_pair_kernel<true>, tmap, p)
                : cudaLaunchKernelEx(&cfg, posterior_tc_pair_kernel<false>, tmap, p);
}

// ---- eager first pairs ------------------------------------------------------------------------------------------------
// sbo_set_observations followed by sbo_step is the end-to-end step, and at n ~ 1000 a third of it is the fp64
// factorisation: a chain of small kernels that keeps one or two SMs busy.  Rows 0 .. 512 P - 1 of L^-1 are final when the
// chain is about half way, and the first P row-block pairs of the posterior need nothing else (no alpha: mu is summed in
// the last pair).  So factor_full launches those pairs on a side stream as soon as their rows (and fp16 planes) exist,
// on fewer CTAs than SMs so that the factor kernels keep finding free ones; the launch stores only the running row
// sums (rs_mode 1).  The next full posterior on the same grid and observation set resumes from them (rs_mode 2: the
// same sums in the same order, so sigma^2 and mu are bit-identical to the one-launch result).  Anything else that
// touches the grid or the observations first waits for the side stream and drops the partial result.
int tc_eager_pairs(sbo_ctx* ctx) {
  const bool off = !ctx->opt.tc_eager, pair_off = !ctx->opt.tc_pair;
  if (off || pair_off || !ctx->tc_ok || ctx->tc_cg != 2 || ctx->cfg.precision != SBO_FP32 || !ctx->grid_set || ctx->m_local == 0)
    return 0;
  if (ctx->cfg.kernel != SBO_KERNEL_AUTO && ctx->cfg.kernel != SBO_KERNEL_TCGEN05) return 0;
  if (ctx->n < 2 * BN || ctx->n > kEagerMaxObs || g_dbg_v) return 0;   // beyond that the factorisation is a few % of the step, and the walk is planned with the measured norm of L^-1
  const long long ntiles = (long long)(ctx->row1 - ctx->row0) * ((ctx->W + BM - 1) / BM);
  if (ntiles < 2 * ctx->sm_count) return 0;                    // a grid that small is latency, not throughput
  // How many pairs: as many as finish about when the factorisation does.  While they run the factor kernels share the SMs
  // that were left free, so a partial launch that outlasts the fit only delays it (measured at n = 4096 on a 4096^2 grid: four
  // eager pairs = 50 ms of posterior turned a 7 ms fit into 57 ms).  Estimates: fit ~ 1.1 ms (n / 1024)^1.35 (measured 1.1 / 2.6 /
  // 7.2 ms at 1024 / 2048 / 4096), of which the share 1 - 512 P / n is left when rows < 512 P are final; P pairs execute
  // ~ 3.2 m (512 P)^2 kept flops at ~ 1.3e15 flop/s on 7/8 of the GPU.  Not worth a launch below 5 % of the whole posterior.
  const int npairs = ((ctx->n + BN - 1) / BN + 1) / 2;
  const double kept = tc_plan_skip(ctx) ? ctx->tc_kept_frac : 1.0;
  const double fit_ms = 1.1 * std::pow((double)ctx->n / 1024.0, 1.35);
  const double m = (double)ctx->m_local;
  for (int P = npairs / 2; P >= 1; --P) {
    const double rows = 512.0 * P;
    const double ms = 3.2 * m * rows * rows * kept / 1.3e15 * 1e3;
    const double all_ms = 3.2 * m * (double)ctx->n * ctx->n * kept / 1.5e15 * 1e3;
    const double left_ms = fit_ms * (1.0 - rows / (double)ctx->n);   // the chain advances about linearly in the rows
    if (ms <= 1.3 * left_ms) return ms >= 0.05 * all_ms ? P : 0;
  }
  return 0;
}

void tc_eager_cancel(sbo_ctx* ctx) {
  if (ctx->eager_inflight) {
    cudaStreamSynchronize(ctx->eager_stream);
    ctx->eager_inflight = false;
  }
}

// called by factor_full once rows [0, 2 BN pairs) of L^-1 and their planes are final (event `ready`)
int32_t tc_eager_launch(sbo_ctx* ctx, int pairs, cudaEvent_t ready) {
  if (!ctx->eager_stream) {
    SBO_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->eager_stream, cudaStreamNonBlocking));
    SBO_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_eager_done, cudaEventDisableTiming));
  }
  if (ctx->rs_cap < ctx->m_local) {
    if (ctx->d_rs_part) cudaFree(ctx->d_rs_part);
    ctx->d_rs_part = nullptr;
    ctx->rs_cap = 0;
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_rs_part, sizeof(float2) * (size_t)ctx->m_local));
    ctx->rs_cap = ctx->m_local;
  }
  cudaStream_t se = ctx->eager_stream;
  SBO_CUDA(ctx, cudaStreamWaitEvent(se, ready, 0));
  SBO_TRY(tc_pack_observations(ctx, se));
  TcParams p;
  tc_fill_params(ctx, p, 0, ctx->n, false);
  p.pi_lo = 0;
  p.pi_hi = pairs;
  p.rs_mode = 1;
  p.rs_part = reinterpret_cast<float2*>(ctx->d_rs_part);
  if (ctx->d_exec_cols) SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_exec_cols, 0, sizeof(unsigned long long), se));
  const long long ntiles = (long long)p.rows_local * ((p.W + BM - 1) / BM);
  // leave ~1/8 of the SMs to the factorisation's own kernels
  const long long clusters =
      std::max<long long>(1, std::min<long long>((ntiles + 1) / 2, ctx->sm_count / 2 - std::max(2, ctx->sm_count / 16)));
  const cudaError_t e = tc_launch_pair(ctx, p, se, clusters);
  if (e != cudaSuccess) return cuda_fail(ctx, e, "posterior_tc_pair_kernel (eager first pairs)");
  ++ctx->launches;
  SBO_CUDA(ctx, cudaEventRecord(ctx->ev_eager_done, se));
  ctx->eager_inflight = true;
  ctx->eager_pairs = pairs;
  ctx->eager_n = ctx->n;
  ctx->eager_grid_epoch = ctx->grid_epoch;
  return SBO_OK;
}

// ---- incremental step as a GEMM (posterior_tc_inc.cuh) ---------------------------------------------------------------
constexpr size_t kIncBandBytes = (size_t)1 << 30;   // planes of P are generated and contracted in bands of at most this size

static void tc_inc_shape(const sbo_ctx* ctx, int k, int* lq, int* nslabs, long long* nitems) {
  int l = 0;
  while ((1 << l) < k) ++l;
  const int rows = ctx->row1 - ctx->row0, slab_rows = BN >> l;
  *lq = l;
  *nslabs = (rows + slab_rows - 1) / slab_rows;
  *nitems = (long long)*nslabs * (((ctx->W + BM - 1) / BM + 1) / 2);
}

bool tc_inc_applicable(sbo_ctx* ctx, int k) {
  const bool off = !ctx->opt.tc_inc;
  if (off || !ctx->tc_ok || ctx->tc_cg != 2 || ctx->cfg.precision != SBO_FP32 || !g_encode_tiled) return false;
  if (k < 1 || k > 16 || ctx->m_local == 0 || ctx->n < 1) return false;
  if ((ctx->n + BK - 1) / BK > kMaskWords * 32) return false;   // chunk mask capacity (32 768 observations)
  int lq, nslabs;
  long long nitems;
  tc_inc_shape(ctx, k, &lq, &nslabs, &nitems);
  // Worth it when the GEMM (waves of (slab, column pair) items over the CTA pairs, ~1 us per 64-observation chunk, plus
  // ~100 us of plane generation, launches, pipeline fill and the last epilogue: measured 0.15 ms at C2 with one new row
  // against 0.116 ms for the strip kernel, 0.10 vs 0.19 ms with two) beats the FMA kernels (~1.2e13 cell-observations/s per row).
  const int clusters = ctx->sm_count / 2, nchunks = (ctx->n + BK - 1) / BK;
  const double gemm_us = (double)((nitems + clusters - 1) / clusters) * nchunks * 1.0 + 100.0;
  const double fma_us = (double)ctx->m_local * ctx->n * k / 1.2e13 * 1e6;
  return gemm_us < fma_us;
}

int32_t launch_posterior_tc_inc(sbo_ctx* ctx, int row_lo) {
  const int k = ctx->n - row_lo;
  int lq, nslabs;
  long long nitems;
  tc_inc_shape(ctx, k, &lq, &nslabs, &nitems);
  tc_eager_cancel(ctx);
  const int ld = ctx->n_cap;
  const size_t slab_bytes = (size_t)2 * BN * ld * sizeof(__half);                       // hi + lo rows of one slab
  const int band_slabs = (int)std::max<size_t>(1, std::min<size_t>((size_t)nslabs, kIncBandBytes / slab_bytes));
  const size_t bytes = (size_t)band_slabs * slab_bytes;
  if (ctx->pplanes_cap < bytes) {
    if (ctx->d_pplanes) cudaFree(ctx->d_pplanes);
    ctx->d_pplanes = nullptr;
    ctx->pplanes_cap = 0;
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_pplanes, bytes));
    ctx->pplanes_cap = bytes;
  }
  SBO_TRY(tc_pack_observations(ctx, ctx->stream));
  const double inv2l2 = 1.0 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale);
  IncParams p;
  p.obs = reinterpret_cast<const float4*>(ctx->d_obs_f32);
  p.w = ctx->d_w;
  p.mu = static_cast<float*>(ctx->d_mu);
  p.s2 = static_cast<float*>(ctx->d_s2);
  p.n = ctx->n; p.row_lo = row_lo; p.k = k;
  p.W = ctx->W; p.rows_local = ctx->row1 - ctx->row0;
  p.x0 = ctx->x0; p.rx = ctx->rx;
  p.y0 = ctx->y0; p.ry = ctx->ry; p.row0 = ctx->row0;
  p.bbox = reinterpret_cast<const float4*>(ctx->d_chunk_bbox);   // written by tc_pack_observations above
  p.skip = (ctx->tc_skip && ctx->n >= 8 * BK) ? 1 : 0;           // block-sparse walk of the items (posterior_tc_inc.cuh)
  p.skip_bits = tc_skip_bits(ctx);
  p.xscale = std::sqrt(1.4426950408889634 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale));
  p.a_scale = (float)std::ldexp(ctx->cfg.sigma_f * ctx->cfg.sigma_f, ctx->shift_a);
  p.u_unscale = std::ldexp(1.0, -(ctx->shift_a + ctx->shift_b));
  p.u_unscale_f = (float)p.u_unscale;
  p.tiles_x = (ctx->W + BM - 1) / BM;
  const int ncp = (p.tiles_x + 1) / 2;
  for (int s0 = 0; s0 < nslabs; s0 += band_slabs) {   // bands run back to back on the stream and reuse the planes buffer
    const int bs = std::min(band_slabs, nslabs - s0), rp = bs * BN;
    p.nslabs = bs; p.rp = rp; p.lr_base = (s0 * BN) >> lq;
    pplanes_kernel<<<dim3((unsigned)((rp >> lq) / kPplaneRows), (unsigned)((ld + 255) / 256)), 256, 0, ctx->stream>>>(
        ctx->d_xy, ctx->d_Linv, ld, ctx->n, row_lo, k, lq, p.rows_local, ctx->row0, ctx->y0, ctx->ry, inv2l2,
        std::ldexp(1.0, ctx->shift_b), rp, p.lr_base, reinterpret_cast<__half*>(ctx->d_pplanes), p.bbox, p.skip, p.skip_bits,
        p.xscale);
    SBO_LAUNCH_CHECK(ctx, "pplanes_kernel");
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)ld, (cuuint64_t)2 * rp};
    const cuuint64_t gstride[1] = {(cuuint64_t)ld * sizeof(__half)};
    const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)(BN / 2)};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = g_encode_tiled(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, ctx->d_pplanes, gdim, gstride, box, estr,
                                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
      return fail(ctx, SBO_ERR_CUDA, "launch_posterior_tc_inc: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
    const long long items = (long long)bs * ncp;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * std::min<long long>(items, ctx->sm_count / 2)));
    cfg.blockDim = dim3(kThreadsTc);
    cfg.dynamicSmemBytes = tc_smem_bytes<2>();
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e;
    switch (lq) {
      case 0: e = cudaLaunchKernelEx(&cfg, posterior_tc_inc_kernel<0>, tmap, p); break;
      case 1: e = cudaLaunchKernelEx(&cfg, posterior_tc_inc_kernel<1>, tmap, p); break;
      case 2: e = cudaLaunchKernelEx(&cfg, posterior_tc_inc_kernel<2>, tmap, p); break;
      case 3: e = cudaLaunchKernelEx(&cfg, posterior_tc_inc_kernel<3>, tmap, p); break;
      default: e = cudaLaunchKernelEx(&cfg, posterior_tc_inc_kernel<4>, tmap, p); break;
    }
    if (e != cudaSuccess) return cuda_fail(ctx, e, "posterior_tc_inc_kernel (cluster launch)");
    SBO_LAUNCH_CHECK(ctx, "posterior_tc_inc_kernel");
  }
  return SBO_OK;
}

int32_t launch_posterior_tc(sbo_ctx* ctx, int j_lo, int j_hi, bool accumulate) {
  if (!ctx->tc_ok) return fail(ctx, SBO_ERR_STATE, "tcgen05 kernel not initialised");
  if (ctx->m_local == 0) return SBO_OK;
  // a partial result of the eager first pairs: usable by a full refresh of the same grid and observation set, once
  bool resume = false;
  if (ctx->eager_inflight) {
    resume = !accumulate && j_lo == 0 && j_hi == ctx->n && ctx->eager_n == ctx->n && ctx->eager_grid_epoch == ctx->grid_epoch;
    if (resume) {
      SBO_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_eager_done, 0));  // also orders the re-pack below behind its reads
      ctx->eager_inflight = false;
    } else {
      tc_eager_cancel(ctx);
    }
  }
  TcParams p;
  tc_fill_params(ctx, p, j_lo, j_hi, accumulate);
  if (resume) {
    p.pi_lo = ctx->eager_pairs;
    p.rs_mode = 2;
    p.rs_part = reinterpret_cast<float2*>(ctx->d_rs_part);
  }
  p.dbg = 0;
#if SBO_TC_CLOCKS
  if (const char* s = std::getenv("SBO_TC_DBG")) p.dbg = std::atoi(s);
#endif
  p.dbg_clk = nullptr;
  static long long* d_clk = nullptr;
#if SBO_TC_CLOCKS
  const bool want_clk = std::getenv("SBO_TC_CLK") != nullptr;
#else
  const bool want_clk = false;
#endif
  if (want_clk) {
    if (!d_clk) cudaMalloc(&d_clk, 64 * sizeof(long long));
    cudaMemsetAsync(d_clk, 0, 64 * sizeof(long long), ctx->stream);
    p.dbg_clk = d_clk;
  }
  if ((ctx->n + BK - 1) / BK > kMaskWords * 32)
    return fail(ctx, SBO_ERR_CAPACITY, "tcgen05 posterior kernel: more than 32 768 observations (chunk mask capacity)");
  SBO_TRY(tc_pack_observations(ctx, ctx->stream));
  const long long ntiles = (long long)p.rows_local * ((p.W + BM - 1) / BM);
  const CUtensorMap& tmap = *static_cast<CUtensorMap*>(ctx->tmap_planes);
  if (ctx->tc_cg == 1) {
    const unsigned grid = (unsigned)std::min<long long>(ntiles, ctx->sm_count);  // persistent: one CTA per SM
    if (p.skip) posterior_tc_kernel<1, true><<<grid, kThreadsTc, tc_smem_bytes<1>(), ctx->stream>>>(tmap, p);
    else posterior_tc_kernel<1, false><<<grid, kThreadsTc, tc_smem_bytes<1>(), ctx->stream>>>(tmap, p);
  } else {
    // persistent CTA pairs: one cluster of 2 per TPC
    const long long pairs = std::min<long long>((ntiles + 1) / 2, ctx->sm_count / 2);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * pairs));
    cfg.blockDim = dim3(kThreadsTc);
    cfg.dynamicSmemBytes = tc_smem_bytes<2>();
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // full refresh over at least two row blocks: the K*-stationary pair kernel (SBO_OPT_TC_PAIR = 0: off)
    bool pair = !p.accumulate && p.j_lo == 0 && p.j_hi == p.n && p.n > BN && p.dbg_v == nullptr;
    pair = pair && ctx->opt.tc_pair != 0;
    if (resume && !pair) return fail(ctx, SBO_ERR_STATE, "eager first pairs without the pair kernel");
    if (pair && !resume && ctx->d_exec_cols)   // executed-work counter of this launch (a resumed launch adds to the eager part's count)
      SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_exec_cols, 0, sizeof(unsigned long long), ctx->stream));
    // Phased walk (large n, block-sparse plan): the row-block pairs are contracted in `phases` launches over consecutive
    // pair ranges, the running row sums travelling through rs_part (8 B per cell and phase) exactly as between the eager
    // first pairs and their resuming launch -- same additions in the same order, bit-identical.  Why: the CTA pairs of ONE
    // launch drift apart along their tiles' 32-pair walks, so together they keep re-reading the whole kept part of the
    // 1 GiB planes (~70-100 MB per patch of tiles at C4) through a 126 MB L2: 58 % hit rate, 1.6 TB/s of DRAM reads that
    // the power-limited tensor pipe pays for in clock.  Inside a phase the B rows in flight are 1 / phases of that.
    int phases = 1;
    if (pair && !resume && p.skip) phases = tc_phase_count(ctx, p.pi_hi);
    cudaError_t e = cudaSuccess;
    if (phases > 1) {
      if (ctx->rs_cap < ctx->m_local) {
        if (ctx->d_rs_part) cudaFree(ctx->d_rs_part);
        ctx->d_rs_part = nullptr;
        ctx->rs_cap = 0;
        SBO_CUDA(ctx, cudaMalloc(&ctx->d_rs_part, sizeof(float2) * (size_t)ctx->m_local));
        ctx->rs_cap = ctx->m_local;
      }
      p.rs_part = reinterpret_cast<float2*>(ctx->d_rs_part);
      const int np_all = p.pi_hi;
      for (int f = 0; f < phases && e == cudaSuccess; ++f) {
        // equal pair counts: beyond the first few row blocks every row of L^-1 a tile walks has the same kept width (the
        // chunks within reach of the tile), so the bytes in flight go with the number of rows
        p.pi_lo = (int)((long long)np_all * f / phases);
        p.pi_hi = (int)((long long)np_all * (f + 1) / phases);
        if (p.pi_hi <= p.pi_lo) continue;
        p.rs_mode = (p.pi_lo > 0 ? 2 : 0) | (p.pi_hi < np_all ? 1 : 0);
        e = tc_launch_pair(ctx, p, ctx->stream, pairs);
        if (e == cudaSuccess && f + 1 < phases) ++ctx->launches;
      }
    } else {
      e = pair     ? tc_launch_pair(ctx, p, ctx->stream, pairs)
          : p.skip ? cudaLaunchKernelEx(&cfg, posterior_tc_kernel<2, true>, tmap, p)
                   : cudaLaunchKernelEx(&cfg, posterior_tc_kernel<2, false>, tmap, p);
    }
    if (e != cudaSuccess) return cuda_fail(ctx, e, "posterior_tc_kernel<2> (cluster launch)");
  }
  SBO_LAUNCH_CHECK(ctx, "posterior_tc_kernel");
  if (want_clk) {
    long long h[64];
    cudaMemcpyAsync(h, d_clk, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream);
    cudaStreamSynchronize(ctx->stream);
    std::fprintf(stderr, "[tc clk] total %lld clk\n", h[63]);
    for (int g = 0; g < 4; ++g)
      std::fprintf(stderr, "[tc clk] cta %d gen warp %2d: iters %lld  math %.0f (stage %.0f sync %.0f)  wait(empty) %.0f  post %.0f (sts1+gen2+sts2 %.0f fence %.0f) clk/iter\n", g / 2, (g & 1) * 15,
                   h[8 * g + 3], (double)h[8 * g] / std::max(1LL, h[8 * g + 3]), (double)h[8 * g + 4] / std::max(1LL, h[8 * g + 3]),
                   (double)h[8 * g + 5] / std::max(1LL, h[8 * g + 3]), (double)h[8 * g + 1] / std::max(1LL, h[8 * g + 3]),
                   (double)h[8 * g + 2] / std::max(1LL, h[8 * g + 3]), (double)h[8 * g + 6] / std::max(1LL, h[8 * g + 3]),
                   (double)h[8 * g + 7] / std::max(1LL, h[8 * g + 3]));
    std::fprintf(stderr, "[tc clk] issuer: iters %lld  wait(full) %.0f  issue %.0f clk/iter, wait(tmem_empty) total %lld\n", h[43],
                 (double)h[40] / std::max(1LL, h[43]), (double)h[41] / std::max(1LL, h[43]), h[42]);
  }
  return SBO_OK;
}

}  // namespace sbo

// Undocumented debug hook (not part of include/sbo.h): V = K* L^-T of cell tile 0 as the tensor core
// produced it, [128][n_cap] floats, for accuracy forensics in scripts/.
extern "C" int32_t sbo_debug_tc_v(sbo_ctx* ctx, float* host_dst) {
  if (!ctx || !host_dst || !ctx->tc_ok || !ctx->grid_set) return SBO_ERR_INVALID_ARG;
  const size_t count = (size_t)128 * ctx->n_cap;
  float* d = nullptr;
  if (cudaMalloc(&d, count * sizeof(float)) != cudaSuccess) return SBO_ERR_CUDA;
  cudaMemsetAsync(d, 0, count * sizeof(float), ctx->stream);
  sbo::g_dbg_v = d;
  const int32_t st = sbo::launch_posterior_tc(ctx, 0, ctx->n, false);
  sbo::g_dbg_v = nullptr;
  cudaMemcpyAsync(host_dst, d, count * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
  cudaStreamSynchronize(ctx->stream);
  cudaFree(d);
  return st;
}
