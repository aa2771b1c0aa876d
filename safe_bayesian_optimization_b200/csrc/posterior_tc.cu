// posterior_tc.cu -- the tcgen05/TMEM GP-posterior kernel (fp32 mode).
//
//   V = K* L^-T   (m x n) . (n x n, lower triangular),   sigma^2 = sigma_f^2 - rowsum(V o V)
//   mu = prior_mean + K* alpha
//
// K* (RBF cross-covariance between every cell and every observation) is never materialised:
// generator warps compute 128 x 32 tiles of it on the fly, split every entry into two TF32-exact
// planes (hi = rna_tf32(x), lo = rna_tf32(x - hi)) and store them straight into the 128B-swizzled
// K-major shared-memory layout the UMMA descriptors expect.  The matching 256 x 32 tiles of the
// pre-split hi/lo planes of L^-1 arrive by TMA.  One elected thread issues
//   D += A_lo.B_hi ; D += A_hi.B_lo ; D += A_hi.B_hi        (tcgen05.mma kind::tf32, M128 N256 K8)
// so the fp32 accumulator in TMEM carries ~2^-22 relative products ("3xTF32").  Epilogue warps
// pull finished 128 x 256 accumulators out of TMEM with tcgen05.ld, square and row-sum them in fp64
// and keep only one number per cell: V never reaches shared or global memory.  Two accumulator
// stages (2 x 256 TMEM columns) let the epilogue of row block jb overlap the MMAs of jb+1.
// L^-1 is lower triangular, so row block jb only walks k < 256 (jb+1).
//
// Warp roles (512 threads, 1 CTA/SM, persistent over cell tiles):
//   0      TMA producer (one lane)          1  MMA issuer (one lane)      2  TMEM alloc/dealloc
//   4-7    epilogue (TMEM lane quarter = warp % 4)
//   8-11   K* generators for even k-chunks  12-15  K* generators for odd k-chunks
#include <cuda.h>

#include <algorithm>
#include <cmath>

#include "common.cuh"

namespace sbo {
namespace {

constexpr int BM = 128;          // cells per tile (UMMA M)
constexpr int BN = 256;          // rows of L^-1 per accumulator (UMMA N)
constexpr int BK = 32;           // k per stage: 32 fp32 = one 128-byte swizzle row
constexpr int UK = 8;            // k per tcgen05.mma kind::tf32
constexpr int kStages = 2;
constexpr int kThreadsTc = 512;
constexpr uint32_t kABytes = BM * BK * 4;   // 16 KB per plane
constexpr uint32_t kBBytes = BN * BK * 4;   // 32 KB per plane
constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytes;  // 96 KB
constexpr uint32_t kTmemCols = 512;

struct __align__(16) TcShared {
  unsigned long long full[kStages];
  unsigned long long empty[kStages];
  unsigned long long tmem_full[2];
  unsigned long long tmem_empty[2];
  uint32_t tmem_base;
  uint32_t pad[3];
  float4 scratch[8][32];        // per generator warp: (x_rel, y-term, alpha, -) of the 32 observations of a chunk
  float mu_part[2][BM];         // odd-chunk group's partial mu, double-buffered by tile parity
};
constexpr size_t kSmemBytes = 1024 /*align slack*/ + (size_t)kStages * kStageBytes + sizeof(TcShared);

struct TcParams {
  const double* xy;        // [n][2] fp64 observation coordinates
  const float* alpha;      // [n_cap] fp32
  float* mu;               // [m_local]
  float* s2;               // [m_local]
  int n, n_cap;
  int j_lo, j_hi;          // rows of L^-1 that count (incremental step: only the new rows)
  int accumulate;
  int W, rows_local, row0; // this shard: grid rows [row0, row0 + rows_local)
  double x0, y0, rx, ry;
  float frx;               // (float) rx
  float neg_c2;            // -log2(e) / (2 l^2)
  float log2_sf2;          // log2(sigma_f^2)
  float sf2, prior_mean;
  float* dbg_v;            // debug: V of tile 0 ([128][n_cap]) when non-null
};

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
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

__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, unsigned long long* bar, void* dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}

// D[tmem] (+)= A[smem] . B[smem]^T, kind::tf32, issued by one thread for the CTA.
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrive when every tcgen05.mma issued so far by this thread has completed.
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
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
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory descriptor: K-major operand, 128-byte swizzle, rows 128 B apart, 8-row groups
// 1024 B apart (SBO), descriptor version 1 (sm_100).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor: D=f32, A=B=tf32, both K-major, N=256, M=128.
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ float tf32_rna_bits(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
#ifndef SBO_TC_TERMS
#define SBO_TC_TERMS 3 // 3: lo.hi + hi.lo + hi.hi   4: + lo.lo   1: hi.hi only (diagnostics)
#endif

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// number of k-chunks of row block jb
__device__ __forceinline__ int chunks_of(int jb, int n_pad32) { return min(n_pad32, (jb + 1) * BN) / BK; }

// ---- the kernel ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreadsTc, 1)
posterior_tc_kernel(const __grid_constant__ CUtensorMap tmap, TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  // 128B-swizzled tiles need 1024-byte alignment; pad by pointer arithmetic so the compiler keeps
  // the shared address space (STS/LDS, not generic stores).
  uint8_t* tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  TcShared* sh = reinterpret_cast<TcShared*>(tiles + (size_t)kStages * kStageBytes);
  // stage s: [A_hi 16K][A_lo 16K][B_hi 32K][B_lo 32K]
  auto a_hi = [&](int s) { return tiles + (size_t)s * kStageBytes; };
  auto a_lo = [&](int s) { return tiles + (size_t)s * kStageBytes + kABytes; };
  auto b_hi = [&](int s) { return tiles + (size_t)s * kStageBytes + 2 * kABytes; };
  auto b_lo = [&](int s) { return tiles + (size_t)s * kStageBytes + 2 * kABytes + kBBytes; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_per_row = (p.W + BM - 1) / BM;
  const long long ntiles = (long long)p.rows_local * tiles_per_row;
  const int n_pad32 = (p.n + BK - 1) / BK * BK;
  const int jb_lo = p.j_lo / BN, jb_hi = (p.j_hi + BN - 1) / BN;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&sh->full[s], 1 + 4);  // TMA producer + one lane of each of the 4 generator warps of the group
      mbar_init(&sh->empty[s], 1);     // tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sh->tmem_full[s], 1);   // tcgen05.commit
      mbar_init(&sh->tmem_empty[s], 4);  // one lane of each epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(&sh->tmem_base, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;

  if (warp == 0) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      uint32_t it = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int jb = jb_lo; jb < jb_hi; ++jb) {
          const int nch = chunks_of(jb, n_pad32);
          for (int kc = 0; kc < nch; ++kc, ++it) {
            const int s = it % kStages;
            const uint32_t ph = (it / kStages) & 1u;
            mbar_wait(&sh->empty[s], ph ^ 1u);
            mbar_arrive_expect_tx(&sh->full[s], 2 * kBBytes);
            tma_load_2d(&tmap, &sh->full[s], b_hi(s), kc * BK, jb * BN);
            tma_load_2d(&tmap, &sh->full[s], b_lo(s), kc * BK, p.n_cap + jb * BN);
          }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      uint32_t it = 0, acc_it = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int jb = jb_lo; jb < jb_hi; ++jb, ++acc_it) {
          const uint32_t as = acc_it & 1u, aph = (acc_it >> 1) & 1u;
          mbar_wait(&sh->tmem_empty[as], aph ^ 1u);  // epilogue has drained this accumulator
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * BN;
          const int nch = chunks_of(jb, n_pad32);
          for (int kc = 0; kc < nch; ++kc, ++it) {
            const int s = it % kStages;
            const uint32_t ph = (it / kStages) & 1u;
            mbar_wait(&sh->full[s], ph);  // TMA bytes landed and generators stored their planes
            tc_fence_after();
            const uint64_t dah = umma_desc_sw128(smem_u32(a_hi(s))), dal = umma_desc_sw128(smem_u32(a_lo(s)));
            const uint64_t dbh = umma_desc_sw128(smem_u32(b_hi(s))), dbl = umma_desc_sw128(smem_u32(b_lo(s)));
#pragma unroll
            for (int k8 = 0; k8 < BK / UK; ++k8) {
              const uint64_t adv = (uint64_t)((k8 * UK * 4) >> 4);  // +32 bytes along K inside the swizzle row
#if SBO_TC_TERMS == 1
              umma_tf32(d_tmem, dah + adv, dbh + adv, kIdesc, (kc | k8) != 0);
#else
#if SBO_TC_TERMS == 4
              umma_tf32(d_tmem, dal + adv, dbl + adv, kIdesc, (kc | k8) != 0);
              umma_tf32(d_tmem, dal + adv, dbh + adv, kIdesc, 1u);
#else
              umma_tf32(d_tmem, dal + adv, dbh + adv, kIdesc, (kc | k8) != 0);
#endif
              umma_tf32(d_tmem, dah + adv, dbl + adv, kIdesc, 1u);
              umma_tf32(d_tmem, dah + adv, dbh + adv, kIdesc, 1u);
#endif
            }
            umma_commit(&sh->empty[s]);  // stage reusable once these MMAs have read it
          }
          umma_commit(&sh->tmem_full[as]);  // accumulator complete
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // =============================== epilogue ===============================
    const int q = warp & 3;
    const int r = q * 32 + lane;  // TMEM lane == row of the tile == cell
    uint32_t acc_it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      double rsum = 0.0;
      for (int jb = jb_lo; jb < jb_hi; ++jb, ++acc_it) {
        const uint32_t as = acc_it & 1u, aph = (acc_it >> 1) & 1u;
        mbar_wait(&sh->tmem_full[as], aph);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
        const int j0 = jb * BN;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          const int jc = j0 + c * 32;
          if (jc >= p.j_hi || jc + 32 <= p.j_lo) continue;  // columns outside the contracted rows (warp-uniform)
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_ld_wait();
          if (p.dbg_v && tile == 0) {
#pragma unroll
            for (int i = 0; i < 32; ++i) p.dbg_v[(size_t)r * p.n_cap + jc + i] = __uint_as_float(v[i]);
          }
          const bool whole = (jc >= p.j_lo) && (jc + 32 <= p.j_hi);
          double s0 = 0.0, s1 = 0.0;
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            double x0 = (double)__uint_as_float(v[i]), x1 = (double)__uint_as_float(v[i + 1]);
            if (!whole) {
              if (jc + i < p.j_lo || jc + i >= p.j_hi) x0 = 0.0;
              if (jc + i + 1 < p.j_lo || jc + i + 1 >= p.j_hi) x1 = 0.0;
            }
            s0 = fma(x0, x0, s0);
            s1 = fma(x1, x1, s1);
          }
          rsum += s0 + s1;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh->tmem_empty[as]);
      }
      const int lrow = (int)(tile / tiles_per_row), col = (int)(tile % tiles_per_row) * BM + r;
      if (col < p.W) {
        const long long lc = (long long)lrow * p.W + col;
        const double prev = p.accumulate ? (double)p.s2[lc] : (double)p.sf2;
        p.s2[lc] = (float)(prev - rsum);
      }
    }
  } else if (warp >= 8) {
    // =============================== K* generators ===============================
    const int gw = warp - 8;       // 0..7
    const int grp = gw >> 2;       // 0: even chunks, 1: odd chunks
    const int r = (gw & 3) * 32 + lane;
    float4* scratch = sh->scratch[gw];
    const uint32_t row_off = (uint32_t)r * 128u;
    const uint32_t swz = (uint32_t)(r & 7);
    uint32_t it = 0;
    uint32_t tile_par = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, tile_par ^= 1u) {
      const int lrow = (int)(tile / tiles_per_row), col0 = (int)(tile % tiles_per_row) * BM;
      const double ox = cell_coord(p.x0, p.rx, col0), oy = cell_coord(p.y0, p.ry, p.row0 + lrow);
      const float cxr = (float)r * p.frx;  // offset of this thread's cell from the tile origin
      double mu_acc = 0.0;  // fp32 inside a 32-observation chunk, fp64 across chunks
      for (int jb = jb_lo; jb < jb_hi; ++jb) {
        const bool last = (jb == jb_hi - 1);
        const int nch = chunks_of(jb, n_pad32);
        for (int kc = 0; kc < nch; ++kc, ++it) {
          if ((it & 1u) != (uint32_t)grp) continue;
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          // per-observation terms of this chunk, one observation per lane
          {
            const int k = kc * BK + lane;
            float4 t = make_float4(0.f, -INFINITY, 0.f, 0.f);
            if (k < p.n) {
              const double2 o = *reinterpret_cast<const double2*>(p.xy + 2 * (size_t)k);
              const float xr = (float)(o.x - ox), yr = (float)(o.y - oy);
              t.x = xr;
              t.y = fmaf(yr * yr, p.neg_c2, p.log2_sf2);
              t.z = p.alpha[k];
            }
            __syncwarp();  // previous chunk's readers are done with the scratch
            scratch[lane] = t;
            __syncwarp();
          }
          mbar_wait(&sh->empty[s], ph ^ 1u);  // MMAs that read this stage have completed
          float mu_chunk = 0.f;
          uint8_t* pa_hi = a_hi(s) + row_off;
          uint8_t* pa_lo = a_lo(s) + row_off;
#pragma unroll
          for (int qd = 0; qd < BK / 4; ++qd) {
            float hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float4 t = scratch[qd * 4 + e];  // broadcast
              const float dx = t.x - cxr;
              // MUFU ex2 is accurate enough here: swapping it for a 1-ulp polynomial changed neither mu nor
              // sigma^2 errors (profiles/accuracy_r01.md); the error budget is set by the accumulator.
              const float kv = ex2_approx(fmaf(dx * dx, p.neg_c2, t.y));
              mu_chunk = fmaf(kv, t.z, mu_chunk);
              hi[e] = tf32_rna_bits(kv);
              lo[e] = tf32_rna_bits(kv - hi[e]);
            }
            const uint32_t off = ((uint32_t)qd ^ swz) << 4;
            *reinterpret_cast<float4*>(pa_hi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<float4*>(pa_lo + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
          }
          if (last) mu_acc += (double)mu_chunk;
          fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
          __syncwarp();
          if (lane == 0) mbar_arrive(&sh->full[s]);
        }
      }
      // combine the two groups' partial mu and write it out
      if (grp == 1) sh->mu_part[tile_par][r] = (float)mu_acc;
      named_bar_sync(1, 256);
      if (grp == 0) {
        const int col = col0 + r;
        if (col < p.W) p.mu[(long long)lrow * p.W + col] = (float)((double)p.prior_mean + mu_acc + (double)sh->mu_part[tile_par][r]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace

// ---- host side -------------------------------------------------------------------------------------
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

}  // namespace

int32_t tc_init(sbo_ctx* ctx) {
  ctx->tc_ok = false;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
    return fail(ctx, SBO_ERR_CUDA, "tc_init: cuTensorMapEncodeTiled not available from the driver");
  CUtensorMap* map = new CUtensorMap;
  const cuuint64_t nc = (cuuint64_t)ctx->n_cap;
  const cuuint64_t gdim[2] = {nc, 2 * nc};                 // inner: k (columns of L^-1), outer: hi rows then lo rows
  const cuuint64_t gstride[1] = {nc * sizeof(float)};      // bytes between rows
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)BN};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = reinterpret_cast<EncodeTiledFn>(fn)(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ctx->d_planes, gdim, gstride,
                                                         box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    delete map;
    return fail(ctx, SBO_ERR_CUDA, "tc_init: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  }
  e = cudaFuncSetAttribute(posterior_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
  if (e != cudaSuccess) {
    delete map;
    return cuda_fail(ctx, e, "tc_init: cudaFuncSetAttribute(max dynamic smem)");
  }
  ctx->tmap_planes = map;
  ctx->tc_ok = true;
  return SBO_OK;
}

void tc_destroy(sbo_ctx* ctx) {
  if (ctx->tmap_planes) delete static_cast<CUtensorMap*>(ctx->tmap_planes);
  ctx->tmap_planes = nullptr;
  ctx->tc_ok = false;
}

static float* g_dbg_v = nullptr;

int32_t launch_posterior_tc(sbo_ctx* ctx, int j_lo, int j_hi, bool accumulate) {
  if (!ctx->tc_ok) return fail(ctx, SBO_ERR_STATE, "tcgen05 kernel not initialised");
  if (ctx->m_local == 0) return SBO_OK;
  TcParams p;
  p.xy = ctx->d_xy;
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
  p.frx = (float)ctx->rx;
  const double sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  p.neg_c2 = (float)(-1.4426950408889634 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale));
  p.log2_sf2 = (float)std::log2(sf2);
  p.sf2 = (float)sf2;
  p.prior_mean = (float)ctx->cfg.prior_mean;
  p.dbg_v = g_dbg_v;
  const long long ntiles = (long long)p.rows_local * ((p.W + BM - 1) / BM);
  const unsigned grid = (unsigned)std::min<long long>(ntiles, ctx->sm_count);  // persistent: one CTA per SM
  posterior_tc_kernel<<<grid, kThreadsTc, kSmemBytes, ctx->stream>>>(*static_cast<CUtensorMap*>(ctx->tmap_planes), p);
  SBO_LAUNCH_CHECK(ctx, "posterior_tc_kernel");
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
