// posterior_tc_inc.cuh -- included by posterior_tc.cu (inside namespace sbo::{anonymous}).
//
// The incremental step (k new rows of L^-1 after a rank-k append) as a tensor-core GEMM.  On a grid the RBF factorises,
//   k*(cell(r, c), j) = sigma_f^2 ex(c, j) ey(r, j),
// so the projection of new row q onto a cell is
//   u_q(r, c) = sum_j ex(c, j) * [sigma_f^2 ey(r, j) Linv[q][j]] = (A P^T)(c, (r, q)),   A = ex (W x n),  P = (H k) x n,
// a dense contraction over the observations whose operands cost (W + H k) n exponentials, not m n.  Both updates follow
// from u:  sigma^2 -= sum_q u_q^2,  mu += sum_q w_q u_q  (w = L^-1 (y - prior_mean), see posterior_thin.cu).
//
// Mapping onto the pipeline of posterior_tc_kernel<2>:  the generators produce A tiles (128 columns x 64 observations of
// ex, fp16 hi/lo) exactly as they produce K*, with the row factor left out; P takes the place of L^-1 -- its fp16 hi/lo
// planes are written by pplanes_kernel once per step (row rho = local_row * KP + q, KP = k rounded up to a power of two)
// and arrive by TMA; one accumulator holds 128 columns x 256 rows of P = 256 / KP grid rows; the epilogue applies the
// two updates per element instead of summing squares.  No triangular structure: every chunk at full N.
// 3 x 2 m n KP executed flops: C4 with one new row ~5 ms instead of the FMA strip kernel's 60.

struct IncParams {
  const float4* obs;   // [n_cap] scaled observation coordinates (x_hi, x_lo, y_hi, y_lo)
  const double* w;     // [n] L^-1 (y - prior_mean); entries row_lo .. row_lo + k - 1 are read
  float* mu;           // [m_local]
  float* s2;           // [m_local]
  int n, row_lo, k;
  int W, rows_local;
  double x0, rx, xscale;
  float a_scale;       // sigma_f^2 2^shift_a
  double u_unscale;    // 2^-(shift_a + shift_b)
  float u_unscale_f;   // ... as fp32 (exact: a power of two)
  int nslabs;          // slabs of 256 rows of P
  int tiles_x;         // 128-column tiles per grid row
  int rp;              // rows of P (multiple of 256); the lo plane starts at row rp of the tensor map
  int lr_base;         // first (local) grid row of this band: P is generated and contracted band by band to bound its size
  // block-sparse walk: an item (256 grid columns x 256 / KP grid rows) only visits the 64-observation chunks within reach
  double y0, ry;       // grid rows in the caller's coordinates (x0, rx above)
  int row0;            // first grid row of this shard
  const float4* bbox;  // [nchunks] scaled chunk boxes (chunk_bbox_kernel)
  int skip;            // 0: every chunk
  float skip_bits;     // chunks whose K* entries are all < 2^-skip_bits sigma_f^2 for the whole item are left out
};

// Chunk mask of one work item (every role of both CTAs derives the same bits: one compiled copy, as build_chunk_mask).
// u_q = sum_j ex ey Linv[q][j]: a left-out term is < 2^-bits sigma_f^2 |Linv[q][j]|, so |du| < 2^-bits sigma_f^2 R1 and
// sigma^2 -= u^2 moves by < 2 sigma_f du -- with the full refresh's cut-off (36 bits at C4) 4e-10 sigma_f^2 per step,
// 1e-7 over the 256 steps between two refreshes.
__device__ __noinline__ void build_item_mask(uint32_t* mask, const IncParams p, int col_first, int col_last, int lr_first,
                                             int lr_last, int nchunks) {
  const int lane = threadIdx.x & 31;
  const double xa = cell_coord(p.x0, p.rx, col_first) * p.xscale, xb = cell_coord(p.x0, p.rx, col_last) * p.xscale;
  const double ya = cell_coord(p.y0, p.ry, p.row0 + lr_first) * p.xscale, yb = cell_coord(p.y0, p.ry, p.row0 + lr_last) * p.xscale;
  const float xlo = (float)fmin(xa, xb) - 1e-3f, xhi = (float)fmax(xa, xb) + 1e-3f;   // margin: the box test runs in fp32
  const float ylo = (float)fmin(ya, yb) - 1e-3f, yhi = (float)fmax(ya, yb) + 1e-3f;
  __syncwarp();  // the previous item's walks of this warp's mask are over
#pragma unroll 1
  for (int w = 0; w < kMaskWords; ++w) {
    const int c = w * 32 + lane;
    bool on = c < nchunks;
    if (on && p.skip) {
      const float4 b = p.bbox[c];
      const float dx = fmaxf(fmaxf(b.x - xhi, xlo - b.y), 0.f), dy = fmaxf(fmaxf(b.z - yhi, ylo - b.w), 0.f);
      on = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) <= p.skip_bits + 0.5f;
    }
    const uint32_t bits = __ballot_sync(0xffffffffu, on);
    if (lane == 0) mask[w] = bits;
  }
  __syncwarp();
}

// fp16 hi/lo planes of P: row rho = lr * KP + q holds 2^shift_b ey(lr, j) Linv[row_lo + q][j]  (sigma_f^2 rides on A).
// One thread per (grid row, observation): the fp64 exp is computed once and serves the KP rows of that grid row.
// Block-sparse walk: a 64-observation chunk that is out of reach in y of a whole slab of 256 rows of P is visited by none of
// the slab's items (build_item_mask adds dx^2 >= 0 to the same dy^2, in the same fp32 arithmetic), so its entries are
// neither computed nor written -- the buffer keeps whatever it held there and nobody reads it.  At C4 (reach 7 length-
// scales of a 45 length-scale domain) that is ~60 % of the fp64 exponentials and of the 512 MiB of stores.
// One thread per observation and 16 grid rows (one thread per (row, observation) spent most of its 0.5 ms starting threads).
constexpr int kPplaneRows = 16;   // grid rows per CTA of pplanes_kernel: divides the 256 >> lq rows of a slab (lq <= 4)
__global__ void __launch_bounds__(256)
pplanes_kernel(const double* __restrict__ xy, const double* __restrict__ Linv, int ld, int n, int row_lo, int k, int lq,
               int rows_local, int row0, double y0, double ry, double inv2l2, double scale, int rp, int lr_base,
               __half* __restrict__ planes, const float4* __restrict__ bbox, int skip, float skip_bits, double xscale) {
  const int j = blockIdx.y * 256 + threadIdx.x;
  const int lb0 = blockIdx.x * kPplaneRows;   // first row of the band this CTA writes, incl. the padding up to a whole slab
  if (j >= ld) return;
  if (skip && j < n) {
    const int sb = (lb0 << lq) / BN;   // the slab these rows of the band belong to, and the grid rows that slab spans
    const int lr_first = min(lr_base + ((sb * BN) >> lq), rows_local - 1);
    const int lr_last = min(lr_base + (((sb + 1) * BN) >> lq) - 1, rows_local - 1);
    const double ya = cell_coord(y0, ry, row0 + lr_first) * xscale, yb = cell_coord(y0, ry, row0 + lr_last) * xscale;
    const float ylo = (float)fmin(ya, yb) - 1e-3f, yhi = (float)fmax(ya, yb) + 1e-3f;
    const float4 b = bbox[j / BK];
    const float dy = fmaxf(fmaxf(b.z - yhi, ylo - b.w), 0.f);
    if (__fmul_rn(dy, dy) > skip_bits + 0.5f) return;
  }
  // one thread = one observation: its y and its entries of the new rows of L^-1 are read once for the CTA's grid rows.
  // The row factor follows a recurrence along the grid rows (pitch ry, dy(g + 1) = dy(g) - ry):
  //   e(g + 1) = e(g) s(g),  s(g) = exp((2 dy(g) ry - ry^2) / 2l^2),  s(g + 1) = s(g) exp(-2 ry^2 / 2l^2),
  // anchored with two fp64 exp at every GLOBAL grid row that is a multiple of 16 -- so the value of a row does not depend
  // on where a band of rows starts (row-band shards and P bands give the same bits) -- instead of one exp per row:
  // <= 15 steps, ~40 roundings, far below the 2^-22 of the fp16 pair it is rounded to.
  const int kp = 1 << lq;
  const double yj = j < n ? xy[2 * (size_t)j + 1] : 0.0;
  double lq_[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) lq_[q] = (q < k && q < kp && j < n) ? Linv[(size_t)(row_lo + q) * ld + j] : 0.0;
  const double q2 = exp(-2.0 * ry * ry * inv2l2);
  double e = 0.0, st = 0.0;
  auto anchor = [&](int grow) {
    const double dy = yj - cell_coord(y0, ry, grow);
    e = exp(-dy * dy * inv2l2) * scale;
    st = e != 0.0 ? exp((2.0 * dy * ry - ry * ry) * inv2l2) : 0.0;   // (a huge ratio must not meet an underflowed anchor)
  };
  const int gr0 = row0 + lr_base + lb0;
  anchor(gr0 & ~15);
  for (int g = gr0 & ~15; g < gr0; ++g) { e *= st; st *= q2; }
#pragma unroll 2
  for (int r = 0; r < kPplaneRows; ++r) {
    const int lb = lb0 + r, lr = lr_base + lb;
    if (r > 0 && ((gr0 + r) & 15) == 0) anchor(gr0 + r);
    const double er = (j < n && lr < rows_local) ? e : 0.0;
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      if (q < kp) {
        const double x = (q < k && er != 0.0) ? er * lq_[q] : 0.0;
        const __half hi = __float2half_rn((float)x);
        const __half lo = __float2half_rn((float)(x - (double)__half2float(hi)));
        const size_t rho = (size_t)lb * kp + q;
        planes[rho * ld + j] = hi;
        planes[((size_t)rp + rho) * ld + j] = lo;
      }
    }
    e *= st;
    st *= q2;
  }
}

constexpr int kPrefetchGroups = 6;   // epilogue: L2 prefetch distance in groups of 16 accumulator columns
__device__ __forceinline__ void prefetch_l2(const void* ptr) { asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr)); }

template <int LQ>
__global__ void __launch_bounds__(kThreadsTc, 1)
posterior_tc_inc_kernel(const __grid_constant__ CUtensorMap tmap, IncParams p) {
  constexpr int CG = 2;
  constexpr int KP = 1 << LQ;
  using Cfg = TcCfg<CG>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  TcShared* sh = reinterpret_cast<TcShared*>(tiles + (size_t)kStages * Cfg::kStageBytes);
  auto a_hi = [&](int s) { return tiles + (size_t)s * Cfg::kStageBytes; };
  auto a_lo = [&](int s) { return tiles + (size_t)s * Cfg::kStageBytes + kABytes; };
  auto b_hi = [&](int s) { return tiles + (size_t)s * Cfg::kStageBytes + 2 * kABytes; };
  auto b_lo = [&](int s) { return tiles + (size_t)s * Cfg::kStageBytes + 2 * kABytes + Cfg::kBBytes; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int ncp = (p.tiles_x + 1) / 2;                       // column-tile pairs per slab
  const long long nitems = (long long)p.nslabs * ncp;        // item = slab * ncp + column pair: neighbours share the slab's P tiles
  const long long group = blockIdx.x / CG, ngroups = gridDim.x / CG;
  const int nchunks = (p.n + BK - 1) / BK;
  uint32_t* cmask = sh->cmask[warp];
  auto item_mask = [&](long long item) {   // the chunks this item visits (all of them when p.skip == 0)
    const int sb = (int)(item / ncp), cp = (int)(item % ncp);
    const int lr_first = min(p.lr_base + ((sb * BN) >> LQ), p.rows_local - 1);
    const int lr_last = min(p.lr_base + (((sb + 1) * BN) >> LQ) - 1, p.rows_local - 1);
    build_item_mask(cmask, p, 2 * cp * BM, min(2 * cp * BM + 2 * BM, p.W) - 1, lr_first, lr_last, nchunks);
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&sh->full[s], 1 + (kGenWarps / 2) * CG);
      mbar_init(&sh->empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sh->tmem_full[s], 1);
      mbar_init(&sh->tmem_empty[s], 4 * CG);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<CG>(&sh->tmem_base, kTmemCols);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;

  if (warp == 0) {
    // =============================== TMA producer: P tiles ===============================
    uint32_t it = 0;
    for (long long item = group; item < nitems; item += ngroups) {
      const int sb = (int)(item / ncp);
      const int row = sb * BN + (int)rank * (int)Cfg::kBRows;
      item_mask(item);
      ChunkIter<true> ci;
      ci.reset(cmask);
      for (int kc = ci.next(); kc < nchunks; kc = ci.next(), ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1u;
        mbar_wait(&sh->empty[s], ph ^ 1u);
        const uint32_t bar = mapa(smem_u32(&sh->full[s]), 0);
        if (elect_one()) {
          if (rank == 0) mbar_arrive_expect_tx(&sh->full[s], 2 * Cfg::kBBytes * CG);
          tma_load_2d<CG>(&tmap, bar, b_hi(s), kc * BK, row);
          tma_load_2d<CG>(&tmap, bar, b_lo(s), kc * BK, p.rp + row);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer (leader CTA) ===============================
    uint32_t it = 0, acc_it = 0;
    for (long long item = group; rank == 0 && item < nitems; item += ngroups) {
      item_mask(item);
      ChunkIter<true> ci;
      ci.reset(cmask);
      int kc = ci.next();
      if (kc >= nchunks) continue;   // nothing within reach: u = 0, the epilogue skips the item too
      const uint32_t as = acc_it & 1u, aph = (acc_it >> 1) & 1u;
      ++acc_it;
      mbar_wait(&sh->tmem_empty[as], aph ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      const uint32_t idesc = idesc_f16<CG>(BN);
      bool first = true;
      for (int kn; kc < nchunks; kc = kn, ++it, first = false) {
        kn = ci.next();   // the next visited chunk: tells whether this visit is the item's last
        const int s = it % kStages;
        mbar_wait(&sh->full[s], (it / kStages) & 1u);
        tc_fence_after();
        const uint64_t dah = umma_desc_sw128(smem_u32(a_hi(s))), dal = umma_desc_sw128(smem_u32(a_lo(s)));
        const uint64_t dbh = umma_desc_sw128(smem_u32(b_hi(s))), dbl = umma_desc_sw128(smem_u32(b_lo(s)));
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < BK / UK; ++ks) {
            const uint64_t adv = (uint64_t)((ks * UK * 2) >> 4);
            umma_f16<CG>(d_tmem, dal + adv, dbh + adv, idesc, (first && ks == 0) ? 0u : 1u);
            umma_f16<CG>(d_tmem, dah + adv, dbl + adv, idesc, 1u);
            umma_f16<CG>(d_tmem, dah + adv, dbh + adv, idesc, 1u);
          }
          umma_commit<CG>(&sh->empty[s]);
          if (kn >= nchunks) umma_commit<CG>(&sh->tmem_full[as]);
        }
        __syncwarp();
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // =============================== epilogue: sigma^2 -= sum_q u_q^2, mu += sum_q w_q u_q ===============================
    const int q4 = warp & 3;
    const int c = q4 * 32 + lane;  // TMEM lane == column of the tile
    float wq[KP];
#pragma unroll
    for (int q = 0; q < KP; ++q) wq[q] = q < p.k ? (float)(p.w[p.row_lo + q] * p.u_unscale) : 0.f;
    const float uu = p.u_unscale_f;
    const long long row_bytes = (long long)p.W * (long long)sizeof(float);
    const long long mu_minus_s2 = reinterpret_cast<const char*>(p.mu) - reinterpret_cast<const char*>(p.s2);
    uint32_t acc_it = 0;
    for (long long item = group; item < nitems; item += ngroups) {
      const int sb = (int)(item / ncp), ct = 2 * (int)(item % ncp) + (int)rank;
      const int col = ct * BM + c;
      const bool live = ct < p.tiles_x && col < p.W;
      {
        item_mask(item);
        ChunkIter<true> ci;
        ci.reset(cmask);
        if (ci.next() >= nchunks) continue;   // no chunk within reach of this item: nothing was accumulated, nothing changes
      }
      const uint32_t as = acc_it & 1u, aph = (acc_it >> 1) & 1u;
      ++acc_it;
      mbar_wait(&sh->tmem_full[as], aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q4 * 32) << 16) + as * BN;
#pragma unroll 1
      for (int g = 0; g < BN / 16; ++g) {
        uint32_t v[16];
        tmem_ld16(taddr + g * 16, v);
        tmem_ld_wait();
        const int lr0 = p.lr_base + ((sb * BN + g * 16) >> LQ);  // first grid row of this group of 16 rows of P
        // The read-modify-write below waits on its loads (long-scoreboard stalls were 40 % of the epilogue's samples): the lines
        // of the group kPrefetchGroups ahead (this item's, or the first ones of this CTA pair's next item) are pulled into the
        // L2 now.  On its own this changed nothing (the epilogue was issue-latency-bound, see below); it is kept because the
        // slimmed epilogue has little else to hide a DRAM round trip behind.
        {
          int pg = g + kPrefetchGroups;
          long long pit = item;
          if (pg >= BN / 16) { pg -= BN / 16; pit += ngroups; }
          if (pit < nitems) {
            const int psb = (int)(pit / ncp), pct = 2 * (int)(pit % ncp) + (int)rank;
            const int pcol = pct * BM + c;
            const int plr0 = p.lr_base + ((psb * BN + pg * 16) >> LQ);
            if (pct < p.tiles_x && pcol < p.W && plr0 + (16 >> LQ) <= p.rows_local) {   // (ragged ends go without)
              const char* ps = reinterpret_cast<const char*>(p.s2 + (long long)plr0 * p.W + pcol);
#pragma unroll
              for (int r = 0; r < (16 >> LQ); ++r) {
                prefetch_l2(ps);
                prefetch_l2(ps + mu_minus_s2);
                ps += row_bytes;
              }
            }
          }
        }
        // The stored sigma^2 / mu of the group's grid rows are fetched before the first of them is rewritten: written as one
        // read-modify-write per row the loads cannot pass the stores (mu and s2 may alias as far as the compiler knows)
        // and a thread's 256 rows of an item become 256 dependent DRAM round trips.  The arithmetic is fp32: one warp per
        // scheduler does this work, and with fp64 conversions and adds it issued one instruction per 12 cycles (ncu r02p/q:
        // 56 instructions per row, fp64-pipe and scoreboard stalls) -- the increment's own rounding (2^-24 of |w u|, of
        // u^2) is below that of the fp32 value it is added to.
        constexpr int kRows = 16 >> LQ;                  // grid rows in this group of 16 TMEM columns (LQ <= 4)
        const int nrows = min(kRows, p.rows_local - lr0);   // (< kRows only in the last slab of the shard)
        if (live && nrows == kRows) {   // every row of the group exists: no predicates, addresses by increment
          constexpr int kSub = kRows < 8 ? kRows : 8;   // rows whose loads are in flight together (16 spill)
          char* base = reinterpret_cast<char*>(p.s2 + (long long)lr0 * p.W + col);
#pragma unroll
          for (int r0 = 0; r0 < kRows; r0 += kSub) {
            float s2v[kSub], muv[kSub];
            {
              const char* ps = base;
#pragma unroll
              for (int r = 0; r < kSub; ++r) {
                s2v[r] = *reinterpret_cast<const float*>(ps);
                muv[r] = *reinterpret_cast<const float*>(ps + mu_minus_s2);
                ps += row_bytes;
              }
            }
#pragma unroll
            for (int r = 0; r < kSub; ++r) {
              float usq = 0.f, dmu = 0.f;
#pragma unroll
              for (int q = 0; q < KP; ++q) {
                const float d = __uint_as_float(v[((r0 + r) << LQ) + q]);
                const float u = d * uu;
                usq = fmaf(u, u, usq);
                dmu = fmaf(wq[q], d, dmu);
              }
              *reinterpret_cast<float*>(base) = s2v[r] - usq;
              *reinterpret_cast<float*>(base + mu_minus_s2) = muv[r] + dmu;
              base += row_bytes;
            }
          }
        } else if (live) {
          float* s2p = p.s2 + (long long)lr0 * p.W + col;
          float* mup = p.mu + (long long)lr0 * p.W + col;
#pragma unroll
          for (int r = 0; r < kRows; ++r) {   // (unrolled: a dynamic index would put the accumulator slice in local memory)
            float usq = 0.f, dmu = 0.f;
#pragma unroll
            for (int q = 0; q < KP; ++q) {
              const float d = __uint_as_float(v[(r << LQ) + q]);
              const float u = d * uu;
              usq = fmaf(u, u, usq);
              dmu = fmaf(wq[q], d, dmu);
            }
            if (r < nrows) {
              s2p[(long long)r * p.W] -= usq;
              mup[(long long)r * p.W] += dmu;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa(smem_u32(&sh->tmem_empty[as]), 0));
    }
  } else if (warp >= 8) {
    // =============================== generators: A = ex(column, observation), the K* generator without the row factor ===============================
    const int gw = warp - 8;
    const int grp = gw >> 3;
    const int oh = (gw >> 2) & 1;
    const int r = (gw & 3) * 32 + lane;
    float* ox = reinterpret_cast<float*>(sh->obs_x[gw]);
    float* oy = reinterpret_cast<float*>(sh->obs_y[gw]);
    const uint32_t row_off = (uint32_t)r * 128u;
    const uint32_t swz = (uint32_t)(r & 7);
    const float cxs = (float)((double)r * p.rx * p.xscale);
    const f32x2 ncx2 = pack2(-cxs, -cxs);
    uint32_t it = 0;
    float4 o_next = make_float4(0.f, 0.f, 0.f, 0.f);
    int c_next = -1;
    for (long long item = group; item < nitems; item += ngroups) {
      const int ct = min(2 * (int)(item % ncp) + (int)rank, p.tiles_x - 1);  // an odd last column tile: the peer regenerates it
      const double ox0 = cell_coord(p.x0, p.rx, ct * BM) * p.xscale;
      const float oxh = (float)ox0, oxl = (float)(ox0 - (double)oxh);
      item_mask(item);
      ChunkIter<true> ci, scout;
      ci.reset(cmask);
      scout.reset(cmask);
      scout.next();
      scout.next();
      for (int kc = ci.next(); kc < nchunks; kc = ci.next(), ++it) {
        const int kn = scout.next();   // the chunk two visits ahead = this warp's next visit inside the item
        if ((it & 1u) != (uint32_t)grp) continue;
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1u;
        {
          const int kobs = kc * BK + oh * 32 + lane;
          if (c_next != kc) o_next = p.obs[kobs];   // (first visits of an item: the look-ahead does not cross items)
          ox[lane] = (o_next.x - oxh) + (o_next.y - oxl);
          oy[lane] = kobs < p.n ? p.a_scale : 0.f;
          c_next = kn < nchunks ? kn : -1;
          if (c_next >= 0) o_next = p.obs[c_next * BK + oh * 32 + lane];
        }
        __syncwarp();
        uint8_t* pa_hi = a_hi(s) + row_off;
        uint8_t* pa_lo = a_lo(s) + row_off;
        f32x2 mu2 = pack2(0.f, 0.f);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t hi[8], lo[8];
          gen16<false>(sh->obs_x[gw] + 4 * half, sh->obs_y[gw] + 4 * half, sh->obs_a[gw] + 4 * half, ncx2, hi, lo, mu2);
          if (half == 0) mbar_wait(&sh->empty[s], ph ^ 1u);
          const uint32_t u0 = (uint32_t)(oh * 4 + half * 2);
          const uint32_t off0 = ((u0 ^ swz) << 4), off1 = (((u0 + 1u) ^ swz) << 4);
          *reinterpret_cast<uint4*>(pa_hi + off0) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<uint4*>(pa_lo + off0) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          *reinterpret_cast<uint4*>(pa_hi + off1) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
          *reinterpret_cast<uint4*>(pa_lo + off1) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa(smem_u32(&sh->full[s]), 0));
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<CG>(tmem_base, kTmemCols);
  }
}
