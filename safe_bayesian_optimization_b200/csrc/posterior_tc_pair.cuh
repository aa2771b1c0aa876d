// posterior_tc_pair.cuh -- included by posterior_tc.cu (inside namespace sbo::{anonymous}).
//
// The "K*-stationary" form of the tcgen05 posterior kernel for the full refresh (all rows of L^-1, CTA pairs; dense or
// block-sparse walk: kSkip as in posterior_tc_kernel, one chunk mask per tile serves both row blocks of a pair).  A generated K* chunk (128 cells x 64 observations, fp16 hi/lo planes) is
// contracted against TWO row blocks of L^-1 -- 512 accumulator columns, the whole of TMEM -- before its stage is
// released, so a tile visits chunk k once per PAIR of row blocks: 24 generations per tile at n = 1024 instead of 40,
// half as many at large n.  The generators (issue slots, MUFU, shared-memory stores) were what paced the
// one-row-block kernel once the MMA issue path was fixed (profiles/ncu_posterior_tc_r01d.md).
//
// Price: no spare accumulator, so the epilogue of a pair's second row block is not hidden behind MMAs of the same
// tile (the first one's is: the pair's last four visits only feed the second row block).  A and B travel in separate
// rings: A stages (generators -> full_a, commit -> empty_a) and B stages (TMA -> full_b, commit -> empty_b), one
// B stage per (visit, row block).
//
// Row-block pair pi = row blocks 2 pi and 2 pi + 1 (the last pair of an odd count has one).  Visit kc of a pair:
//   row block jb0 = 2 pi     takes part while kc < chunks_of(jb0)       (narrowed in its diagonal block)
//   row block jb1 = 2 pi + 1 takes part in every visit                  (kc < chunks_of(jb1))

#ifndef SBO_PAIR_SA
#define SBO_PAIR_SA 3
#define SBO_PAIR_SB 3
#endif
constexpr int kPairStagesA = SBO_PAIR_SA;
constexpr int kPairStagesB = SBO_PAIR_SB;
constexpr uint32_t kPairABytes = 2 * kABytes;              // hi + lo planes of one K* chunk: 32 KB
constexpr uint32_t kPairBBytes = 2 * TcCfg<2>::kBBytes;    // hi + lo half-tiles of L^-1 held by one CTA: 32 KB

struct __align__(16) TcPairShared {
  unsigned long long full_a[4];    // leader CTA's copy is the one in use
  unsigned long long empty_a[4];   // local (tcgen05.commit multicasts to both CTAs)
  unsigned long long full_b[4];    // leader CTA's copy is the one in use
  unsigned long long empty_b[4];   // local
  unsigned long long tmem_full[2]; // local
  unsigned long long tmem_empty[2];// leader CTA's copy is the one in use
  uint32_t tmem_base;
  uint32_t pad[3];
  float4 obs_x[kGenWarps][8];
  float4 obs_y[kGenWarps][8];
  float4 obs_a[kGenWarps][8];
  float mu_part[2][3][BM];
  uint32_t cmask[kThreadsTc / 32][2 * kMaskWords];  // per warp: chunk masks of the current tile pair (contracted / all three products), see build_chunk_mask
};

constexpr size_t tc_pair_smem_bytes() {
  return 1024 + (size_t)kPairStagesA * kPairABytes + (size_t)kPairStagesB * kPairBBytes + sizeof(TcPairShared);
}

// visits of one tile in pipeline order: for every pair, the contracted chunks below chunks_of(last row block of the pair)
template <bool kSkip> struct PairWalk {
  ChunkIter<kSkip> ci;
  const uint32_t* mask;
  int pi, npairs, nblocks, n_pad, nch;
  __device__ __forceinline__ int nch_of(int pair) const { return chunks_of(min(2 * pair + 1, nblocks - 1), n_pad); }
  // pairs [pi_lo, pi_hi): `npairs` is the END of the walk, not the pair count of the observation set
  __device__ __forceinline__ void start(const uint32_t* m, int nblocks_, int n_pad_, int pi_lo, int pi_hi) {
    mask = m; nblocks = nblocks_; n_pad = n_pad_; npairs = pi_hi; pi = pi_lo; nch = pi < npairs ? nch_of(pi) : 0;
    ci.reset(mask);
  }
  // chunk of the next visit (it belongs to pair `pi`), or -1 when the tile is done
  __device__ __forceinline__ int next() {
    while (pi < npairs) {
      const int kc = ci.next();
      if (kc < nch) return kc;
      if (++pi < npairs) {
        nch = nch_of(pi);
        ci.reset(mask);
      }
    }
    return -1;
  }
};

template <bool kSkip>
__global__ void __launch_bounds__(kThreadsTc, 1)
posterior_tc_pair_kernel(const __grid_constant__ CUtensorMap tmap, TcParams p) {
  constexpr int CG = 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* tiles_b = tiles + (size_t)kPairStagesA * kPairABytes;
  TcPairShared* sh = reinterpret_cast<TcPairShared*>(tiles_b + (size_t)kPairStagesB * kPairBBytes);
  auto a_hi = [&](int s) { return tiles + (size_t)s * kPairABytes; };
  auto a_lo = [&](int s) { return tiles + (size_t)s * kPairABytes + kABytes; };
  auto b_hi = [&](int s) { return tiles_b + (size_t)s * kPairBBytes; };
  auto b_lo = [&](int s) { return tiles_b + (size_t)s * kPairBBytes + TcCfg<2>::kBBytes; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int tiles_per_row = (p.W + BM - 1) / BM;
  const long long ntiles = (long long)p.rows_local * tiles_per_row;
  const long long group = blockIdx.x / CG, ngroups = gridDim.x / CG;
  const long long ngtiles = (ntiles + CG - 1) / CG;
  const int n_pad = (p.n + BK - 1) / BK * BK;
  const int nblocks = (p.n + BN - 1) / BN;
  const int npairs = (nblocks + 1) / 2;
  const int nchunks = n_pad / BK;
  uint32_t* cmask = sh->cmask[warp];
  const TileMap tmap_tiles = {tiles_per_row, p.rows_local, p.strip_tiles};
  const MaskGeom geom = {p.x0, p.y0, p.rx, p.ry, p.xscale, p.bbox, p.row0, p.skip, p.skip_bits, p.ring_bits, tmap_tiles};

  if (threadIdx.x == 0) {
    for (int s = 0; s < 4; ++s) {
      mbar_init(&sh->full_a[s], (kGenWarps / 2) * CG);  // the 8 generator warps per CTA that own the visit
      mbar_init(&sh->empty_a[s], 1);                    // tcgen05.commit
      mbar_init(&sh->full_b[s], 1);                     // leader's TMA producer (+ transaction bytes of both CTAs)
      mbar_init(&sh->empty_b[s], 1);                    // tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sh->tmem_full[s], 1);        // tcgen05.commit
      mbar_init(&sh->tmem_empty[s], 4 * CG);  // every epilogue warp of the pair
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<CG>(&sh->tmem_base, kTmemCols);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;

  if (warp == 0) {
    // =============================== TMA producer: B stages in consumption order ===============================
    uint32_t itb = 0;
    auto load_b = [&](int jb, int kc, bool fresh, bool full3) {
      const int s = itb % kPairStagesB;
      const uint32_t ph = (itb / kPairStagesB) & 1u;
      // diagonal block: rows < 64 c hold zeros, the tile starts below them.  An accumulator's first contracted chunk
      // runs at full N (it initialises every column), so its tile starts at the row block's first row.
      const int c = fresh ? 0 : max(kc - jb * (BN / BK), 0);
      const int row = jb * BN + c * BK + (int)rank * ((int)TcCfg<2>::kBRows - c * (BK / 2));
      mbar_wait(&sh->empty_b[s], ph ^ 1u);
      const uint32_t bar = mapa(smem_u32(&sh->full_b[s]), 0);
      if (elect_one()) {
        if (rank == 0) mbar_arrive_expect_tx(&sh->full_b[s], (full3 ? kPairBBytes : TcCfg<2>::kBBytes) * CG);
        tma_load_2d<CG>(&tmap, bar, b_hi(s), kc * BK, row);
        if (full3) tma_load_2d<CG>(&tmap, bar, b_lo(s), kc * BK, p.n_cap + row);   // (hi.hi only: the lo plane is not read)
      }
      __syncwarp();
      ++itb;
    };
    for (long long gt = group; gt < ngtiles; gt += ngroups) {
      if (kSkip) build_chunk_mask<CG>(cmask, geom, gt, ntiles, nchunks);
      for (int pi = p.pi_lo; pi < p.pi_hi; ++pi) {
        const int jb0 = 2 * pi;
        const bool has1 = jb0 + 1 < nblocks;
        const int nch0 = chunks_of(jb0, n_pad), nchl = has1 ? chunks_of(jb0 + 1, n_pad) : nch0;
        bool fresh0 = true, fresh1 = true;
        ChunkIter<kSkip> ci;
        ci.reset(cmask);
        for (int kc = ci.next(); kc < nchl; kc = ci.next()) {
          const bool full3 = !kSkip || chunk_full(cmask, kc);
          if (kc < nch0) { load_b(jb0, kc, fresh0, full3); fresh0 = false; }
          if (has1) { load_b(jb0 + 1, kc, fresh1, full3); fresh1 = false; }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer (leader CTA; converged warp, one elected lane issues) ===============================
    uint32_t ita = 0, itb = 0, use0 = 0, use1 = 0;
    unsigned long long issued_cols = 0;   // sum over the MMAs issued of their N (each is M 256 x N x K 16): the executed tensor work
    for (long long gt = group; rank == 0 && gt < ngtiles; gt += ngroups) {
      if (kSkip) build_chunk_mask<CG>(cmask, geom, gt, ntiles, nchunks);
      for (int pi = p.pi_lo; pi < p.pi_hi; ++pi) {
        const int jb0 = 2 * pi;
        const bool has1 = jb0 + 1 < nblocks;
        const int nch0 = chunks_of(jb0, n_pad), nchl = has1 ? chunks_of(jb0 + 1, n_pad) : nch0;
        ChunkIter<kSkip> ci;
        ci.reset(cmask);
        int kc = ci.next();
        bool fresh0 = true, fresh1 = true;
        for (int kn; kc < nchl; kc = kn, ++ita) {
          kn = ci.next();  // the next contracted chunk: tells whether this visit is an accumulator's last
          const bool full3 = !kSkip || chunk_full(cmask, kc);   // graded split depth: far chunks run hi.hi only
          const int sa = ita % kPairStagesA;
          mbar_wait(&sh->full_a[sa], (ita / kPairStagesA) & 1u);  // generators stored their planes
          tc_fence_after();
          const uint64_t dah = umma_desc_sw128(smem_u32(a_hi(sa))), dal = umma_desc_sw128(smem_u32(a_lo(sa)));
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            if (half == 0 ? kc >= nch0 : !has1) continue;
            const int jb = jb0 + half;
            const bool fresh = half == 0 ? fresh0 : fresh1;
            if (fresh) {  // first MMA into this accumulator: the epilogue must have drained its previous contents
              mbar_wait(&sh->tmem_empty[half], ((half == 0 ? use0 : use1) & 1u) ^ 1u);
              tc_fence_after();
              if (half == 0) { ++use0; fresh0 = false; } else { ++use1; fresh1 = false; }
            }
            const int sb = itb % kPairStagesB;
            mbar_wait(&sh->full_b[sb], (itb / kPairStagesB) & 1u);  // TMA bytes of both CTAs landed
            tc_fence_after();
            const uint64_t dbh = umma_desc_sw128(smem_u32(b_hi(sb))), dbl = umma_desc_sw128(smem_u32(b_lo(sb)));
            // diagonal block, chunk c: columns < 64 c of the accumulator get no contribution (zeros of L^-1)
            const uint32_t c = fresh ? 0u : (uint32_t)max(kc - jb * (BN / BK), 0);
            const uint32_t idesc = idesc_f16<CG>(BN - c * BK), d_col = tmem_base + half * BN + c * BK;
            issued_cols += (unsigned long long)((BK / UK) * (full3 ? 3 : 1)) * (BN - c * BK);
            if (elect_one()) {
#pragma unroll
              for (int ks = 0; ks < BK / UK; ++ks) {
                const uint64_t adv = (uint64_t)((ks * UK * 2) >> 4);
                if (full3) {
                  umma_f16<CG>(d_col, dal + adv, dbh + adv, idesc, (fresh && ks == 0) ? 0u : 1u);
                  umma_f16<CG>(d_col, dah + adv, dbl + adv, idesc, 1u);
                  umma_f16<CG>(d_col, dah + adv, dbh + adv, idesc, 1u);
                } else {
                  umma_f16<CG>(d_col, dah + adv, dbh + adv, idesc, (fresh && ks == 0) ? 0u : 1u);
                }
              }
              umma_commit<CG>(&sh->empty_b[sb]);
              if (half == 1 || !has1) umma_commit<CG>(&sh->empty_a[sa]);  // last use: K* stage reusable once these MMAs have read it
              if (half == 0 && kn >= nch0) umma_commit<CG>(&sh->tmem_full[0]);  // first row block complete
              if (half == 1 && kn >= nchl) umma_commit<CG>(&sh->tmem_full[1]);
            }
            __syncwarp();
            ++itb;
          }
        }
      }
    }
    if (p.exec_cols && rank == 0 && lane == 0 && issued_cols) atomicAdd(p.exec_cols, issued_cols);
  } else if (warp >= 4 && warp < 8) {
    // =============================== epilogue ===============================
    const int q = warp & 3;
    const int r = q * 32 + lane;  // TMEM lane == row of the tile == cell
    uint32_t use0 = 0, use1 = 0;
    for (long long gt = group; gt < ngtiles; gt += ngroups) {
      const long long tile = gt * CG + rank;
      if (kSkip) build_chunk_mask<CG>(cmask, geom, gt, ntiles, nchunks);
      float rs_hi = 0.f, rs_lo = 0.f;  // sum of squares of this cell's row of V (scaled), double-word fp32
      const long long e_tile = min(tile, ntiles - 1);
      int e_lrow, e_col0;
      tmap_tiles.locate(e_tile, e_lrow, e_col0);
      const int e_col = e_col0 + r;
      const bool e_live = tile < ntiles && e_col < p.W;
      const long long e_lc = (long long)e_lrow * p.W + e_col;
      if ((p.rs_mode & 2) && e_live) {  // resume: pairs < pi_lo were contracted by an earlier launch (same sums, same order)
        const float2 v0 = p.rs_part[e_lc];
        rs_hi = v0.x; rs_lo = v0.y;
      }
      ChunkIter<kSkip> ci0;
      ci0.reset(cmask);
      const int first_chunk = ci0.next();
      for (int pi = p.pi_lo; pi < p.pi_hi; ++pi) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int jb = 2 * pi + half;
          if (jb >= nblocks) continue;
          if (first_chunk >= chunks_of(jb, n_pad)) continue;  // row block skipped by the issuer: V = 0
          const uint32_t use = half == 0 ? use0++ : use1++;
          mbar_wait(&sh->tmem_full[half], use & 1u);
          tc_fence_after();
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + half * BN;
          const int j0 = jb * BN;
#pragma unroll 1
          for (int c = 0; c < BN / 32; ++c) {
            const int jc = j0 + c * 32;
            if (jc >= p.n) continue;  // rows of L^-1 beyond n are zero (warp-uniform)
            uint32_t v[32];
            tmem_ld32(taddr + c * 32, v);
            tmem_ld_wait();
            if (jc + 32 > p.n) {  // the group that holds row n - 1 (warp-uniform): rows >= n are zero by invariant, masked all the same
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (jc + i >= p.n) v[i] = 0u;
            }
            // same arithmetic as posterior_tc_kernel's epilogue: exact squares, four 8-term fp32 partial sums, 2Sum
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
          if (lane == 0) mbar_arrive_cluster(mapa(smem_u32(&sh->tmem_empty[half]), 0));
        }
      }
      if (e_live) {
        if (p.rs_mode & 1) p.rs_part[e_lc] = make_float2(rs_hi, rs_lo);  // partial launch: the running sum, nothing else
        else p.s2[e_lc] = (float)((double)p.sf2 - ((double)rs_hi + (double)rs_lo) * p.v_unscale2);
      }
    }
  } else if (warp >= 8) {
    // =============================== K* generators (as in posterior_tc_kernel, one generation per pair visit) ===============================
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
    uint32_t ita = 0;
    uint32_t tile_par = 0;
    float4 o_next = make_float4(0.f, 0.f, 0.f, 0.f);
    float a_next = 0.f;
    int c_next = -1;
    for (long long gt = group; gt < ngtiles; gt += ngroups, tile_par ^= 1u) {
      const long long tile = min(gt * CG + (long long)rank, ntiles - 1);  // an odd last tile: the peer recomputes it, no store
      const bool live = gt * CG + rank < ntiles;
      int lrow, col0;
      tmap_tiles.locate(tile, lrow, col0);
      const double ox0 = cell_coord(p.x0, p.rx, col0) * p.xscale, oy0 = cell_coord(p.y0, p.ry, p.row0 + lrow) * p.xscale;
      const float oxh = (float)ox0, oxl = (float)(ox0 - (double)oxh);
      const float oyh = (float)oy0, oyl = (float)(oy0 - (double)oyh);
      float mu_hi = 0.f, mu_lo = 0.f;
      if (kSkip) build_chunk_mask<CG>(cmask, geom, gt, ntiles, nchunks);
      PairWalk<kSkip> w, scout;
      w.start(cmask, nblocks, n_pad, p.pi_lo, p.pi_hi);
      scout = w;
      const int first0 = scout.next(), first1 = scout.next();
      // Visit parity (which group of warps generates a visit, hence how the observations are grouped in the partial sums
      // of mu) restarts with every pair: it must depend neither on the tiles this CTA processed before (row-band shards)
      // nor on how many visits the pairs before had, which a launch that resumes at pair pi_lo does not see (block-sparse
      // walk: any count).  The look-ahead below then guesses wrong across a pair of odd length: checked at use.
      uint32_t vi = 0;
      int pi_seen = -1;
      for (int kc = w.next(); kc >= 0; kc = w.next(), ++ita, ++vi) {
        if (w.pi != pi_seen) { pi_seen = w.pi; vi = 0; }
        int kn = scout.next();  // the chunk two visits ahead
        if (kn < 0) kn = grp == 0 ? first0 : first1;
        if ((vi & 1u) != (uint32_t)grp) continue;
        const bool last = (w.pi == npairs - 1);  // the last pair walks every chunk: mu is summed there
        const bool full3 = !kSkip || chunk_full(cmask, kc);
        const int s = ita % kPairStagesA;
        const uint32_t ph = (ita / kPairStagesA) & 1u;
        {
          const int k = kc * BK + oh * 32 + lane;
          if (c_next != kc) {  // warp-uniform
            o_next = p.obs[k];
            a_next = p.alpha[k];
          }
          const float yr = (o_next.z - oyh) + (o_next.w - oyl);
          ox[lane] = (o_next.x - oxh) + (o_next.y - oxl);
          oy[lane] = k < p.n ? ex2_approx(-yr * yr) * p.a_scale : 0.f;
          oa[lane] = a_next;
          c_next = kn >= 0 ? kn : kc;
          const int kp = c_next * BK + oh * 32 + lane;  // < n_cap: the buffers are padded
          o_next = p.obs[kp];
          a_next = p.alpha[kp];
        }
        __syncwarp();
        uint8_t* pa_hi = a_hi(s) + row_off;
        uint8_t* pa_lo = a_lo(s) + row_off;
        f32x2 mu2 = pack2(0.f, 0.f);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t hi[8], lo[8];
          if (last) gen16<true>(sh->obs_x[gw] + 4 * half, sh->obs_y[gw] + 4 * half, sh->obs_a[gw] + 4 * half, ncx2, hi, lo, mu2);
          else gen16<false>(sh->obs_x[gw] + 4 * half, sh->obs_y[gw] + 4 * half, sh->obs_a[gw] + 4 * half, ncx2, hi, lo, mu2);
          if (half == 0) mbar_wait(&sh->empty_a[s], ph ^ 1u);  // MMAs that read this stage have completed
          const uint32_t u0 = (uint32_t)(oh * 4 + half * 2);
          const uint32_t off0 = ((u0 ^ swz) << 4), off1 = (((u0 + 1u) ^ swz) << 4);
          *reinterpret_cast<uint4*>(pa_hi + off0) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<uint4*>(pa_hi + off1) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
          if (full3) {   // (hi.hi only: the lo plane of this stage is not read)
            *reinterpret_cast<uint4*>(pa_lo + off0) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            *reinterpret_cast<uint4*>(pa_lo + off1) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
          }
        }
        if (last) {
          float m0, m1;
          unpack2(mu2, m0, m1);
          two_sum_acc(mu_hi, mu_lo, m0 + m1);
        }
        fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa(smem_u32(&sh->full_a[s]), 0));
      }
      const double mu_acc = (double)mu_hi + (double)mu_lo;
      if (pq > 0) sh->mu_part[tile_par][pq - 1][r] = (float)mu_acc;
      named_bar_sync(1, kGenWarps * 32);
      if (pq == 0 && live && p.pi_hi == npairs) {  // (a partial launch that stops before the last pair has no mu)
        const int col = col0 + r;
        if (col < p.W) {
          const double part = (double)sh->mu_part[tile_par][0][r] + (double)sh->mu_part[tile_par][1][r] +
                              (double)sh->mu_part[tile_par][2][r];
          p.mu[(long long)lrow * p.W + col] = (float)((double)p.prior_mean + (mu_acc + part) * (double)p.mu_unscale);
        }
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
