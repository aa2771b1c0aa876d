// factor_dist.cuh -- the fit shared between the GPUs of a box (included by factor.cu, inside namespace sbo).
//
// "L^-1 broadcast vs redundant factorisation -- pick by measurement".  Measured (round 2, C4 cut into
// bands): every rank repeating the 0.22 s fp64 fit at n = 16 384 is the serial term of the strong-scaled step
// (posterior 4.05 s / N + 0.22 s).  Here the O(n^3) work is split by BLOCK COLUMNS, dealt to the ranks in snake
// order (0 1 .. W-1 W-1 .. 1 0 ...: column c costs ~(nblk - c) tiles per step, the snake evens that out):
//
//   Cholesky, step k:  the owner of column k factors A_kk (D_k = L_kk^-1) and computes the panel L_ik = A_ik D_k^T,
//                      then sends [status, D_k, panel k] to everybody -- the ONE exchange per step;
//                      every rank applies A_ij -= L_ik L_jk^T to the block columns j it owns.
//                      Look-ahead of one step: the owner of column k + 1 updates that column first, factors it and
//                      starts its exchange before the rest of its step-k update, so the exchange of step k + 1
//                      travels under the trailing update of step k.
//   L^-1, step k:      X_kc = D_k B_kc and B_ic -= L_ik X_kc for the block columns c <= k a rank owns (columns of
//                      L^-1 are independent of each other: no exchange until the end);
//   end:               every rank sends its block columns of L^-1 to every rank; alpha, the row norm and the fp16
//                      planes are then computed from the full L^-1 on each rank as in the replicated fit.
//
// Every entry of L and L^-1 is produced by the same sequence of fp64 operations as in the replicated fit where that runs
// the 128 x 128 DMMA tile routine too (n > 2944: factor_dmma.cuh; same kernels with W = 1), so the result is bit-identical
// to it -- tests/test_gpu_group.py; against the 64 x 64 FMA tiles of a small replicated fit it agrees to rounding.
//
// Two transports behind the same schedule:
//   * one process per GPU (sbo_comm_init):   ncclBroadcast of a packed panel message per step, ncclAllGather of the
//                                            packed block columns at the end (comm.cu);
//   * one process, several GPUs (sbo_group): peer copies (cudaMemcpy2DAsync straight out of the owner's d_A / d_Dinv /
//                                            d_Linv, which are final once written) ordered by cross-device events;
//                                            the host threads of the members meet in a FitPeers object.

struct FitPeers {
  std::mutex mu;
  std::condition_variable cv;
  int world = 0;
  std::vector<sbo_ctx*> member;      // by rank
  std::vector<int> posted;           // [nblk] the owner has recorded "panel k ready" (reset by the entry barrier)
  std::vector<int> info;             // [world] Cholesky status of each member
  int arrived = 0;
  uint64_t generation = 0;
  bool aborted = false;
};

FitPeers* fit_peers_create(int world) {
  FitPeers* p = new (std::nothrow) FitPeers();
  if (!p) return nullptr;
  p->world = world;
  p->member.assign((size_t)world, nullptr);
  p->info.assign((size_t)world, 0);
  return p;
}
void fit_peers_destroy(FitPeers* p) { delete p; }
void fit_peers_attach(FitPeers* p, sbo_ctx* ctx, int rank) {
  p->member[(size_t)rank] = ctx;
  ctx->fit_peers = p;
  ctx->fit_rank = rank;
}

namespace {

// panel message of step k (NCCL transport): [0] Cholesky status so far, [16 ..) D_k, then the rem x 128-row panel
constexpr int kMsgHeader = 16;   // doubles

__global__ void __launch_bounds__(256)
pack_panel_kernel(const double* __restrict__ A, int ld, int k, int rem, const double* __restrict__ Dk,
                  const int* __restrict__ info, double* __restrict__ msg) {
  const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x, nd = (size_t)NB * NB, total = nd + (size_t)rem * nd;
  if (e == 0) msg[0] = (double)info[0];
  if (e >= total) return;
  if (e < nd) {
    msg[kMsgHeader + e] = Dk[e];
  } else {
    const size_t p = e - nd, row = p / NB, c = p - row * NB;
    msg[kMsgHeader + e] = A[((size_t)(k + 1) * NB + row) * ld + (size_t)k * NB + c];
  }
}

__global__ void __launch_bounds__(256)
unpack_panel_kernel(double* __restrict__ A, int ld, int k, int rem, double* __restrict__ Dk, int* __restrict__ info,
                    const double* __restrict__ msg) {
  const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x, nd = (size_t)NB * NB, total = nd + (size_t)rem * nd;
  if (e == 0 && info[0] == 0 && msg[0] != 0.0) info[0] = (int)msg[0];   // first failing row, in step order as on one GPU
  if (e >= total) return;
  if (e < nd) {
    Dk[e] = msg[kMsgHeader + e];
  } else {
    const size_t p = e - nd, row = p / NB, c = p - row * NB;
    A[((size_t)(k + 1) * NB + row) * ld + (size_t)k * NB + c] = msg[kMsgHeader + e];
  }
}

// packed block columns of L^-1 of one rank: column j (block column c = map.col(j)) holds rows [c NB, np), 128 wide;
// columns back to back.  off(j) = sum_{j' < j} (np - col(j') NB) * NB.
__host__ __device__ inline size_t packed_off(const ColMap& map, int j, int np) {
  size_t o = 0;
  for (int q = 0; q < j; ++q) o += (size_t)(np - map.col(q) * NB) * NB;
  return o;
}

template <bool kPack>
__global__ void __launch_bounds__(256)
linv_cols_kernel(double* __restrict__ X, int ld, int np, ColMap map, double* __restrict__ buf) {
  const int j = blockIdx.y, c = map.col(j);
  const int rows = np - c * NB;
  const size_t off = packed_off(map, j, np);
  for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < (size_t)rows * NB; e += (size_t)gridDim.x * 256) {
    const size_t row = e / NB, cc = e - row * NB;
    double* x = X + ((size_t)c * NB + row) * ld + (size_t)c * NB + cc;
    if (kPack) buf[off + e] = *x;
    else *x = buf[off + e];
  }
}

size_t packed_size(int W, int r, int nblk, int np) {
  ColMap m{W, r};
  return packed_off(m, m.count_le(nblk - 1), np);
}

// ---- same-process rendezvous ----------------------------------------------------------------------------------------
// A member that is fitted on its own (not through the group) would leave the others waiting: every wait gives up after
// this long and fails the group's fit with SBO_ERR_COMM instead of hanging the host.
constexpr std::chrono::seconds kPeerTimeout(120);

int32_t peers_barrier(sbo_ctx* ctx, bool reset_posts, int nblk) {
  FitPeers* p = ctx->fit_peers;
  std::unique_lock<std::mutex> lk(p->mu);
  if (p->aborted) return fail(ctx, SBO_ERR_COMM, "shared fit: a member of the group failed or did not arrive");
  const uint64_t gen = p->generation;
  if (++p->arrived == p->world) {
    p->arrived = 0;
    if (reset_posts) p->posted.assign((size_t)nblk, 0);
    ++p->generation;
    p->cv.notify_all();
  } else {
    if (!p->cv.wait_for(lk, kPeerTimeout, [&] { return p->generation != gen || p->aborted; })) p->aborted = true;
  }
  if (p->aborted) return fail(ctx, SBO_ERR_COMM, "shared fit: a member of the group failed or did not arrive");
  return SBO_OK;
}

void peers_post(FitPeers* p, int k) {
  std::lock_guard<std::mutex> lk(p->mu);
  p->posted[(size_t)k] = 1;
  p->cv.notify_all();
}

int32_t peers_wait_posted(sbo_ctx* ctx, int k) {
  FitPeers* p = ctx->fit_peers;
  std::unique_lock<std::mutex> lk(p->mu);
  if (!p->cv.wait_for(lk, kPeerTimeout, [&] { return p->posted[(size_t)k] != 0 || p->aborted; })) p->aborted = true;
  if (p->aborted) return fail(ctx, SBO_ERR_COMM, "shared fit: a member of the group failed or did not arrive");
  return SBO_OK;
}

void peers_abort(FitPeers* p) {
  std::lock_guard<std::mutex> lk(p->mu);
  p->aborted = true;
  p->cv.notify_all();
}

int32_t ensure_dist_streams(sbo_ctx* ctx, int nblk) {
  if (!ctx->fit_hi) {
    int lo = 0, hi = 0;
    SBO_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));   // lo = least priority (largest number)
    SBO_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->fit_hi, cudaStreamNonBlocking, hi));
    SBO_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->fit_cm, cudaStreamNonBlocking, hi));
    SBO_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->fit_lo, cudaStreamNonBlocking, lo));
  }
  while ((int)ctx->ev_factor.size() < 4 + 2 * nblk) {
    cudaEvent_t e;
    SBO_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ctx->ev_factor.push_back(e);
  }
  return SBO_OK;
}

int32_t ensure_buffer(sbo_ctx* ctx, double** buf, size_t* cap, size_t need) {
  if (*cap >= need) return SBO_OK;
  if (*buf) SBO_CUDA(ctx, cudaFree(*buf));
  *buf = nullptr;
  *cap = 0;
  SBO_CUDA(ctx, cudaMalloc(buf, sizeof(double) * need));
  *cap = need;
  return SBO_OK;
}

// Is the next full fit shared?  (n is known; every rank / member takes the same decision from the same numbers.)
int dist_world(const sbo_ctx* ctx) {
  const int W = ctx->fit_peers ? ctx->fit_peers->world : (ctx->nccl_comm ? ctx->world : 1);
  if (W <= 1 || ctx->opt.dist_fit == 0) return 1;
  const int nblk = (ctx->n + NB - 1) / NB;
  if (nblk < 2 || (nblk + W - 1) / W > 256) return 1;
  if (ctx->opt.dist_fit == 2 && ctx->n < 8192) return 1;   // below that the fit is a chain of one-CTA launches: nothing to share
  return W;
}

int32_t factor_full_dist_impl(sbo_ctx* ctx, int W) {
  const int n = ctx->n, ld = ctx->n_cap;
  const int np = (n + NB - 1) / NB * NB, nblk = np / NB;
  const double sf2 = ctx->cfg.sigma_f * ctx->cfg.sigma_f;
  const double inv2l2 = 1.0 / (2.0 * ctx->cfg.length_scale * ctx->cfg.length_scale);
  // W = 1: the same schedule on one GPU (look-ahead of the next panel, prioritised streams), nothing to exchange
  const bool solo = W == 1;
  FitPeers* peers = solo ? nullptr : ctx->fit_peers;
  const int me = solo ? 0 : (peers ? ctx->fit_rank : ctx->rank);
  const ColMap map{W, me};
  cudaStream_t st = ctx->stream;

  SBO_TRY(ensure_factor_attr(ctx));
  SBO_TRY(ensure_dist_streams(ctx, nblk));
  cudaStream_t hi = ctx->fit_hi, lo = ctx->fit_lo, cm = ctx->fit_cm;
  const size_t msg_max = kMsgHeader + (size_t)NB * NB * (size_t)nblk;
  size_t slot = 0;
  if (!peers && !solo) {
    SBO_TRY(ensure_buffer(ctx, &ctx->d_fit_stage, &ctx->fit_stage_cap, msg_max));
    for (int r = 0; r < W; ++r) slot = std::max(slot, packed_size(W, r, nblk, np));
    SBO_TRY(ensure_buffer(ctx, &ctx->d_fit_send, &ctx->fit_send_cap, slot));
    SBO_TRY(ensure_buffer(ctx, &ctx->d_fit_recv, &ctx->fit_recv_cap, slot * (size_t)W));
  }
  auto ev_panel = [&](int k) { return ctx->ev_factor[1 + 2 * k]; };
  auto ev_recv = [&](int k) { return ctx->ev_factor[2 + 2 * k]; };
  cudaEvent_t ev_fork = ctx->ev_factor[0], ev_lo = ctx->ev_factor[1 + 2 * nblk], ev_cm = ctx->ev_factor[2 + 2 * nblk],
              ev_hi = ctx->ev_factor[3 + 2 * nblk];

  ctx->linv_r1 = 0.0;
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_info, 0, sizeof(InfoRecord), st));
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_Linv, 0, sizeof(double) * (size_t)ld * ld, st));
  SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_alpha, 0, sizeof(double) * (size_t)ld, st));
  if (ctx->d_alpha_f32) SBO_CUDA(ctx, cudaMemsetAsync(ctx->d_alpha_f32, 0, sizeof(float) * (size_t)ld, st));
  {
    dim3 b(32, 8), g((np + 31) / 32, (np + 7) / 8);
    build_K_kernel<<<g, b, 0, st>>>(ctx->d_xy, n, np, ld, sf2, inv2l2, ctx->cfg.noise_var, ctx->d_A);
    SBO_LAUNCH_CHECK(ctx, "build_K_kernel");
  }
  SBO_CUDA(ctx, cudaEventRecord(ev_fork, st));
  SBO_CUDA(ctx, cudaStreamWaitEvent(hi, ev_fork, 0));
  SBO_CUDA(ctx, cudaStreamWaitEvent(lo, ev_fork, 0));
  SBO_CUDA(ctx, cudaStreamWaitEvent(cm, ev_fork, 0));
  // nobody may read a peer's buffers (or reuse the post flags) before every member is past its memsets' enqueue AND the
  // previous fit's copies are done (the exit barrier of that fit); the events below order the device side
  if (peers) SBO_TRY(peers_barrier(ctx, true, nblk));

  auto diag_and_panel = [&](int k) -> int32_t {   // owner of column k, on `hi`
    double* Dk = ctx->d_Dinv + (size_t)k * NB * NB;
    diag_chol_inv_kernel<<<1, kDiagThreads, kDiagSmem, hi>>>(ctx->d_A + (size_t)k * NB * ld + (size_t)k * NB, ld, NB, Dk, NB,
                                                             k * NB, ctx->d_info);
    SBO_LAUNCH_CHECK(ctx, "diag_chol_inv_kernel");
    const int rem = nblk - k - 1;
    if (rem > 0) {
      panel_trsm_dm_kernel<<<rem, kDmThreads, kDmSmem, hi>>>(ctx->d_A, ld, k, Dk);
      SBO_LAUNCH_CHECK(ctx, "panel_trsm_dm_kernel");
    }
    SBO_CUDA(ctx, cudaEventRecord(ev_panel(k), hi));
    if (peers) peers_post(peers, k);
    return SBO_OK;
  };

  std::vector<char> panel_done((size_t)nblk, 0);
  for (int k = 0; k < nblk; ++k) {
    const int own = ColMap::owner(k, W), rem = nblk - k - 1;
    double* Dk = ctx->d_Dinv + (size_t)k * NB * NB;
    if (own == me && !panel_done[(size_t)k]) SBO_TRY(diag_and_panel(k));
    // ---- the exchange of step k ----
    const size_t msg = kMsgHeader + (size_t)NB * NB * (size_t)(1 + rem);
    const unsigned msg_blocks = (unsigned)(((size_t)NB * NB * (size_t)(1 + rem) + 255) / 256);
    if (solo) {
      // nothing to exchange
    } else if (!peers) {
      if (own == me) {
        SBO_CUDA(ctx, cudaStreamWaitEvent(cm, ev_panel(k), 0));
        pack_panel_kernel<<<msg_blocks, 256, 0, cm>>>(ctx->d_A, ld, k, rem, Dk, ctx->d_info, ctx->d_fit_stage);
        SBO_LAUNCH_CHECK(ctx, "pack_panel_kernel");
      }
      SBO_TRY(comm_broadcast(ctx, ctx->d_fit_stage, sizeof(double) * msg, own, cm));
      if (own != me) {
        unpack_panel_kernel<<<msg_blocks, 256, 0, cm>>>(ctx->d_A, ld, k, rem, Dk, ctx->d_info, ctx->d_fit_stage);
        SBO_LAUNCH_CHECK(ctx, "unpack_panel_kernel");
      }
      SBO_CUDA(ctx, cudaEventRecord(ev_recv(k), cm));
    } else if (own != me) {
      SBO_TRY(peers_wait_posted(ctx, k));
      sbo_ctx* o = peers->member[(size_t)own];
      SBO_CUDA(ctx, cudaStreamWaitEvent(cm, o->ev_factor[1 + 2 * k], 0));
      SBO_CUDA(ctx, cudaMemcpyAsync(Dk, o->d_Dinv + (size_t)k * NB * NB, sizeof(double) * NB * NB, cudaMemcpyDefault, cm));
      if (rem > 0) {
        const size_t off = (size_t)(k + 1) * NB * ld + (size_t)k * NB;
        SBO_CUDA(ctx, cudaMemcpy2DAsync(ctx->d_A + off, sizeof(double) * ld, o->d_A + off, sizeof(double) * (size_t)o->n_cap,
                                        sizeof(double) * NB, (size_t)rem * NB, cudaMemcpyDefault, cm));
      }
      SBO_CUDA(ctx, cudaEventRecord(ev_recv(k), cm));
    }
    cudaEvent_t ready = (own == me) ? ev_panel(k) : ev_recv(k);
    if (own != me) SBO_CUDA(ctx, cudaStreamWaitEvent(hi, ready, 0));
    SBO_CUDA(ctx, cudaStreamWaitEvent(lo, ready, 0));
    // ---- trailing update of my block columns; the next panel first ----
    if (rem > 0) {
      const int j_next = map.count_le(k);                 // index of my first column > k
      const int j_end = map.count_le(nblk - 1);
      int j0 = j_next;
      if (ColMap::owner(k + 1, W) == me) {                // my column j_next IS k + 1
        syrk_dm_kernel<<<dim3(rem, 1), kDmThreads, kDmSmem, hi>>>(ctx->d_A, ld, k, map, j_next);
        SBO_LAUNCH_CHECK(ctx, "syrk_dm_kernel");
        SBO_TRY(diag_and_panel(k + 1));
        panel_done[(size_t)(k + 1)] = 1;
        j0 = j_next + 1;
      }
      if (j_end > j0) {
        syrk_dm_kernel<<<dim3(rem, j_end - j0), kDmThreads, kDmSmem, hi>>>(ctx->d_A, ld, k, map, j0);
        SBO_LAUNCH_CHECK(ctx, "syrk_dm_kernel");
      }
    }
    // ---- rows of L^-1 for my block columns <= k ----
    const int jc = map.count_le(k);
    if (jc > 0) {
      trtri_row_dm_kernel<<<jc, kDmThreads, kDmSmem, lo>>>(ctx->d_Linv, ld, k, Dk, map);
      SBO_LAUNCH_CHECK(ctx, "trtri_row_dm_kernel");
      if (rem > 0) {
        trtri_update_dm_kernel<<<dim3(jc, rem), kDmThreads, kDmSmem, lo>>>(ctx->d_Linv, ld, ctx->d_A, k, map);
        SBO_LAUNCH_CHECK(ctx, "trtri_update_dm_kernel");
      }
    }
  }
  // ---- every rank's block columns of L^-1 to every rank ----
  const int jmine = map.count_le(nblk - 1);
  SBO_CUDA(ctx, cudaEventRecord(ev_lo, lo));
  if (solo) {
    // every column is here already
  } else if (!peers) {
    if (jmine > 0) {
      linv_cols_kernel<true><<<dim3(256, jmine), 256, 0, lo>>>(ctx->d_Linv, ld, np, map, ctx->d_fit_send);
      SBO_LAUNCH_CHECK(ctx, "linv_cols_kernel");
    }
    SBO_CUDA(ctx, cudaEventRecord(ev_lo, lo));
    SBO_CUDA(ctx, cudaStreamWaitEvent(cm, ev_lo, 0));
    SBO_TRY(comm_all_gather(ctx, ctx->d_fit_send, ctx->d_fit_recv, sizeof(double) * slot, cm));
    for (int r = 0; r < W; ++r) {
      const ColMap pm{W, r};
      const int jr = pm.count_le(nblk - 1);
      if (r == me || jr == 0) continue;
      linv_cols_kernel<false><<<dim3(256, jr), 256, 0, cm>>>(ctx->d_Linv, ld, np, pm, ctx->d_fit_recv + slot * (size_t)r);
      SBO_LAUNCH_CHECK(ctx, "linv_cols_kernel");
    }
  } else {
    SBO_TRY(peers_barrier(ctx, false, nblk));             // every member has recorded its "my columns are final" event
    for (int r = 0; r < W; ++r) {
      if (r == me) continue;
      sbo_ctx* o = peers->member[(size_t)r];
      const ColMap pm{W, r};
      const int jr = pm.count_le(nblk - 1);
      SBO_CUDA(ctx, cudaStreamWaitEvent(cm, o->ev_factor[1 + 2 * nblk], 0));
      for (int j = 0; j < jr; ++j) {
        const int c = pm.col(j);
        const size_t off = (size_t)c * NB * ld + (size_t)c * NB;
        SBO_CUDA(ctx, cudaMemcpy2DAsync(ctx->d_Linv + off, sizeof(double) * ld, o->d_Linv + off, sizeof(double) * (size_t)o->n_cap,
                                        sizeof(double) * NB, (size_t)(np - c * NB), cudaMemcpyDefault, cm));
      }
    }
  }
  SBO_CUDA(ctx, cudaEventRecord(ev_cm, cm));
  SBO_CUDA(ctx, cudaEventRecord(ev_hi, hi));
  SBO_CUDA(ctx, cudaStreamWaitEvent(st, ev_lo, 0));
  SBO_CUDA(ctx, cudaStreamWaitEvent(st, ev_cm, 0));
  SBO_CUDA(ctx, cudaStreamWaitEvent(st, ev_hi, 0));
  if (np > n) {
    clear_pad_kernel<<<(np - n + 127) / 128, 128, 0, st>>>(ctx->d_Linv, ld, n, np);
    SBO_LAUNCH_CHECK(ctx, "clear_pad_kernel");
  }
  row_l1_max_kernel<<<(n * 32 + 255) / 256, 256, 0, st>>>(ctx->d_Linv, ld, 0, n, reinterpret_cast<double*>(ctx->d_info) + 1);
  SBO_LAUNCH_CHECK(ctx, "row_l1_max_kernel");
  InfoRecord rec;
  int32_t verdict = check_info(ctx, "sbo_set_observations", &rec);   // synchronises the stream: my copies are done
  if (peers) {   // the verdict of the group = the first failing member's; nobody leaves while a peer still copies from it
    {
      std::lock_guard<std::mutex> lk(peers->mu);
      peers->info[(size_t)me] = rec.info;
    }
    SBO_TRY(peers_barrier(ctx, false, nblk));
    int first = 0;
    {
      std::lock_guard<std::mutex> lk(peers->mu);
      for (int r = 0; r < W; ++r)
        if (peers->info[(size_t)r] != 0 && (first == 0 || peers->info[(size_t)r] < first)) first = peers->info[(size_t)r];
    }
    SBO_TRY(peers_barrier(ctx, false, nblk));             // (info[] is rewritten by the next fit only after this)
    if (first != 0 && verdict == SBO_OK)
      verdict = fail(ctx, SBO_ERR_NOT_POSDEF, "sbo_set_observations: covariance not positive definite at row " + std::to_string(first - 1));
  }
  SBO_TRY(verdict);
  ctx->linv_r1 = rec.r1_full;
  SBO_TRY(compute_alpha(ctx, n));
  SBO_TRY(refresh_planes(ctx, 0, ld));
  return SBO_OK;
}

}  // namespace

// 0: the caller runs the small-n fit (64 x 64 tiles, eager first pairs); otherwise this schedule ran -- shared between W ranks,
// or on this GPU alone when the fit is large enough for 128 x 128 tiles -- and *status holds its result
int factor_full_dist(sbo_ctx* ctx, int32_t* status) {
  const int W = dist_world(ctx);
  const int nblk = (ctx->n + NB - 1) / NB;
  const bool large = nblk * (nblk + 1) / 2 > 2 * ctx->sm_count;   // more 128-tiles per launch than 2 x SMs (n > 2944 on 148 SMs)
  ctx->fit_dist_used = 0;
  if (W <= 1 && !large) return 0;
  *status = factor_full_dist_impl(ctx, W);
  if (*status != SBO_OK && *status != SBO_ERR_NOT_POSDEF && ctx->fit_peers) peers_abort(ctx->fit_peers);
  if (*status != SBO_OK) {   // leave the side streams idle whatever happened
    cudaStreamSynchronize(ctx->fit_hi);
    cudaStreamSynchronize(ctx->fit_lo);
    cudaStreamSynchronize(ctx->fit_cm);
  }
  ctx->fit_dist_used = W > 1 ? W : 0;
  return 1;
}
