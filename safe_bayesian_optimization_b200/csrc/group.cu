// group.cu -- one context that fans out to several GPUs of the box ("ctx internally fans out to <= 8
// GPUs").  The reference host is ONE process with a SingleThreadedExecutor (src/safe_bayesian_optimization_node.cpp
// :1500-1508): it cannot launch one process per GPU, so the drop-in for it is a single object it calls from its one
// thread.  A group owns one sbo_ctx per device, cuts the grid into bands of rows of equal estimated work
// (sbo_plan_row_bands) and drives the members from one host thread per device for the duration of a call; the members
// write their bands of the safe mask straight into the caller's one H x W image, and the per-band records are folded on
// the host with the library's total order (value desc, index asc) -- inside one process no collective is needed for 56
// bytes per GPU.  The one-process-per-GPU entry (sbo_comm_*, NCCL) stays for torchrun-style hosts.
#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

struct sbo_group {
  sbo_config cfg{};
  std::vector<sbo_ctx*> ctx;
  std::vector<int32_t> cuts;          // [n + 1] rows
  bool grid_set = false;
  double x0 = 0, y0 = 0, rx = 0, ry = 0;
  int W = 0, H = 0;
  std::vector<double> xy, y;          // the observations in arrival order (band planning)
  sbo::FitPeers* peers = nullptr;     // rendezvous of the members' host threads during a shared fit (factor_dist.cuh)
  std::string err;
};

namespace {

thread_local std::string g_group_create_error;

int32_t gfail(sbo_group* g, int32_t st, const std::string& msg) {
  if (g) g->err = msg;
  else g_group_create_error = msg;
  return st;
}

// run fn(i) for every member on its own host thread; first non-OK status wins, its context's text is kept
template <typename F> int32_t for_each_member(sbo_group* g, F fn) {
  const int n = (int)g->ctx.size();
  std::vector<int32_t> st((size_t)n, SBO_OK);
  if (n == 1) {
    st[0] = fn(0);
  } else {
    std::vector<std::thread> th;
    th.reserve((size_t)n);
    for (int i = 0; i < n; ++i) th.emplace_back([&, i] { st[i] = fn(i); });
    for (auto& t : th) t.join();
  }
  for (int i = 0; i < n; ++i)
    if (st[i] != SBO_OK) {
      g->err = "device " + std::to_string(g->ctx[i]->device) + ": " + sbo_last_error(g->ctx[i]);
      return st[i];
    }
  return SBO_OK;
}

int32_t apply_bands(sbo_group* g) {
  const int n = (int)g->ctx.size();
  std::vector<int32_t> cuts((size_t)n + 1);
  const int32_t nobs = (int32_t)g->y.size();
  const int32_t st = sbo_plan_row_bands(&g->cfg, g->xy.data(), nobs, g->x0, g->y0, g->rx, g->ry, g->W, g->H, n, 0, cuts.data(), nullptr);
  if (st != SBO_OK) return gfail(g, st, "sbo_group: band planning failed");
  if (cuts == g->cuts) return SBO_OK;
  g->cuts = cuts;
  for (int i = 0; i < n; ++i) {
    const int32_t s = sbo_set_grid(g->ctx[i], g->x0, g->y0, g->rx, g->ry, g->W, g->H, cuts[i], cuts[i + 1]);
    if (s != SBO_OK) return gfail(g, s, std::string("sbo_group: ") + sbo_last_error(g->ctx[i]));
  }
  return SBO_OK;
}

}  // namespace

extern "C" {

int32_t sbo_group_create(const sbo_config* cfg, const int32_t* devices, int32_t n_devices, sbo_group** out) {
  if (!cfg || !devices || !out || n_devices < 1 || n_devices > 64) return gfail(nullptr, SBO_ERR_INVALID_ARG, "sbo_group_create: bad arguments");
  *out = nullptr;
  sbo_group* g = new (std::nothrow) sbo_group();
  if (!g) return gfail(nullptr, SBO_ERR_CUDA, "sbo_group_create: out of host memory");
  g->cfg = *cfg;
  for (int i = 0; i < n_devices; ++i) {
    sbo_config c = *cfg;   // every member can hold the whole grid (9 B per cell): bands of equal work are not bands of equal rows
    c.device = devices[i];
    sbo_ctx* ctx = nullptr;
    const int32_t st = sbo_create(&c, &ctx);
    if (st != SBO_OK) {
      g_group_create_error = std::string("device ") + std::to_string(devices[i]) + ": " + sbo_last_error(nullptr);
      for (sbo_ctx* m : g->ctx) sbo_destroy(m);
      delete g;
      return st;
    }
    g->ctx.push_back(ctx);
  }
  if (n_devices > 1) {   // the full fit is shared between the members: peer copies between their buffers
    g->peers = sbo::fit_peers_create(n_devices);
    if (!g->peers) {
      for (sbo_ctx* m : g->ctx) sbo_destroy(m);
      delete g;
      return gfail(nullptr, SBO_ERR_CUDA, "sbo_group_create: out of host memory");
    }
    for (int i = 0; i < n_devices; ++i) {
      sbo::fit_peers_attach(g->peers, g->ctx[i], i);
      cudaSetDevice(devices[i]);
      for (int j = 0; j < n_devices; ++j) {
        int can = 0;
        if (devices[j] == devices[i] || cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) != cudaSuccess || !can) continue;
        cudaDeviceEnablePeerAccess(devices[j], 0);   // (already enabled: harmless error, cleared below)
      }
      cudaGetLastError();
    }
  }
  *out = g;
  return SBO_OK;
}

int32_t sbo_group_destroy(sbo_group* g) {
  if (!g) return SBO_ERR_INVALID_ARG;
  for (sbo_ctx* m : g->ctx) sbo_destroy(m);
  if (g->peers) sbo::fit_peers_destroy(g->peers);
  delete g;
  return SBO_OK;
}

const char* sbo_group_last_error(const sbo_group* g) { return g ? g->err.c_str() : g_group_create_error.c_str(); }
int32_t sbo_group_size(const sbo_group* g) { return g ? (int32_t)g->ctx.size() : SBO_ERR_INVALID_ARG; }
sbo_ctx* sbo_group_member(sbo_group* g, int32_t i) { return (g && i >= 0 && i < (int32_t)g->ctx.size()) ? g->ctx[i] : nullptr; }

int32_t sbo_group_bands(const sbo_group* g, int32_t* row_cuts) {
  if (!g || !row_cuts) return SBO_ERR_INVALID_ARG;
  if (g->cuts.empty()) return SBO_ERR_STATE;
  std::copy(g->cuts.begin(), g->cuts.end(), row_cuts);
  return SBO_OK;
}

int32_t sbo_group_set_option(sbo_group* g, int32_t option, int32_t value) {
  if (!g) return SBO_ERR_INVALID_ARG;
  for (sbo_ctx* m : g->ctx) {
    const int32_t st = sbo_set_option(m, option, value);
    if (st != SBO_OK) return gfail(g, st, sbo_last_error(m));
  }
  return SBO_OK;
}

int32_t sbo_group_set_grid(sbo_group* g, double x0, double y0, double rx, double ry, int32_t W, int32_t H) {
  if (!g) return SBO_ERR_INVALID_ARG;
  if (W < 1 || H < 1 || (int64_t)W * H > g->cfg.max_cells) return gfail(g, SBO_ERR_CAPACITY, "sbo_group_set_grid: W * H exceeds max_cells");
  g->x0 = x0; g->y0 = y0; g->rx = rx; g->ry = ry; g->W = W; g->H = H;
  g->grid_set = true;
  g->cuts.clear();
  return apply_bands(g);
}

int32_t sbo_group_set_observations(sbo_group* g, const double* xy, const double* y, int32_t n) {
  if (!g) return SBO_ERR_INVALID_ARG;
  if (n < 0 || (n > 0 && (!xy || !y))) return gfail(g, SBO_ERR_INVALID_ARG, "sbo_group_set_observations: null input or n < 0");
  const std::vector<double> old_xy = g->xy, old_y = g->y;
  g->xy.assign(xy, xy + 2 * (size_t)n);
  g->y.assign(y, y + (size_t)n);
  if (g->grid_set) SBO_TRY(apply_bands(g));   // the bands follow the observations (block sparsity of K*)
  const int32_t st = for_each_member(g, [&](int i) { return sbo_set_observations(g->ctx[i], xy, y, n); });
  if (st != SBO_OK) {   // every member restored its previous observations (same data, same verdict): so does the group
    const std::string why = g->err;
    g->xy = old_xy; g->y = old_y;
    if (g->grid_set) apply_bands(g);
    g->err = why;
  }
  return st;
}

int32_t sbo_group_append_observations(sbo_group* g, const double* xy, const double* y, int32_t k) {
  if (!g) return SBO_ERR_INVALID_ARG;
  if (k < 0 || (k > 0 && (!xy || !y))) return gfail(g, SBO_ERR_INVALID_ARG, "sbo_group_append_observations: null input or k < 0");
  const int32_t st = for_each_member(g, [&](int i) { return sbo_append_observations(g->ctx[i], xy, y, k); });
  if (st == SBO_OK) {   // bands stay: an incremental step needs the stored posterior of the same cells
    g->xy.insert(g->xy.end(), xy, xy + 2 * (size_t)k);
    g->y.insert(g->y.end(), y, y + (size_t)k);
  }
  return st;
}

int32_t sbo_group_num_observations(const sbo_group* g) { return g ? sbo_num_observations(g->ctx[0]) : SBO_ERR_INVALID_ARG; }

static int32_t group_step(sbo_group* g, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out, bool incremental) {
  if (!g) return SBO_ERR_INVALID_ARG;
  if (!g->grid_set || g->cuts.empty()) return gfail(g, SBO_ERR_STATE, "sbo_group_step: grid not set");
  const int n = (int)g->ctx.size();
  std::vector<sbo_step_out> parts((size_t)n);
  const int32_t st = for_each_member(g, [&](int i) {
    uint8_t* dst = mask_host ? mask_host + (size_t)g->cuts[i] * g->W : nullptr;   // this band of the one H x W image
    return incremental ? sbo_step_incremental(g->ctx[i], beta, f_min, dst, &parts[i]) : sbo_step(g->ctx[i], beta, f_min, dst, &parts[i]);
  });
  SBO_TRY(st);
  if (out) sbo_merge_step_out(parts.data(), n, out);
  return SBO_OK;
}

int32_t sbo_group_step(sbo_group* g, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out) {
  return group_step(g, beta, f_min, mask_host, out, false);
}
int32_t sbo_group_step_incremental(sbo_group* g, double beta, double f_min, uint8_t* mask_host, sbo_step_out* out) {
  return group_step(g, beta, f_min, mask_host, out, true);
}

static int32_t group_copy(sbo_group* g, double* host_dst, bool mu) {
  if (!g || !host_dst) return SBO_ERR_INVALID_ARG;
  if (g->cuts.empty()) return gfail(g, SBO_ERR_STATE, "sbo_group_copy_*: grid not set");
  return for_each_member(g, [&](int i) {
    double* dst = host_dst + (size_t)g->cuts[i] * g->W;
    if (g->cuts[i + 1] == g->cuts[i]) return (int32_t)SBO_OK;
    return mu ? sbo_copy_mu_f64(g->ctx[i], dst) : sbo_copy_sigma2_f64(g->ctx[i], dst);
  });
}
int32_t sbo_group_copy_mu_f64(sbo_group* g, double* host_dst) { return group_copy(g, host_dst, true); }
int32_t sbo_group_copy_sigma2_f64(sbo_group* g, double* host_dst) { return group_copy(g, host_dst, false); }

// GetNextSubgoal's rule (reference :508-549) over a frontier that crosses the bands: every member computes the
// (distance key, CI width) of the frontier cells it owns (select_kernel mode 1), the host gathers the two arrays by
// owner, and member 0 runs the selection itself (radix select of the closest quarter + masked argmax, mode 2).
int32_t sbo_group_select_subgoal(sbo_group* g, const int32_t* frontier, int64_t F, double goal_x, double goal_y, int32_t* out_cell) {
  if (!g || !out_cell) return SBO_ERR_INVALID_ARG;
  *out_cell = -1;
  if (F < 0 || (F > 0 && !frontier)) return gfail(g, SBO_ERR_INVALID_ARG, "sbo_group_select_subgoal: null frontier or F < 0");
  if (F == 0) return SBO_OK;
  if (g->cuts.empty()) return gfail(g, SBO_ERR_STATE, "sbo_group_select_subgoal: grid not set");
  const int n = (int)g->ctx.size();
  if (n == 1) return sbo_select_subgoal(g->ctx[0], frontier, F, goal_x, goal_y, out_cell);
  std::vector<std::vector<unsigned long long>> keys((size_t)n);
  std::vector<std::vector<double>> width((size_t)n);
  SBO_TRY(for_each_member(g, [&](int i) -> int32_t {
    sbo_ctx* c = g->ctx[i];
    if (c->source != sbo::PosteriorSource::kStep) return sbo::fail(c, SBO_ERR_STATE, "sbo_group_select_subgoal: no step yet");
    SBO_TRY(sbo::select_upload(c, frontier, F));
    SBO_TRY(sbo::launch_select(c, c->d_frontier, F, goal_x, goal_y, 1));
    keys[i].resize((size_t)F);
    width[i].resize((size_t)F);
    SBO_CUDA(c, cudaMemcpyAsync(keys[i].data(), c->d_sel_keys, sizeof(unsigned long long) * (size_t)F, cudaMemcpyDeviceToHost, c->stream));
    SBO_CUDA(c, cudaMemcpyAsync(width[i].data(), c->d_sel_width, sizeof(double) * (size_t)F, cudaMemcpyDeviceToHost, c->stream));
    SBO_CUDA(c, cudaStreamSynchronize(c->stream));
    return SBO_OK;
  }));
  for (int64_t e = 0; e < F; ++e) {   // owner of entry e = the band its cell's row lies in
    const int64_t cell = frontier[e];
    const int64_t row = cell / g->W;
    if (cell < 0 || row >= g->H) return gfail(g, SBO_ERR_INVALID_ARG, "sbo_group_select_subgoal: frontier holds a cell outside the grid");
    const int owner = (int)(std::upper_bound(g->cuts.begin(), g->cuts.end(), (int32_t)row) - g->cuts.begin()) - 1;
    if (owner != 0) { keys[0][e] = keys[owner][e]; width[0][e] = width[owner][e]; }
  }
  sbo_ctx* c = g->ctx[0];
  SBO_CUDA(c, cudaSetDevice(c->device));
  SBO_CUDA(c, cudaMemcpyAsync(c->d_sel_keys, keys[0].data(), sizeof(unsigned long long) * (size_t)F, cudaMemcpyHostToDevice, c->stream));
  SBO_CUDA(c, cudaMemcpyAsync(c->d_sel_width, width[0].data(), sizeof(double) * (size_t)F, cudaMemcpyHostToDevice, c->stream));
  SBO_TRY(sbo::launch_select(c, c->d_frontier, F, goal_x, goal_y, 2));
  SBO_CUDA(c, cudaMemcpyAsync(c->h_sel_out, c->d_sel_out, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  SBO_CUDA(c, cudaStreamSynchronize(c->stream));
  *out_cell = *c->h_sel_out;
  return SBO_OK;
}

}  // extern "C"
