// host_example/group_main.cpp -- the single-process, multi-GPU form of the step: ONE object, one
// caller thread, several GPUs.  The node (reference src/safe_bayesian_optimization_node.cpp:1500-1508) is one process
// with a SingleThreadedExecutor; this is what it would call at :631-647 on a multi-GPU box.
//
//   group_example N_MEMBERS [n W H] [--same-device]
//
// Runs one full BO step through an sbo_group of N_MEMBERS devices (0 .. N_MEMBERS-1, or device 0 repeated with
// --same-device) and the same step through a single sbo_ctx on device 0, and checks that the two agree: identical
// safe-mask image, safe count and argmax cells; then an append + incremental step through the group.  Plain C++17.
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "../include/sbo.h"

static double terrain(double x, double y, double L) {
  return 5.0 + 2.5 * std::sin(6.0 * M_PI * x / L) * std::cos(4.0 * M_PI * y / L) +
         1.5 * std::exp(-((x - 0.3 * L) * (x - 0.3 * L) + (y - 0.6 * L) * (y - 0.6 * L)) / (0.02 * L * L));
}
static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char** argv) {
  const int members = argc > 1 ? std::atoi(argv[1]) : 2;
  const int n = argc > 4 ? std::atoi(argv[2]) : 2304;
  const int W = argc > 4 ? std::atoi(argv[3]) : 1024, H = argc > 4 ? std::atoi(argv[4]) : 768;
  bool same = false;
  for (int i = 1; i < argc; ++i) same = same || !std::strcmp(argv[i], "--same-device");
  const double len = 0.4, L = len * std::sqrt(n / 8.0), beta = 2.0, f_min = 4.2;

  sbo_config cfg{};
  cfg.sigma_f = 1.0; cfg.length_scale = len; cfg.noise_var = 0.1; cfg.prior_mean = 5.0;
  cfg.precision = SBO_FP32; cfg.kernel = SBO_KERNEL_AUTO; cfg.device = 0;
  cfg.max_obs = n + 64; cfg.max_cells = (int64_t)W * H;
  std::vector<int32_t> devices((size_t)members);
  for (int i = 0; i < members; ++i) devices[i] = same ? 0 : i;

  std::mt19937_64 rng(99);
  std::uniform_real_distribution<double> U(0.0, L);
  std::normal_distribution<double> N(0.0, std::sqrt(cfg.noise_var));
  std::vector<double> xy(2 * (size_t)n), y((size_t)n);
  for (int i = 0; i < n; ++i) {
    xy[2 * i] = U(rng); xy[2 * i + 1] = U(rng);
    y[i] = terrain(xy[2 * i], xy[2 * i + 1], L) + N(rng);
  }

  sbo_group* g = nullptr;
  if (sbo_group_create(&cfg, devices.data(), members, &g) != SBO_OK) {
    std::fprintf(stderr, "sbo_group_create: %s\n", sbo_group_last_error(nullptr));
    return 2;
  }
#define GCHECK(call)                                                                                         \
  do {                                                                                                       \
    const int32_t st__ = (call);                                                                             \
    if (st__ != SBO_OK) {                                                                                    \
      std::fprintf(stderr, "%s failed: %s (%s)\n", #call, sbo_status_string(st__), sbo_group_last_error(g)); \
      return 1;                                                                                              \
    }                                                                                                        \
  } while (0)
  GCHECK(sbo_group_set_grid(g, 0.0, 0.0, L / W, L / H, W, H));
  std::vector<uint8_t> image((size_t)W * H), image1((size_t)W * H);
  sbo_step_out out{}, warm{};
  double t0 = now_ms();
  GCHECK(sbo_group_set_observations(g, xy.data(), y.data(), n));
  const double t_fit = now_ms() - t0;
  GCHECK(sbo_group_step(g, beta, f_min, image.data(), &warm));
  t0 = now_ms();
  GCHECK(sbo_group_step(g, beta, f_min, image.data(), &out));
  const double t_step = now_ms() - t0;
  std::vector<int32_t> cuts((size_t)members + 1);
  GCHECK(sbo_group_bands(g, cuts.data()));
  std::printf("group of %d: bands", members);
  for (int i = 0; i <= members; ++i) std::printf(" %d", cuts[i]);
  std::printf(" | fit %.2f ms, step %.2f ms wall (slowest member: posterior %.2f ms), safe=%lld argmax_ucb=(%lld, %.5f)\n", t_fit,
              t_step, out.ms_posterior, (long long)out.safe_count, (long long)out.argmax_ucb.idx, out.argmax_ucb.value);

  // the same step on one context
  sbo_ctx* ctx = nullptr;
  if (sbo_create(&cfg, &ctx) != SBO_OK) { std::fprintf(stderr, "sbo_create: %s\n", sbo_last_error(nullptr)); return 2; }
  sbo_step_out one{};
  if (sbo_set_grid(ctx, 0.0, 0.0, L / W, L / H, W, H, 0, H) != SBO_OK || sbo_set_observations(ctx, xy.data(), y.data(), n) != SBO_OK ||
      sbo_step(ctx, beta, f_min, image1.data(), &one) != SBO_OK) {
    std::fprintf(stderr, "single context: %s\n", sbo_last_error(ctx));
    return 1;
  }
  long long diff = 0;
  for (size_t i = 0; i < image.size(); ++i) diff += image[i] != image1[i];
  std::printf("single context: posterior %.2f ms, safe=%lld argmax_ucb=(%lld, %.5f); mask bytes that differ: %lld\n", one.ms_posterior,
              (long long)one.safe_count, (long long)one.argmax_ucb.idx, one.argmax_ucb.value, diff);
  if (diff != 0 || one.safe_count != out.safe_count || one.argmax_ucb.idx != out.argmax_ucb.idx ||
      one.argmax_width.idx != out.argmax_width.idx || out.n_cells != (int64_t)W * H) {
    std::fprintf(stderr, "group and single context disagree\n");
    return 3;
  }
  // subgoal rule over a frontier that crosses the bands: every 97th safe cell, against the single context
  std::vector<int32_t> frontier;
  for (size_t i = 0; i < image.size(); i += 97)
    if (image[i]) frontier.push_back((int32_t)i);
  int32_t cg = -2, c1 = -3;
  GCHECK(sbo_group_select_subgoal(g, frontier.data(), (int64_t)frontier.size(), 0.9 * L, 0.9 * L, &cg));
  if (sbo_select_subgoal(ctx, frontier.data(), (int64_t)frontier.size(), 0.9 * L, 0.9 * L, &c1) != SBO_OK || cg != c1) {
    std::fprintf(stderr, "subgoal: group %d, single %d\n", cg, c1);
    return 3;
  }
  // the steady state: one new measurement, incremental step on every band
  const double nx[2] = {0.37 * L, 0.61 * L}, ny = terrain(nx[0], nx[1], L);
  GCHECK(sbo_group_append_observations(g, nx, &ny, 1));
  sbo_step_out inc{}, inc1{};
  GCHECK(sbo_group_step_incremental(g, beta, f_min, image.data(), &inc));
  if (sbo_append_observations(ctx, nx, &ny, 1) != SBO_OK || sbo_step_incremental(ctx, beta, f_min, image1.data(), &inc1) != SBO_OK) return 1;
  diff = 0;
  for (size_t i = 0; i < image.size(); ++i) diff += image[i] != image1[i];
  std::printf("incremental step: group safe=%lld, single safe=%lld, subgoal cell %d, mask bytes that differ: %lld\n",
              (long long)inc.safe_count, (long long)inc1.safe_count, cg, diff);
  if (diff != 0 || inc.safe_count != inc1.safe_count || inc.n_obs != n + 1) return 3;
  sbo_destroy(ctx);
  sbo_group_destroy(g);
  std::printf("group_example OK\n");
  return 0;
}
