// host_example/main.cpp -- ROS-free stand-in for OptimizerNode: walks the call sequence of
// process_terrain_map (reference src/safe_bayesian_optimization_node.cpp:625-723) against the C-ABI of
// include/sbo.h.  Plain C++17, no CUDA headers, no torch: what a maintainer's node links against.
//
//   ctor (:48-77)                         -> sbo_create + sbo_set_grid
//   measurements (publisher.py:137-145)   -> sbo_set_observations / sbo_append_observations
//   request + unpack + ComputeSets        -> sbo_step / sbo_step_incremental (mask lands in the "cv::Mat" buffer)
//   FindSafetyContourIndices (:418-497)   -> stays on the host (OpenCV in the node; a 4-neighbour border
//                                            scan stands in for cv::findContours here, OpenCV-C++ is absent)
//   GetNextSubgoal (:499-550)             -> sbo_select_subgoal
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../include/sbo.h"

static double terrain(double x, double y, double L) {
  return 5.0 + 2.5 * std::sin(6.0 * M_PI * x / L) * std::cos(4.0 * M_PI * y / L) +
         1.5 * std::exp(-((x - 0.3 * L) * (x - 0.3 * L) + (y - 0.6 * L) * (y - 0.6 * L)) / (0.02 * L * L));
}

#define CHECK(call)                                                                                   \
  do {                                                                                                \
    const int32_t st__ = (call);                                                                      \
    if (st__ != SBO_OK) {                                                                             \
      std::fprintf(stderr, "%s failed: %s (%s)\n", #call, sbo_status_string(st__), sbo_last_error(ctx)); \
      return 1;                                                                                       \
    }                                                                                                 \
  } while (0)

int main(int argc, char** argv) {
  const int n = argc > 1 ? std::atoi(argv[1]) : 300;
  const int W = argc > 2 ? std::atoi(argv[2]) : 256, H = argc > 3 ? std::atoi(argv[3]) : 192;
  const double len = 0.4, L = len * std::sqrt(n / 8.0), beta = 2.0, f_min = 4.2;
  const double rx = L / W, ry = L / H;

  sbo_ctx* ctx = nullptr;
  sbo_config cfg{};
  cfg.sigma_f = 1.0; cfg.length_scale = len; cfg.noise_var = 0.1; cfg.prior_mean = 5.0;  // config/lpsc.yaml:35-37
  cfg.precision = SBO_FP32; cfg.kernel = SBO_KERNEL_AUTO; cfg.device = 0;
  cfg.max_obs = n + 64; cfg.max_cells = (int64_t)W * H;
  if (sbo_create(&cfg, &ctx) != SBO_OK) {
    std::fprintf(stderr, "sbo_create: %s\n", sbo_last_error(nullptr));  // no GPU -> the node cannot start (:61-66)
    return 2;
  }
  CHECK(sbo_set_grid(ctx, 0.0, 0.0, rx, ry, W, H, 0, H));

  std::mt19937_64 rng(1234);
  std::uniform_real_distribution<double> U(0.0, L);
  std::normal_distribution<double> N(0.0, std::sqrt(cfg.noise_var));
  std::vector<double> xy(2 * (size_t)n), y((size_t)n);
  for (int i = 0; i < n; ++i) {
    xy[2 * i] = U(rng); xy[2 * i + 1] = U(rng);
    y[i] = terrain(xy[2 * i], xy[2 * i + 1], L) + N(rng);
  }
  CHECK(sbo_set_observations(ctx, xy.data(), y.data(), n));

  std::vector<uint8_t> safety_image((size_t)W * H);  // cv::Mat(H, W, CV_8UC1) in the node
  const double goal[2] = {0.9 * L, 0.9 * L};
  for (int step = 0; step < 3; ++step) {
    sbo_step_out out{};
    if (step == 0) CHECK(sbo_step(ctx, beta, f_min, safety_image.data(), &out));
    else {
      double nxy[2] = {U(rng), U(rng)};
      double ny = terrain(nxy[0], nxy[1], L) + N(rng);
      CHECK(sbo_append_observations(ctx, nxy, &ny, 1));          // one measurement per second (:43)
      CHECK(sbo_step_incremental(ctx, beta, f_min, safety_image.data(), &out));
    }
    // frontier: safe cells with an unsafe (or outside) 4-neighbour -- the node uses cv::findContours here
    std::vector<int32_t> frontier;
    for (int iy = 0; iy < H; ++iy)
      for (int ix = 0; ix < W; ++ix) {
        if (!safety_image[(size_t)iy * W + ix]) continue;
        const bool edge = ix == 0 || iy == 0 || ix == W - 1 || iy == H - 1 || !safety_image[(size_t)iy * W + ix - 1] ||
                          !safety_image[(size_t)iy * W + ix + 1] || !safety_image[(size_t)(iy - 1) * W + ix] ||
                          !safety_image[(size_t)(iy + 1) * W + ix];
        if (edge) frontier.push_back(iy * W + ix);
      }
    int32_t cell = -1;
    CHECK(sbo_select_subgoal(ctx, frontier.data(), (int64_t)frontier.size(), goal[0], goal[1], &cell));
    long long count = 0;
    for (uint8_t v : safety_image) count += v == 255;
    std::printf("step %d: n=%d safe=%lld (mask count %lld) argmax_ucb=(%lld, %.4f) frontier=%zu subgoal cell=%d "
                "(%.3f, %.3f) kernel=%d posterior %.3f ms total %.3f ms\n",
                step, out.n_obs, (long long)out.safe_count, count, (long long)out.argmax_ucb.idx, out.argmax_ucb.value,
                frontier.size(), cell, cell >= 0 ? (cell % W) * rx : NAN, cell >= 0 ? (cell / W) * ry : NAN,
                out.kernel_used, out.ms_posterior, out.ms_total);
    if (count != out.safe_count) { std::fprintf(stderr, "mask / safe_count mismatch\n"); return 3; }
    if (frontier.empty() != (cell < 0)) { std::fprintf(stderr, "subgoal sentinel mismatch\n"); return 3; }
  }
  sbo_destroy(ctx);
  std::printf("host_example OK\n");
  return 0;
}
