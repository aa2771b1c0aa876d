// host_example/main.cpp -- ROS-free stand-in for OptimizerNode: walks the call sequence of
// process_terrain_map (reference src/safe_bayesian_optimization_node.cpp:625-723) against the C-ABI of
// include/sbo.h.  Plain C++17, no CUDA headers, no torch: what a maintainer's node links against.
//
//   ctor (:48-77)                         -> sbo_create + sbo_set_grid
//   measurements (publisher.py:137-145)   -> sbo_set_observations, then sbo_ingest_measurements one at a time
//   request + unpack + ComputeSets        -> sbo_step / sbo_step_incremental (mask lands in the "cv::Mat" buffer)
//   FindSafetyContourIndices (:418-497)   -> stays on the host (OpenCV in the node; a 4-neighbour border
//                                            scan stands in for cv::findContours here, OpenCV-C++ is absent)
//   GetNextSubgoal (:499-550) + goal-in-polygon test and polydist projection (:650-706)
//                                         -> sbo_subgoal_point (a fixed quadrilateral stands in for the eroded alpha
//                                            shape: CGAL / Boost are absent here and stay on the host in the node)
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
      sbo_measurement msg{};                                     // one SpatialMeasurement per second (:43, :137-145)
      msg.x = U(rng); msg.y = U(rng); msg.z = 0.0;
      msg.value = terrain(msg.x, msg.y, L) + N(rng);
      msg.uncertainty = 0.1; msg.stamp_sec = step;
      int32_t size = 0, refit = 0;
      CHECK(sbo_ingest_measurements(ctx, &msg, 1, &size, &refit));
      if (size != n + step || refit) { std::fprintf(stderr, "ingestion: size %d refit %d\n", size, refit); return 3; }
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
    // stand-in for eroded_concave_polygon_: a quadrilateral in the lower-left of the domain (the goal lies outside it,
    // so the frontier subgoal is chosen and projected; step 2 moves the goal inside to take the other branch)
    const double poly[8] = {0.1 * L, 0.1 * L, 0.6 * L, 0.15 * L, 0.55 * L, 0.6 * L, 0.12 * L, 0.5 * L};
    const int32_t ring_size = 4;
    const double gx = step == 2 ? 0.3 * L : goal[0], gy = step == 2 ? 0.3 * L : goal[1];
    int32_t cell = -1, kind = SBO_SUBGOAL_NONE;
    double subgoal[2] = {NAN, NAN}, pdist = 0.0;
    CHECK(sbo_subgoal_point(ctx, frontier.data(), (int64_t)frontier.size(), gx, gy, poly, &ring_size, 1, subgoal, &cell,
                            &kind, &pdist));
    if (step == 2 && (kind != SBO_SUBGOAL_GOAL || subgoal[0] != gx || subgoal[1] != gy)) {
      std::fprintf(stderr, "goal inside the polygon must be its own subgoal\n");
      return 3;
    }
    long long count = 0;
    for (uint8_t v : safety_image) count += v == 255;
    std::printf("step %d: n=%d safe=%lld (mask count %lld) argmax_ucb=(%lld, %.4f) frontier=%zu subgoal kind=%d cell=%d "
                "raw (%.3f, %.3f) -> published (%.3f, %.3f) dist %.3f kernel=%d posterior %.3f ms total %.3f ms\n",
                step, out.n_obs, (long long)out.safe_count, count, (long long)out.argmax_ucb.idx, out.argmax_ucb.value,
                frontier.size(), kind, cell, cell >= 0 ? (cell % W) * rx : NAN, cell >= 0 ? (cell / W) * ry : NAN,
                subgoal[0], subgoal[1], pdist, out.kernel_used, out.ms_posterior, out.ms_total);
    if (count != out.safe_count) { std::fprintf(stderr, "mask / safe_count mismatch\n"); return 3; }
    if (step != 2 && frontier.empty() != (cell < 0)) { std::fprintf(stderr, "subgoal sentinel mismatch\n"); return 3; }
  }
  sbo_destroy(ctx);
  std::printf("host_example OK\n");
  return 0;
}
