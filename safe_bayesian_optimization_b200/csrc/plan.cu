// plan.cu -- host-side planning shared by the fit, the tcgen05 launchers and the multi-GPU band split.
//
//   * spatial_order: the internal observation order (sbo_copy_order)
//   * chunk_boxes / tile_visit_cost: the block-sparsity estimate of K* for one cell tile -- the same box test the
//     kernel's build_chunk_mask runs, evaluated on the host in fp64
//   * sbo_plan_row_bands: bands of grid rows of equal estimated WORK for G GPUs (SURVEY 8(e) G1 shards by contiguous
//     row bands; under the block-sparse walk the kept fraction varies 0.13 .. 0.24 between bands of C4, so equal
//     row counts finish up to 1.8x apart)
//
// No device code: everything here is a pure function of its arguments, so every rank of a multi-process job derives
// the same cuts from the same observations without a collective.
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

namespace sbo {

namespace {
constexpr int kBM = 128, kBN = 256, kBK = 64;   // tile shape of posterior_tc.cu (cells, rows of L^-1, observations per chunk)
}

// Bands of two length-scales in y, by x inside a band; stable.
void spatial_order(double length_scale, const double* xy, int n, std::vector<int32_t>& order) {
  order.resize((size_t)n);
  for (int i = 0; i < n; ++i) order[i] = i;
  if (n < 2) return;
  double ymin = xy[1];
  for (int i = 1; i < n; ++i) ymin = std::min(ymin, xy[2 * i + 1]);
  const double inv_band = 1.0 / (2.0 * length_scale);
  std::vector<long long> band((size_t)n);
  for (int i = 0; i < n; ++i) band[i] = (long long)std::floor((xy[2 * i + 1] - ymin) * inv_band);
  std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
    if (band[a] != band[b]) return band[a] < band[b];
    return xy[2 * a] < xy[2 * b];
  });
}

double tc_xscale(const sbo_config& cfg) { return std::sqrt(1.4426950408889634 / (2.0 * cfg.length_scale * cfg.length_scale)); }

// Bounding boxes (xmin, xmax, ymin, ymax; scaled coordinates) of the 64-observation chunks of xy_sorted.
void chunk_boxes(const double* xy_sorted, int n, double xs, std::vector<double>& box) {
  const int nchunks = (n + kBK - 1) / kBK;
  box.assign(4 * (size_t)nchunks, 0.0);
  for (int c = 0; c < nchunks; ++c) {
    double x0 = 1e300, x1 = -1e300, y0 = 1e300, y1 = -1e300;
    for (int k = c * kBK; k < std::min(n, (c + 1) * kBK); ++k) {
      const double x = xy_sorted[2 * (size_t)k] * xs, y = xy_sorted[2 * (size_t)k + 1] * xs;
      x0 = std::min(x0, x); x1 = std::max(x1, x); y0 = std::min(y0, y); y1 = std::max(y1, y);
    }
    box[4 * c] = x0; box[4 * c + 1] = x1; box[4 * c + 2] = y0; box[4 * c + 3] = y1;
  }
}

// One cell tile (scaled x range [xlo, xhi], scaled y): MMA work of its walk in units of (row block, chunk) visits.
//   total = dense walk;  kept = visits that survive the cut-off `bits`;  passes = split products issued, as a multiple of one
//   (3 per visit inside `ring`, 1 in the ring between `ring` and `bits`).
void tile_visit_cost(const std::vector<double>& box, int n, double xlo, double xhi, double y, double bits, double ring,
                     double* total, double* kept, double* passes) {
  const int nchunks = (n + kBK - 1) / kBK, nblocks = (n + kBN - 1) / kBN;
  double t = 0.0, k = 0.0, p = 0.0;
  for (int c = 0; c < nchunks; ++c) {
    const double wgt = nblocks - c / (kBN / kBK);  // row blocks that walk chunk c
    const double dx = std::max(std::max(box[4 * c] - xhi, xlo - box[4 * c + 1]), 0.0);
    const double dy = std::max(std::max(box[4 * c + 2] - y, y - box[4 * c + 3]), 0.0);
    const double d2 = dx * dx + dy * dy;
    t += wgt;
    if (d2 <= bits + 0.5) {
      k += wgt;
      p += d2 <= ring + 0.5 ? 3.0 * wgt : wgt;
    }
  }
  *total = t; *kept = k; *passes = p;
}

// Per-term error budget of the block-sparse walk, relative to sigma_f^2.  A term K*[k] L^-1[j][k] that is left out (or
// computed with error) by e sigma_f^2 |L^-1[j][k]| moves v_j by at most e sigma_f^2 R1 (R1 = the largest row 1-norm of
// L^-1), |v| by sqrt(n) times that, and sigma^2 = sigma_f^2 - |v|^2 by at most 2 |v| |dv| <= 2 sigma_f sqrt(n) e sigma_f^2 R1.
// e is chosen so that this stays under 1e-7 sigma_f^2 (1 % of the fp32 tolerance).  R1 is measured after the fit
// (factor.cu) where the walk is planned after it; its a-priori bound is sqrt(n / noise_var) (||L^-1||_2 <= noise_var^-1/2).
double tc_term_budget(const sbo_config& cfg, int n, double r1) {
  const double nn = std::max(n, 1);
  if (!(r1 > 0.0)) r1 = std::sqrt(nn / std::max(cfg.noise_var, 1e-300));
  return 1e-7 / (2.0 * std::sqrt(nn) * cfg.sigma_f * r1);
}

// Cut-off of the walk in bits: chunks whose K* entries are all < 2^-bits sigma_f^2 for a tile (pair) are skipped.
double tc_skip_bits_of(const sbo_config& cfg, int n, double r1) {
  return std::min(60.0, std::max(30.0, std::ceil(-std::log2(tc_term_budget(cfg, n, r1)))));
}

// The same cut-off for the fp64 kernel (posterior_f64.cu): per-term budget for 1e-12 sigma_f^2 on sigma^2 (1 % of the fp64
// tolerance 1e-10); R1 measured by the fit, else its a-priori bound.
double f64_skip_bits_of(const sbo_config& cfg, int n, double r1) {
  const double nn = std::max(n, 1);
  if (!(r1 > 0.0)) r1 = std::sqrt(nn / std::max(cfg.noise_var, 1e-300));
  const double e = 1e-12 / (2.0 * std::sqrt(nn) * cfg.sigma_f * r1);
  return std::min(90.0, std::max(40.0, std::ceil(-std::log2(e))));
}

// Graded split depth: visits whose K* entries are all < 2^-ring sigma_f^2 issue hi.hi only.  Leaving out A_lo.B_hi and
// A_hi.B_lo costs, per term, at most (max(2^-11 K*, q) + 2^-11 K*) |L^-1[j][k]| with q = the fp16 subnormal quantum of the
// scaled K* plane (2^-25 2^-shift_a / sigma_f^2: far entries are subnormal in the hi plane); ring is the smallest integer
// with 2^-10 2^-ring + q <= the per-term budget.  Returns `bits` (no ring) when the budget is below q.
double tc_ring_bits_of(const sbo_config& cfg, int n, double r1, int shift_a) {
  const double e = tc_term_budget(cfg, n, r1), bits = tc_skip_bits_of(cfg, n, r1);
  const double q = std::ldexp(1.0, -25 - shift_a) / (cfg.sigma_f * cfg.sigma_f);
  if (!(e > q)) return bits;
  return std::min(bits, std::max(12.0, std::ceil(-std::log2((e - q) * 1024.0))));
}

}  // namespace sbo

extern "C" int32_t sbo_plan_row_bands(const sbo_config* cfg, const double* xy, int32_t n, double x0, double y0, double rx,
                                      double ry, int32_t W, int32_t H, int32_t parts, int32_t max_rows, int32_t* row_cuts,
                                      double* cost_share) {
  using namespace sbo;
  if (!cfg || !row_cuts || parts < 1 || W < 1 || H < 1 || n < 0 || (n > 0 && !xy)) return SBO_ERR_INVALID_ARG;
  if (max_rows > 0 && (int64_t)max_rows * parts < H) return SBO_ERR_CAPACITY;
  if (parts == 1) {   // nothing to cut
    row_cuts[0] = 0; row_cuts[1] = H;
    if (cost_share) cost_share[0] = 1.0;
    return SBO_OK;
  }
  // cost of one grid row: estimated on every `stride`-th row (the cost varies on the scale of a length-scale, i.e. many rows)
  const int nsamp = std::min(H, 512);
  std::vector<double> cost((size_t)nsamp, 1.0);
  const bool f64_path = cfg->precision == SBO_FP64 && cfg->kernel != SBO_KERNEL_SIMT;
  const bool tensor_path = (cfg->precision == SBO_FP32 || f64_path) && cfg->kernel != SBO_KERNEL_SIMT && n > 8 * kBK && cfg->noise_var > 0.0;
  if (tensor_path) {
    std::vector<int32_t> order;
    spatial_order(cfg->length_scale, xy, n, order);
    std::vector<double> sxy(2 * (size_t)n);
    for (int i = 0; i < n; ++i) { sxy[2 * i] = xy[2 * (size_t)order[i]]; sxy[2 * i + 1] = xy[2 * (size_t)order[i] + 1]; }
    const double xs = tc_xscale(*cfg);
    std::vector<double> box;
    chunk_boxes(sxy.data(), n, xs, box);
    int shift_a, shift_b;
    tc_plane_shifts(*cfg, &shift_a, &shift_b);
    // (fp64: the DMMA kernel walks the same chunks, no ring, its own cut-off; row blocks of 128 instead of 256 change the
    //  weights by a constant factor only)
    const double bits = f64_path ? f64_skip_bits_of(*cfg, n, 0.0) : tc_skip_bits_of(*cfg, n, 0.0);
    const double ring = f64_path ? bits : tc_ring_bits_of(*cfg, n, 0.0, shift_a);
    const int tiles_per_row = (W + kBM - 1) / kBM;
    const int tstep = std::max(1, tiles_per_row / 64);
    // The box test of tile_visit_cost, separated: the x distance depends on (tile column, chunk), the y distance on (row,
    // chunk), and the chunks are banded in y, so most of them fail on dy alone for a given row.  Every term is a small
    // integer (row blocks that walk the chunk, times 1 or 3), so the sums are exact whatever their order.
    const int nchunks = (n + kBK - 1) / kBK, nblocks = (n + kBN - 1) / kBN;
    std::vector<double> wgt((size_t)nchunks);
    double wsum = 0.0;
    for (int c = 0; c < nchunks; ++c) { wgt[c] = nblocks - c / (kBN / kBK); wsum += wgt[c]; }
    std::vector<int> tcols;
    for (int b = 0; b < tiles_per_row; b += tstep) tcols.push_back(b);
    // Per chunk, the squared x distances to the sampled tile columns in ascending order: d2 = fl(dx2 + dy2) is monotone in
    // dx2, so the number of tile columns within a threshold is one binary search on the exact predicate per (row, chunk)
    // instead of a pass over the columns -- the same counts, hence the same cuts, in a fifth of the time (10 -> 2 ms at C4).
    const size_t nt = tcols.size();
    std::vector<double> dx2((size_t)nchunks * nt);
    for (int c = 0; c < nchunks; ++c) {
      double* row = &dx2[(size_t)c * nt];
      for (size_t t = 0; t < nt; ++t) {
        const double tx0 = (x0 + (double)tcols[t] * kBM * rx) * xs, tx1 = tx0 + (kBM - 1) * rx * xs;
        const double xlo = std::min(tx0, tx1), xhi = std::max(tx0, tx1);
        const double dx = std::max(std::max(box[4 * c] - xhi, xlo - box[4 * c + 1]), 0.0);
        row[t] = dx * dx;
      }
      std::sort(row, row + nt);
    }
    auto count_within = [&](const double* row, size_t hi, double dy2, double thr) {   // #t < hi with row[t] + dy2 <= thr
      size_t lo = 0;
      while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (row[mid] + dy2 <= thr) lo = mid + 1; else hi = mid;
      }
      return lo;
    };
    double dense = 0.0, sparse = 0.0;
    for (int a = 0; a < nsamp; ++a) {
      const int row = (int)(((2LL * a + 1) * H) / (2LL * nsamp));
      const double y = (y0 + row * ry) * xs;
      double cst = 0.0;
      for (int c = 0; c < nchunks; ++c) {
        const double dy = std::max(std::max(box[4 * c + 2] - y, y - box[4 * c + 3]), 0.0);
        const double dy2 = dy * dy;
        if (dy2 > bits + 0.5) continue;
        const double* drow = &dx2[(size_t)c * nt];
        const size_t kept = count_within(drow, nt, dy2, bits + 0.5);
        const size_t full = count_within(drow, kept, dy2, ring + 0.5);   // ring <= bits: among the kept ones
        cst += wgt[c] * (3.0 * (double)full + (double)(kept - full));
      }
      cost[a] = cst;
      dense += 3.0 * wsum * (double)nt;
      sparse += cst;
    }
    // the launcher walks densely when little can be skipped (tc_plan_skip): equal cost per row then.  Otherwise a fixed
    // cost per tile (mask build, pipeline fill and drain, epilogues, mu) rides on the MMA passes: calibrated on the two
    // bands of the 2-GPU C4 run of round 2 (t = a passes + b rows gave b H = 6.5 % of the step; profiles/bench_c4_2gpu_r02b.json).
    if (sparse >= 0.90 * dense) std::fill(cost.begin(), cost.end(), 1.0);   // (the launcher's rule: posterior_tc.cu kSkipWorthIt)
    else {
      const double floor_c = 0.07 * sparse / nsamp;
      for (double& c : cost) c += floor_c;
    }
  }
  // prefix sums over rows (sample a covers rows [a H / nsamp, (a + 1) H / nsamp)), cuts at equal shares
  std::vector<double> row_cost((size_t)H);
  for (int r = 0; r < H; ++r) row_cost[r] = cost[std::min(nsamp - 1, (int)((long long)r * nsamp / H))];
  double total = 0.0;
  for (double c : row_cost) total += c;
  row_cuts[0] = 0;
  double acc = 0.0;
  int r = 0;
  for (int g = 1; g < parts; ++g) {
    const double target = total * g / parts;
    while (r < H && acc + 0.5 * row_cost[r] < target) acc += row_cost[r++];
    row_cuts[g] = r;
  }
  row_cuts[parts] = H;
  if (max_rows > 0) {   // capacity of a shard: push excess rows to the neighbours (forward, then backward)
    for (int g = 0; g < parts; ++g)
      if (row_cuts[g + 1] - row_cuts[g] > max_rows) row_cuts[g + 1] = row_cuts[g] + max_rows;
    row_cuts[parts] = H;
    for (int g = parts - 1; g >= 0; --g)
      if (row_cuts[g + 1] - row_cuts[g] > max_rows) row_cuts[g] = row_cuts[g + 1] - max_rows;
  }
  if (cost_share) {
    for (int g = 0; g < parts; ++g) {
      double c = 0.0;
      for (int q = row_cuts[g]; q < row_cuts[g + 1]; ++q) c += row_cost[q];
      cost_share[g] = total > 0.0 ? c / total : 0.0;
    }
  }
  return SBO_OK;
}
