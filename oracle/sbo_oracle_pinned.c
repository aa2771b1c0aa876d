/*
 * sbo_oracle_pinned.c -- the part of the oracle that IS pinned by reference
 * source lines: confidence intervals, safe set and subgoal selection of
 * OptimizerNode (src/safe_bayesian_optimization_node.cpp).  See
 * sbo_oracle.h: TEST INFRASTRUCTURE ONLY.
 *
 * Built with -ffp-contract=off: the reference is plain C++14 with no -march
 * flag (CMakeLists.txt:4-9), i.e. baseline x86-64 without FMA, so every
 * product and sum below is rounded on its own, exactly as Eigen evaluates
 * them there.  The GPU path is compared bit-for-bit against this file.
 */
#include "sbo_oracle.h"

#include <math.h>
#include <stdlib.h>

/* :411-416 -- `confidence = beta_ * std_` is a materialised temporary (:412),
 * then two separate column assignments (:414, :415). */
void oracle_confidence_intervals(const double* mu, const double* std_, double beta, int64_t m,
                                 double* q0, double* q1) {
  double* conf = (double*)malloc(sizeof(double) * (size_t)(m > 0 ? m : 1));
  for (int64_t i = 0; i < m; ++i) conf[i] = beta * std_[i];
  for (int64_t i = 0; i < m; ++i) q0[i] = mu[i] - conf[i];
  for (int64_t i = 0; i < m; ++i) q1[i] = mu[i] + conf[i];
  free(conf);
}

/* :409 -- strict `>`; a NaN lower bound compares false, hence unsafe.  The
 * byte value is what :454 writes into the CV_8UC1 image. */
int64_t oracle_safe_set(const double* q0, double f_min, int64_t m, uint8_t* mask) {
  int64_t count = 0;
  for (int64_t i = 0; i < m; ++i) {
    const int safe = q0[i] > f_min;
    mask[i] = safe ? 255 : 0;
    count += safe;
  }
  return count;
}

void oracle_cell_xy(const oracle_grid* g, int64_t cell, double* x, double* y) {
  const int64_t iy = cell / g->W;
  const int64_t ix = cell - iy * g->W;
  const double px = (double)ix * g->rx;
  const double py = (double)iy * g->ry;
  *x = g->x0 + px;
  *y = g->y0 + py;
}

typedef struct {
  double dist;
  int64_t pos;
} dist_pos;

/* std::pair<double,size_t> operator< (:530): first, then second. */
static int cmp_dist_pos(const void* a, const void* b) {
  const dist_pos* p = (const dist_pos*)a;
  const dist_pos* q = (const dist_pos*)b;
  if (p->dist < q->dist) return -1;
  if (q->dist < p->dist) return 1;
  if (p->pos < q->pos) return -1;
  if (q->pos < p->pos) return 1;
  return 0;
}

/* :499-550.  D_.row(idx) is the cell coordinate (:515), width = Q1 - Q0
 * (:516), distance = rowwise().norm() = sqrt(dx*dx + dy*dy) (:519-523). */
int32_t oracle_get_next_subgoal(const int32_t* frontier, int64_t F, const oracle_grid* g,
                                const double* q0, const double* q1, double goal_x, double goal_y) {
  if (F <= 0) return -1; /* :503-506 */
  dist_pos* dp = (dist_pos*)malloc(sizeof(dist_pos) * (size_t)F);
  double* width = (double*)malloc(sizeof(double) * (size_t)F);
  for (int64_t i = 0; i < F; ++i) {
    const int32_t idx = frontier[i];
    double x, y;
    oracle_cell_xy(g, idx, &x, &y);
    width[i] = q1[idx] - q0[idx];
    const double dx = x - goal_x;
    const double dy = y - goal_y;
    const double dx2 = dx * dx;
    const double dy2 = dy * dy;
    dp[i].dist = sqrt(dx2 + dy2);
    dp[i].pos = i;
  }
  qsort(dp, (size_t)F, sizeof(dist_pos), cmp_dist_pos); /* keys are unique: order is total */
  int64_t top = F / 4; /* :533 */
  if (top < 1) top = 1;
  double best_w = -1.0; /* :536 */
  int64_t best = -1;
  for (int64_t i = 0; i < top; ++i) {
    const int64_t p = dp[i].pos;
    if (width[p] > best_w) { /* :543 strict */
      best_w = width[p];
      best = p;
    }
  }
  /* :549 indexes frontier_indices[best_index] unconditionally; best stays -1
   * only when every width in the window is NaN or <= -1, which is undefined
   * behaviour there.  Report -1 instead. */
  const int32_t out = best >= 0 ? frontier[best] : -1;
  free(dp);
  free(width);
  return out;
}

int64_t oracle_argmax_safe(const double* val, const uint8_t* mask, int64_t m, double* best) {
  int64_t bi = -1;
  double bv = 0.0;
  for (int64_t i = 0; i < m; ++i) {
    if (!mask[i]) continue;
    if (bi < 0 || val[i] > bv) {
      bv = val[i];
      bi = i;
    }
  }
  if (best) *best = bv;
  return bi;
}
