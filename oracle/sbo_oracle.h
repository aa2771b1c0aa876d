/*
 * sbo_oracle.h -- CPU restatement (fp64) of the safe-BO hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker or as
 * the timed CPU baseline.  The product path (safe_bayesian_optimization_b200/)
 * never links, imports or calls it and has no CPU fallback.
 *
 * Parity status
 *   - confidence intervals, safe set, subgoal selection: PINNED by the
 *     reference source itself (src/safe_bayesian_optimization_node.cpp:409-416
 *     and :499-550); the reference ships no golden vectors or tests
 *     (SURVEY.md section 4), so the pin is the line-by-line semantics plus
 *     the known-answer cases in tests/.
 *   - GP posterior (mu, sigma^2): PARITY UNPINNED.  The reference receives
 *     mu/sigma from the external, un-vendored ROS 2 package `mapping_package`
 *     (terrain_mapping_node; no version pinned anywhere -- only named in
 *     launch/safe_bayesian_optimization.launch.py:111-117) through the
 *     GetTerrainMapWithUncertainty service
 *     (src/safe_bayesian_optimization_node.cpp:75-77, :568-644).  The
 *     arithmetic restated here is textbook GP regression (Rasmussen &
 *     Williams, Algorithm 2.1) with the hyper-parameters of
 *     config/lpsc.yaml:35-37, cross-checked against scipy and scikit-learn.
 */
#ifndef SBO_ORACLE_H
#define SBO_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  double sigma_f;    /* config/lpsc.yaml:37  sigma_f       */
  double length;     /* config/lpsc.yaml:36  length_scale  */
  double noise_var;  /* config/lpsc.yaml:35  noise_level (taken as a variance) */
  double prior_mean; /* constant mean subtracted from y, added back to mu */
} oracle_gp_hyper;

typedef struct {
  double x0, y0;   /* coordinates of cell (ix=0, iy=0)            */
  double rx, ry;   /* cell pitch: terrain_map.resolution (:50-51) */
  int32_t W, H;    /* n_width_cells, n_height_cells (:612-613)    */
} oracle_grid;

/* ---- GP regression (external mapping node; unpinned) -------------------- */

/* K[i*n+j] = sigma_f^2 exp(-|xi-xj|^2 / (2 l^2)) + noise_var * (i==j). */
void oracle_build_K(const double* xy, int32_t n, const oracle_gp_hyper* hp, double* K);

/* In-place lower Cholesky of a row-major n x n matrix (upper part zeroed).
 * Returns 0, or 1+index of the first non-positive pivot. */
int32_t oracle_cholesky(double* A, int32_t n);

/* Linv = L^-1 for lower-triangular row-major L (upper part of Linv zeroed). */
void oracle_tri_inverse(const double* L, int32_t n, double* Linv);

/* Full fit: L, Linv (n x n row-major) and alpha = K^-1 (y - prior_mean).
 * Returns the oracle_cholesky status. */
int32_t oracle_gp_fit(const double* xy, const double* y, int32_t n, const oracle_gp_hyper* hp,
                      double* L, double* Linv, double* alpha);

/* Posterior over the cells [cell0, cell1) of the grid (cell = iy*W + ix):
 *   mu  = prior_mean + k*^T alpha
 *   s2  = sigma_f^2 - |Linv k*|^2        (latent variance, not clamped)
 * Writes mu[c-cell0], sigma2[c-cell0].  Uses OpenMP over cells. */
void oracle_posterior(const double* xy, int32_t n, const oracle_gp_hyper* hp, const double* Linv,
                      const double* alpha, const oracle_grid* g, int64_t cell0, int64_t cell1,
                      double* mu, double* sigma2);

/* ---- Reference arithmetic (pinned) -------------------------------------- */

/* src/safe_bayesian_optimization_node.cpp:411-416  ComputeConfidenceIntervals
 *   conf = beta * std ; Q0 = mu - conf ; Q1 = mu + conf   (three passes, fp64) */
void oracle_confidence_intervals(const double* mu, const double* std_, double beta, int64_t m,
                                 double* q0, double* q1);

/* src/safe_bayesian_optimization_node.cpp:409  UpdateSafeSet
 *   S = Q0 > f_min (strict; NaN -> unsafe).  mask byte = 255 / 0 as written
 *   into the cv::Mat at :454.  Returns the safe count (S_.count(), :426). */
int64_t oracle_safe_set(const double* q0, double f_min, int64_t m, uint8_t* mask);

/* Cell coordinates: x = x0 + ix*rx, y = y0 + iy*ry, each product and sum
 * rounded separately (no FMA). */
void oracle_cell_xy(const oracle_grid* g, int64_t cell, double* x, double* y);

/* src/safe_bayesian_optimization_node.cpp:499-550  GetNextSubgoal, given the
 * frontier index list of :501.  width = Q1 - Q0 (:516), dist = sqrt(dx^2+dy^2)
 * (:519-523), sort (dist, position) pairs (:526-530), T = max(1, F/4) (:533),
 * strictly-greater scan from -1.0 (:536-547).  Returns frontier[best] or -1
 * when F == 0 (:503-506). */
int32_t oracle_get_next_subgoal(const int32_t* frontier, int64_t F, const oracle_grid* g,
                                const double* q0, const double* q1, double goal_x, double goal_y);

/* argmax over safe cells of `val`, ties to the lowest index; -1 if none. */
int64_t oracle_argmax_safe(const double* val, const uint8_t* mask, int64_t m, double* best);

int32_t oracle_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif
