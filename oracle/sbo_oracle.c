/*
 * sbo_oracle.c -- GP-regression part of the CPU oracle (fp64).  See
 * sbo_oracle.h for the scope statement: TEST INFRASTRUCTURE ONLY, and for
 * this file PARITY UNPINNED (the reference obtains mu/sigma from the external
 * mapping_package over a ROS 2 service, src/safe_bayesian_optimization_node.cpp
 * :568-644; the algorithm below is Rasmussen & Williams Alg. 2.1 with the
 * hyper-parameters of config/lpsc.yaml:35-37).
 *
 * Compiled with -O3 and OpenMP; floating-point contraction is allowed here
 * because nothing in this file is compared bit-for-bit.
 */
#include "sbo_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int32_t oracle_num_threads(void) {
#ifdef _OPENMP
  return (int32_t)omp_get_max_threads();
#else
  return 1;
#endif
}

static inline double rbf(double dx, double dy, double sf2, double inv2l2) {
  return sf2 * exp(-(dx * dx + dy * dy) * inv2l2);
}

void oracle_build_K(const double* xy, int32_t n, const oracle_gp_hyper* hp, double* K) {
  const double sf2 = hp->sigma_f * hp->sigma_f;
  const double inv2l2 = 1.0 / (2.0 * hp->length * hp->length);
#pragma omp parallel for schedule(static)
  for (int32_t i = 0; i < n; ++i) {
    for (int32_t j = 0; j < n; ++j) {
      double v = rbf(xy[2 * i] - xy[2 * j], xy[2 * i + 1] - xy[2 * j + 1], sf2, inv2l2);
      if (i == j) v += hp->noise_var;
      K[(size_t)i * n + j] = v;
    }
  }
}

/* Row-oriented (Cholesky-Banachiewicz) factorisation: row i of L needs rows
 * 0..i-1 only, so the inner dot products run over contiguous memory. */
int32_t oracle_cholesky(double* A, int32_t n) {
  for (int32_t i = 0; i < n; ++i) {
    double* Ai = A + (size_t)i * n;
    for (int32_t j = 0; j <= i; ++j) {
      const double* Aj = A + (size_t)j * n;
      double s = Ai[j];
      for (int32_t k = 0; k < j; ++k) s -= Ai[k] * Aj[k];
      if (j < i) {
        Ai[j] = s / Aj[j];
      } else {
        if (!(s > 0.0)) return i + 1;
        Ai[i] = sqrt(s);
      }
    }
    for (int32_t j = i + 1; j < n; ++j) Ai[j] = 0.0;
  }
  return 0;
}

/* Column c of L^-1 by forward substitution on e_c; columns are independent. */
void oracle_tri_inverse(const double* L, int32_t n, double* Linv) {
  memset(Linv, 0, sizeof(double) * (size_t)n * n);
#pragma omp parallel for schedule(dynamic, 8)
  for (int32_t c = 0; c < n; ++c) {
    Linv[(size_t)c * n + c] = 1.0 / L[(size_t)c * n + c];
    for (int32_t i = c + 1; i < n; ++i) {
      const double* Li = L + (size_t)i * n;
      double s = 0.0;
      for (int32_t k = c; k < i; ++k) s += Li[k] * Linv[(size_t)k * n + c];
      Linv[(size_t)i * n + c] = -s / Li[i];
    }
  }
}

int32_t oracle_gp_fit(const double* xy, const double* y, int32_t n, const oracle_gp_hyper* hp,
                      double* L, double* Linv, double* alpha) {
  oracle_build_K(xy, n, hp, L);
  int32_t st = oracle_cholesky(L, n);
  if (st != 0) return st;
  oracle_tri_inverse(L, n, Linv);
  /* alpha = L^-T (L^-1 r), r = y - prior_mean; both solves by substitution on L. */
  double* w = (double*)malloc(sizeof(double) * (size_t)n);
  for (int32_t i = 0; i < n; ++i) {
    double s = y[i] - hp->prior_mean;
    const double* Li = L + (size_t)i * n;
    for (int32_t k = 0; k < i; ++k) s -= Li[k] * w[k];
    w[i] = s / Li[i];
  }
  for (int32_t i = n - 1; i >= 0; --i) {
    double s = w[i];
    for (int32_t k = i + 1; k < n; ++k) s -= L[(size_t)k * n + i] * alpha[k];
    alpha[i] = s / L[(size_t)i * n + i];
  }
  free(w);
  return 0;
}

#define CELL_BLOCK 16

void oracle_posterior(const double* xy, int32_t n, const oracle_gp_hyper* hp, const double* Linv,
                      const double* alpha, const oracle_grid* g, int64_t cell0, int64_t cell1,
                      double* mu, double* sigma2) {
  const double sf2 = hp->sigma_f * hp->sigma_f;
  const double inv2l2 = 1.0 / (2.0 * hp->length * hp->length);
  const int64_t ncell = cell1 - cell0;
  const int64_t nblk = (ncell + CELL_BLOCK - 1) / CELL_BLOCK;
#pragma omp parallel
  {
    double* ks = (double*)malloc(sizeof(double) * (size_t)CELL_BLOCK * (size_t)(n > 0 ? n : 1));
#pragma omp for schedule(dynamic, 4)
    for (int64_t b = 0; b < nblk; ++b) {
      const int64_t c0 = cell0 + b * CELL_BLOCK;
      const int nc = (int)((c0 + CELL_BLOCK <= cell1) ? CELL_BLOCK : (cell1 - c0));
      double m_acc[CELL_BLOCK], q_acc[CELL_BLOCK];
      for (int c = 0; c < CELL_BLOCK; ++c) {
        m_acc[c] = 0.0;
        q_acc[c] = 0.0;
        double* kc = ks + c; /* ks is k-major: ks[k*CELL_BLOCK + c] */
        if (c < nc) {
          double cx, cy;
          oracle_cell_xy(g, c0 + c, &cx, &cy);
          double s = 0.0;
          for (int32_t k = 0; k < n; ++k) {
            const double kv = rbf(cx - xy[2 * k], cy - xy[2 * k + 1], sf2, inv2l2);
            kc[(size_t)k * CELL_BLOCK] = kv;
            s += kv * alpha[k];
          }
          m_acc[c] = s;
        } else {
          for (int32_t k = 0; k < n; ++k) kc[(size_t)k * CELL_BLOCK] = 0.0;
        }
      }
      /* v_j = Linv[j, 0..j] . k*[0..j]; one Linv row serves the whole block. */
      for (int32_t j = 0; j < n; ++j) {
        const double* row = Linv + (size_t)j * n;
        double v[CELL_BLOCK];
        for (int c = 0; c < CELL_BLOCK; ++c) v[c] = 0.0;
        for (int32_t k = 0; k <= j; ++k) {
          const double r = row[k];
#pragma omp simd
          for (int c = 0; c < CELL_BLOCK; ++c) v[c] += r * ks[(size_t)k * CELL_BLOCK + c];
        }
        for (int c = 0; c < CELL_BLOCK; ++c) q_acc[c] += v[c] * v[c];
      }
      for (int c = 0; c < nc; ++c) {
        mu[c0 + c - cell0] = hp->prior_mean + m_acc[c];
        sigma2[c0 + c - cell0] = sf2 - q_acc[c];
      }
    }
    free(ks);
  }
}
