"""Sampled parity of a BO step against the fp64 scipy restatement -- TEST INFRASTRUCTURE ONLY.

SURVEY.md 8(d) M6 report for configurations whose full grid the CPU oracle cannot finish (C3-C5): an fp64 Cholesky of
the observation covariance (scipy / OpenBLAS, all host threads) once per observation set, then mu / sigma^2 / LCB of a
random sample of cells by triangular solves.  Imported by tests/ and by bench.py's parity leg, never by the product.
Same GP definition as oracle.py (textbook regression: RBF, K + noise_var I, zero mean on y - prior, latent variance;
PARITY UNPINNED by the reference, see oracle.py).
"""
from __future__ import annotations

import time

import numpy as np

from . import oracle as O


class SampledReference:
    """fp64 factorisation of one observation set; evaluates sampled cells of a grid on demand."""

    def __init__(self, p, xy=None, y=None):
        from scipy.linalg import cho_solve
        self.p = p
        self.xy = np.asarray(p.xy if xy is None else xy, dtype=np.float64)
        yv = np.asarray(p.y if y is None else y, dtype=np.float64)
        self.hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
        self.grid = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        t0 = time.perf_counter()
        n = self.xy.shape[0]
        K = np.empty((n, n))
        for i0 in range(0, n, 2048):                      # blocked: rbf_np's temporaries are n x block, not n x n x 3
            K[i0:i0 + 2048] = O.rbf_np(self.xy[i0:i0 + 2048], self.xy, self.hp)
        K[np.diag_indices(n)] += self.hp.noise_var
        self.L = np.linalg.cholesky(K)
        del K
        self.alpha = cho_solve((self.L, True), yv - self.hp.prior_mean)
        self.fit_seconds = time.perf_counter() - t0

    def evaluate(self, cells):
        """(mu, sigma^2) fp64 at global cell indices `cells`."""
        from scipy.linalg import solve_triangular
        cells = np.asarray(cells, dtype=np.int64)
        mu = np.empty(cells.size); s2 = np.empty(cells.size)
        for b0 in range(0, cells.size, 2048):
            c = cells[b0:b0 + 2048]
            cx, cy = self.grid.cell_xy(c)
            Ks = O.rbf_np(np.stack([cx, cy], axis=1), self.xy, self.hp)
            mu[b0:b0 + c.size] = self.hp.prior_mean + Ks @ self.alpha
            V = solve_triangular(self.L, Ks.T, lower=True)
            s2[b0:b0 + c.size] = self.hp.sigma_f ** 2 - np.einsum("ij,ij->j", V, V)
        return mu, s2


def sample_cells(p, rows, per_band, rng):
    """`per_band` random cells from each band of grid rows in `rows` = [(r0, r1), ...]; sorted global indices per band."""
    out = []
    for r0, r1 in rows:
        m = (r1 - r0) * p.W
        out.append(r0 * p.W + np.sort(rng.choice(m, min(per_band, m), replace=False)))
    return out


def parity_block(p, mu, s2, mask, mu_ref, s2_ref, tol):
    """SURVEY 8(d) M6 fields for one set of sampled cells.  mu / s2 / mask: the step's values AT those cells."""
    mu = np.asarray(mu, dtype=np.float64); s2 = np.asarray(s2, dtype=np.float64)
    sf2 = p.sigma_f ** 2
    mu_scale = max(1.0, np.abs(mu_ref).max())
    rel_mu = np.abs(mu - mu_ref) / np.maximum(np.abs(mu_ref), 1e-3 * np.abs(mu_ref).max())
    ds2 = np.abs(s2 - s2_ref)
    lcb_ref = mu_ref - p.beta * np.sqrt(np.maximum(s2_ref, 0.0))
    lcb = mu - p.beta * np.sqrt(np.maximum(s2, 0.0))
    mref = lcb_ref > p.f_min                                     # strict, reference :409
    mm = (np.asarray(mask) != 0) != mref
    # north_star's tolerance band: cells whose margin to f_min lies within the tolerance of the LCB itself
    band = tol * mu_scale + p.beta * tol * sf2 / (2.0 * np.sqrt(np.maximum(s2_ref, 1e-12)))
    inside = mm & (np.abs(lcb_ref - p.f_min) <= band)
    return {"cells": int(mu.size),
            "mu_rel_max": float(rel_mu.max()), "mu_rel_mean": float(rel_mu.mean()),
            "s2_h1_max": float((ds2 / np.maximum(s2_ref, 1e-3 * sf2)).max()),
            "s2_abs_max_over_sf2": float(ds2.max() / sf2),
            "s2_plain_rel_max": float((ds2 / np.maximum(np.abs(s2_ref), 1e-300)).max()),
            "s2_ref_min_over_sf2": float(s2_ref.min() / sf2),
            "lcb_abs_max_over_mu_scale": float(np.abs(lcb - lcb_ref).max() / mu_scale),
            "mask_mismatch_inside_band": int(inside.sum()), "mask_mismatch_outside_band": int(mm.sum() - inside.sum())}


def merge_blocks(blocks):
    """Fold parity blocks of several bands (max of maxima, sums of counts, weighted mean)."""
    out = {}
    n = sum(b["cells"] for b in blocks)
    for k in blocks[0]:
        if k == "cells":
            out[k] = n
        elif k.endswith("_mean"):
            out[k] = float(sum(b[k] * b["cells"] for b in blocks) / max(n, 1))
        elif k.startswith("mask_"):
            out[k] = int(sum(b[k] for b in blocks))
        elif k == "s2_ref_min_over_sf2":
            out[k] = float(min(b[k] for b in blocks))
        else:
            out[k] = float(max(b[k] for b in blocks))
    return out
