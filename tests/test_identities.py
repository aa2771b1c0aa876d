"""The algebra the incremental kernels rely on, checked in fp64 on the CPU against the oracle's full refit
(oracle/oracle.py: textbook GP regression, Rasmussen & Williams Alg. 2.1).  No GPU.

  after appending k observations, with rows_new = the k new rows of L^-1 and w = L^-1 (y - prior_mean):
    alpha_new   = [alpha_old; 0] + rows_new^T w_new                       (factor.cu append_alpha_kernel)
    u_q(cell)   = rows_new[q] . k*(cell)                                  (posterior_thin.cu, posterior_tc_inc.cuh)
    sigma^2_new = sigma^2_old - sum_q u_q^2,   mu_new = mu_old + sum_q w_q u_q
  and on a grid, with the separable RBF k*(cell(r, c), j) = sigma_f^2 ex(c, j) ey(r, j):
    U_q = A P_q^T,  A = ex (W x n),  P_q[r, j] = sigma_f^2 ey(r, j) rows_new[q][j]          (the incremental GEMM)
"""
import numpy as np
import pytest

from oracle import oracle as O
from safe_bayesian_optimization_b200 import synthetic as S


@pytest.mark.parametrize("n0,k", [(40, 1), (57, 3), (100, 16)])
def test_incremental_identities(n0, k):
    p = S.make_problem("C3", n=n0 + k, W=37, H=23)
    hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    g = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    _, Linv0, alpha0 = O.gp_fit_scipy(p.xy[:n0], p.y[:n0], hp)
    _, Linv1, alpha1 = O.gp_fit_scipy(p.xy, p.y, hp)
    mu0, s20 = O.gp_posterior_scipy(p.xy[:n0], hp, Linv0, alpha0, g)
    mu1, s21 = O.gp_posterior_scipy(p.xy, hp, Linv1, alpha1, g)
    n = n0 + k
    # L^-1 only gains rows
    assert np.abs(Linv1[:n0, :n0] - Linv0).max() < 1e-9
    rows_new = Linv1[n0:, :]                       # k x n
    w = Linv1 @ (p.y - hp.prior_mean)
    alpha_upd = np.concatenate([alpha0, np.zeros(k)]) + rows_new.T @ w[n0:]
    assert np.abs(alpha_upd - alpha1).max() < 1e-8 * max(1.0, np.abs(alpha1).max())
    cx, cy = g.cell_xy(np.arange(g.m))
    Ks = O.rbf_np(np.stack([cx, cy], axis=1), p.xy, hp)          # m x n
    U = Ks @ rows_new.T                                          # m x k
    assert np.abs((s20 - (U ** 2).sum(axis=1)) - s21).max() < 1e-9
    assert np.abs((mu0 + U @ w[n0:]) - mu1).max() < 1e-8 * max(1.0, np.abs(mu1).max())
    # separability: U_q as a (rows x columns) GEMM over the observations
    xs = p.x0 + p.rx * np.arange(p.W)
    ys = p.y0 + p.ry * np.arange(p.H)
    A = np.exp(-(xs[:, None] - p.xy[None, :, 0]) ** 2 / (2.0 * hp.length ** 2))          # W x n
    E = np.exp(-(ys[:, None] - p.xy[None, :, 1]) ** 2 / (2.0 * hp.length ** 2))          # H x n
    for q in range(k):
        Pq = hp.sigma_f ** 2 * E * rows_new[q][None, :]                                   # H x n
        Uq = Pq @ A.T                                                                     # H x W, row-major like the grid
        assert np.abs(Uq.reshape(-1) - U[:, q]).max() < 1e-10


def test_skip_bound_is_conservative():
    """The block-sparse cut-off (posterior_tc.cu tc_skip_bits): dropping every K* entry below 2^-b sigma_f^2 moves sigma^2 by
    less than 2 n sqrt(sigma_f^2 / noise_var) 2^-b sigma_f^2.  Checked on a spread-out problem with b small enough to bite."""
    p = S.make_problem("C3", n=300, W=41, H=29)
    p.xy = p.xy * 5.0
    p.rx, p.ry = p.rx * 5.0, p.ry * 5.0
    hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    g = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    _, Linv, alpha = O.gp_fit_scipy(p.xy, p.y, hp)
    cx, cy = g.cell_xy(np.arange(g.m))
    Ks = O.rbf_np(np.stack([cx, cy], axis=1), p.xy, hp)
    V = Ks @ Linv.T
    s2 = hp.sigma_f ** 2 - (V ** 2).sum(axis=1)
    for b in (10, 16, 24):
        Kc = np.where(Ks < 2.0 ** -b * hp.sigma_f ** 2, 0.0, Ks)
        assert (Kc != Ks).any()
        Vc = Kc @ Linv.T
        s2c = hp.sigma_f ** 2 - (Vc ** 2).sum(axis=1)
        bound = 2.0 * p.n * np.sqrt(hp.sigma_f ** 2 / hp.noise_var) * 2.0 ** -b * hp.sigma_f ** 2
        assert np.abs(s2c - s2).max() <= bound


def test_dmma_tile_staging_recurrence():
    """Staging of posterior_thin2d_dmma_kernel (posterior_thin.cu): along a tile edge of 128 cells with pitch h, from three
    exp per (observation, factor) -- e(0), g = exp(-2 d0 h / 2l^2), G_0 = exp(-(32 d0 h + 256 h^2) / 2l^2) -- the anchors
    e(c) = e(0) g^c T[c] (c < 16, g^c from g, g^2, g^4, g^8) and eight recurrence steps e(c + 16 (i + 1)) = e(c + 16 i) G_i,
    G_(i+1) = G_i exp(-512 h^2 / 2l^2) with G_0 scaled by R[c] = exp(-32 c h^2 / 2l^2) reproduce exp(-(d0 + c h)^2 / 2l^2)
    to a few tens of roundings, for observations on either side of the edge and at the reach of the block-sparse walk."""
    length = 0.4
    inv2l2 = 1.0 / (2.0 * length * length)
    worst = 0.0
    for h in (4.0 * length / 128.0, 0.0078 * length, -0.011 * length):      # the host's bound 128 |h| <= 4 l, C5's pitch, a negative pitch
        T = np.exp(-(np.arange(16.0) ** 2) * h * h * inv2l2)
        R = np.exp(-32.0 * np.arange(16.0) * h * h * inv2l2)
        q16 = np.exp(-512.0 * h * h * inv2l2)
        for d0 in (-9.0 * length, -1.3 * length, 0.0, 0.37 * length, 8.5 * length, -13.0):
            e0 = np.exp(-d0 * d0 * inv2l2)
            g1 = np.exp(-2.0 * d0 * h * inv2l2)
            g16 = np.exp(-(32.0 * d0 * h + 256.0 * h * h) * inv2l2)
            pw = [g1, g1 * g1, (g1 * g1) ** 2, ((g1 * g1) ** 2) ** 2]
            got = np.zeros(128)
            for c in range(16):
                e = e0 * T[c]
                for b in range(4):
                    if c >> b & 1:
                        e *= pw[b]
                gg = g16 * R[c]
                for i in range(8):
                    got[c + 16 * i] = e
                    e *= gg
                    gg *= q16
            ref = np.exp(-((d0 + np.arange(128.0) * h) ** 2) * inv2l2)
            live = ref > 1e-280                                                 # (an underflowing anchor zeroes its segment)
            worst = max(worst, float(np.abs(got[live] / ref[live] - 1.0).max()))
    assert worst < 2e-13, worst


def test_pplanes_chunk_test_is_implied_by_the_item_mask():
    """pplanes_kernel (posterior_tc_inc.cuh) leaves out the chunks whose box is farther than the cut-off in y alone from a slab
    of rows; build_item_mask keeps a chunk for an item of that slab when dx^2 + dy^2 <= bits + 0.5 in fp32 with the same dy.
    Whatever an item reads was therefore written: fl(dx^2 + dy^2) >= dy^2 for dx^2 >= 0, in round-to-nearest fp32."""
    rng = np.random.default_rng(7)
    dx = rng.uniform(0.0, 9.0, 200000).astype(np.float32)
    dy = rng.uniform(0.0, 9.0, 200000).astype(np.float32)
    bits = np.float32(36.0) + np.float32(0.5)
    item_keeps = (dx * dx + dy * dy) <= bits             # numpy float32 arithmetic rounds each operation to nearest, like __fmul_rn / __fadd_rn
    plane_written = (dy * dy) <= bits
    assert not np.any(item_keeps & ~plane_written)
    assert np.all((dx * dx + dy * dy) >= dy * dy)


def test_pplanes_row_recurrence():
    """pplanes_kernel (posterior_tc_inc.cuh): the row factor exp(-(y_j - y(g))^2 / 2l^2) along consecutive grid rows g by
    e(g + 1) = e(g) s(g), s(g + 1) = s(g) exp(-2 ry^2 / 2l^2), re-anchored with two exp at every global row that is a
    multiple of 16; a band that starts anywhere reproduces the same values, bit for bit, as one that starts at row 0."""
    length, y0 = 0.4, 0.25
    inv2l2 = 1.0 / (2.0 * length * length)

    def rows(gr0, count, yj, ry):
        out = np.zeros(count)
        q2 = np.exp(-2.0 * ry * ry * inv2l2)

        def anchor(g):
            dy = yj - (y0 + g * ry)
            e = np.exp(-dy * dy * inv2l2)
            return e, (np.exp((2.0 * dy * ry - ry * ry) * inv2l2) if e != 0.0 else 0.0)
        e, st = anchor(gr0 & ~15)
        for _ in range(gr0 & ~15, gr0):
            e, st = e * st, st * q2
        for r in range(count):
            if r > 0 and (gr0 + r) & 15 == 0:
                e, st = anchor(gr0 + r)
            out[r] = e
            e, st = e * st, st * q2
        return out

    for ry in (0.0022, 0.05, -0.013):
        for yj in (0.3, 7.9, -2.0):
            full = rows(0, 96, yj, ry)
            ref = np.exp(-((yj - (y0 + np.arange(96.0) * ry)) ** 2) * inv2l2)
            live = ref > 1e-250
            assert np.abs(full[live] / ref[live] - 1.0).max() < 5e-13
            for start in (5, 16, 37):
                assert np.array_equal(rows(start, 48, yj, ry), full[start:start + 48])
