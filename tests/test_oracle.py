"""CPU tests of the oracle itself: three independent statements agree, closed-form known answers hold,
the pinned arithmetic (reference :409-416, :499-550) matches a second numpy restatement bit for bit,
and the committed golden fixtures are reproduced."""
import glob
import os

import numpy as np
import pytest

from safe_bayesian_optimization_b200 import synthetic as S

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _setup(oracle, name="C2", n=150, W=48, H=32):
    p = S.make_problem(name, n=n, W=W, H=H)
    return p, oracle.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean), oracle.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)


def test_three_oracles_agree(oracle, coracle):
    p, hp, g = _setup(oracle, n=300, W=64, H=64)
    _, Linv, alpha = coracle.gp_fit(p.xy, p.y, hp)
    mu, s2 = coracle.posterior(p.xy, hp, Linv, alpha, g)
    _, Linv2, alpha2 = oracle.gp_fit_scipy(p.xy, p.y, hp)
    mu2, s22 = oracle.gp_posterior_scipy(p.xy, hp, Linv2, alpha2, g)
    mu3, s23 = oracle.gp_posterior_sklearn(p.xy, p.y, hp, g, np.arange(g.m))
    assert np.abs(Linv - Linv2).max() < 1e-11 and np.abs(alpha - alpha2).max() < 1e-10
    for a, b in ((mu, mu2), (mu, mu3)):
        assert np.abs(a - b).max() < 1e-11
    for a, b in ((s2, s22), (s2, s23)):
        assert np.abs(a - b).max() < 1e-12


def test_known_answers_n0_n1(oracle, coracle):
    hp = oracle.Hyper(1.3, 0.4, 0.1, 2.0)
    g = oracle.Grid(0.0, 0.0, 0.1, 0.1, 8, 4)
    # n = 0: prior
    mu, s2 = coracle.posterior(np.zeros((0, 2)), hp, np.zeros((0, 0)), np.zeros(0), g)
    assert np.all(mu == 2.0) and np.allclose(s2, 1.3 ** 2, rtol=0, atol=1e-15)
    # n = 1: closed form
    xy = np.array([[0.3, 0.2]]); y = np.array([3.5])
    _, Linv, alpha = coracle.gp_fit(xy, y, hp)
    mu, s2 = coracle.posterior(xy, hp, Linv, alpha, g)
    cx, cy = g.cell_xy()
    k = 1.3 ** 2 * np.exp(-((cx - 0.3) ** 2 + (cy - 0.2) ** 2) / (2 * 0.4 ** 2))
    assert np.allclose(mu, 2.0 + k * (3.5 - 2.0) / (1.3 ** 2 + 0.1), rtol=0, atol=1e-13)
    assert np.allclose(s2, 1.3 ** 2 - k ** 2 / (1.3 ** 2 + 0.1), rtol=0, atol=1e-13)
    # the cell that coincides with the observation: sigma^2 = sf2*noise/(sf2+noise)
    c = 2 * 8 + 3
    assert abs(s2[c] - 1.3 ** 2 * 0.1 / (1.3 ** 2 + 0.1)) < 1e-12


def test_cholesky_rejects_non_posdef(oracle, coracle):
    hp = oracle.Hyper(1.0, 0.4, 0.0, 0.0)
    xy = np.array([[0.1, 0.1], [0.5, 0.5], [0.1, 0.1]])   # duplicate point, zero noise
    with pytest.raises(np.linalg.LinAlgError):
        coracle.gp_fit(xy, np.zeros(3), hp)


def test_pinned_ci_and_safe_set_bit_exact(oracle, coracle):
    rng = np.random.default_rng(7)
    mu = rng.normal(5, 3, 10007); std = np.abs(rng.normal(0, 1, 10007))
    mu[5] = np.nan; std[9] = np.inf; std[11] = np.nan; mu[13] = np.inf
    q0, q1 = coracle.confidence_intervals(mu, std, 2.0)
    q0n, q1n = oracle.confidence_intervals_np(mu, std, 2.0)
    assert np.array_equal(q0, q0n, equal_nan=True) and np.array_equal(q1, q1n, equal_nan=True)
    f_min = float(q0[100])          # exactly on the threshold: strict > means unsafe
    mask, cnt = coracle.safe_set(q0, f_min)
    maskn, cntn = oracle.safe_set_np(q0, f_min)
    assert np.array_equal(mask, maskn) and cnt == cntn
    assert mask[100] == 0 and mask[5] == 0 and mask[11] == 0
    assert set(np.unique(mask)) <= {0, 255}


@pytest.mark.parametrize("F", [0, 1, 3, 4, 5, 17, 400])
def test_pinned_subgoal_selection(oracle, coracle, F):
    p, hp, g = _setup(oracle)
    rng = np.random.default_rng(F)
    q0 = rng.normal(4, 1, g.m); q1 = q0 + np.round(rng.uniform(0, 2, g.m), 1)   # rounded widths -> many ties
    fr = rng.integers(0, g.m, F).astype(np.int32)
    if F >= 5:
        fr[1] = fr[0]; fr[-1] = fr[2]                # duplicates, as CHAIN_APPROX_NONE produces
    a = coracle.get_next_subgoal(fr, g, q0, q1, p.goal)
    b = oracle.get_next_subgoal_np(fr, g, q0, q1, p.goal)
    assert a == b
    if F == 0:
        assert a == -1
    else:
        # brute-force restatement with python's stable sort on (dist, position)
        x, y = g.cell_xy(fr)
        d = np.sqrt((x - p.goal[0]) * (x - p.goal[0]) + (y - p.goal[1]) * (y - p.goal[1]))
        order = sorted(range(F), key=lambda i: (d[i], i))[:max(1, F // 4)]
        best = max(order, key=lambda i: (q1[fr[i]] - q0[fr[i]], -order.index(i)))
        assert a == fr[best]


def test_subgoal_equidistant_ties_break_on_position(oracle, coracle):
    g = oracle.Grid(0.0, 0.0, 1.0, 1.0, 5, 5)
    q0 = np.zeros(25); q1 = np.ones(25)
    fr = np.array([1 * 5 + 2, 3 * 5 + 2, 2 * 5 + 1, 2 * 5 + 3, 0, 24, 4, 20], dtype=np.int32)   # 4 cells at distance 1 of (2,2)
    # T = 2: the two first equidistant entries by list position; equal widths -> the first wins
    assert coracle.get_next_subgoal(fr, g, q0, q1, (2.0, 2.0)) == fr[0]
    q1[fr[1]] = 1.5
    assert coracle.get_next_subgoal(fr, g, q0, q1, (2.0, 2.0)) == fr[1]
    q1[fr[2]] = 9.0                                   # outside the closest quarter: ignored
    assert coracle.get_next_subgoal(fr, g, q0, q1, (2.0, 2.0)) == fr[1]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(x) for x in GOLDEN])
def test_golden_fixtures_reproduced(oracle, coracle, path):
    z = np.load(path)
    sf, ell, nv, pm = z["hyper"]
    hp = oracle.Hyper(sf, ell, nv, pm)
    x0, y0, rx, ry = z["grid"]
    g = oracle.Grid(x0, y0, rx, ry, int(z["W"]), int(z["H"]))
    # independent path: scipy factorisation + BLAS posterior
    _, Linv, alpha = oracle.gp_fit_scipy(z["xy"], z["y"], hp)
    mu, s2 = oracle.gp_posterior_scipy(z["xy"], hp, Linv, alpha, g)
    assert np.abs(mu - z["mu"]).max() < 1e-10 and np.abs(s2 - z["s2"]).max() < 1e-12
    # pinned part, from the stored posterior: bit exact
    q0, q1 = coracle.confidence_intervals(z["mu"], np.sqrt(np.maximum(z["s2"], 0.0)), float(z["beta"]))
    assert np.array_equal(q0, z["q0"]) and np.array_equal(q1, z["q1"])
    mask, cnt = coracle.safe_set(q0, float(z["f_min"]))
    assert np.array_equal(mask, z["mask"]) and cnt == int(z["safe_count"])
    fr = oracle.find_safety_contour_indices(mask, g)
    assert np.array_equal(fr, z["frontier"])
    assert coracle.get_next_subgoal(fr, g, q0, q1, z["goal"]) == int(z["subgoal"])
