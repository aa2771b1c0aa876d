"""GPU parity: ComputeSets (reference :399-416) and GetNextSubgoal's selection (:499-550) through the
C-ABI against the pinned oracle.  Integer/byte/index work and the fp64 intervals: bit exact."""
import numpy as np
import pytest

from conftest import make_ctx
from safe_bayesian_optimization_b200 import synthetic as S

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.SboContext(precision="fp64", max_obs=8, max_cells=1 << 22)
    yield c
    c.close()


@pytest.mark.parametrize("m", [0, 1, 3, 4, 5, 1023, 4096, 100003, 1 << 20])
def test_compute_sets_bit_exact(ctx, coracle, m):
    rng = np.random.default_rng(m + 1)
    mu = rng.normal(5, 3, m); std = np.abs(rng.normal(0, 1, m))
    if m > 64:
        mu[5] = np.nan; std[9] = np.inf; std[11] = np.nan; mu[13] = np.inf; mu[17] = -np.inf; std[21] = 0.0
    beta, f_min = 2.0, 4.25
    if m > 100:
        f_min = float((mu - beta * std)[100])     # a cell exactly on the threshold (strict >)
    res, mask, q0, q1 = ctx.compute_sets(mu, std, beta, f_min)
    r0, r1 = coracle.confidence_intervals(mu, std, beta)
    rmask, rcnt = coracle.safe_set(r0, f_min)
    assert np.array_equal(q0, r0, equal_nan=True) and np.array_equal(q1, r1, equal_nan=True)
    assert np.array_equal(mask, rmask) and res.safe_count == rcnt and res.n_cells == m
    ui, uv = coracle.argmax_safe(r1, rmask)
    wi, wv = coracle.argmax_safe(r1 - r0, rmask)
    assert res.argmax_ucb[0] == ui and (ui < 0 or res.argmax_ucb[1] == uv)
    assert res.argmax_width[0] == wi and (wi < 0 or res.argmax_width[1] == wv)


def test_argmax_ties_go_to_lowest_index(ctx):
    m = 50000
    mu = np.full(m, 3.0); std = np.full(m, 0.5)
    mu[[40001, 777, 31000]] = 9.0                   # three equal maxima of the UCB
    std[[45000, 1234]] = 0.75                       # two equal maxima of the width
    res, mask, *_ = ctx.compute_sets(mu, std, 2.0, 0.0)
    assert res.argmax_ucb == (777, 10.0) and res.argmax_width == (1234, 3.0) and res.safe_count == m


def test_empty_safe_set(ctx):
    res, mask, *_ = ctx.compute_sets(np.zeros(1000), np.ones(1000), 2.0, 500.0)   # yaml f_min = 500 on a 0-10 terrain
    assert res.safe_count == 0 and res.argmax_ucb[0] == -1 and res.argmax_width[0] == -1 and not mask.any()


@pytest.mark.parametrize("F", [0, 1, 3, 4, 5, 17, 400, 5000, 70000])
def test_select_subgoal_matches_reference_rule(ctx, oracle, coracle, F):
    W, H = 300, 257
    g = oracle.Grid(0.0, 0.0, 0.013, 0.021, W, H)
    ctx.set_grid(g.x0, g.y0, g.rx, g.ry, W, H)
    rng = np.random.default_rng(F)
    mu = rng.normal(4, 1, g.m); std = np.round(rng.uniform(0, 1, g.m), 2)    # rounded -> many equal widths
    res, mask, q0, q1 = ctx.compute_sets(mu, std, 2.0, 3.0)
    fr = rng.integers(0, g.m, F).astype(np.int32)
    if F >= 5:
        fr[1] = fr[0]; fr[-1] = fr[2]
    goal = (0.9 * W * g.rx, 0.1 * H * g.ry)
    assert ctx.select_subgoal(fr, goal) == coracle.get_next_subgoal(fr, g, q0, q1, goal)


def test_select_subgoal_equidistant(ctx, oracle, coracle):
    g = oracle.Grid(0.0, 0.0, 1.0, 1.0, 5, 5)
    ctx.set_grid(0.0, 0.0, 1.0, 1.0, 5, 5)
    mu = np.zeros(25); std = np.full(25, 0.25)
    fr = np.array([7, 17, 11, 13, 0, 24, 4, 20], dtype=np.int32)
    for bump in (None, 17, 11):
        if bump is not None:
            std[bump] += 0.25
        _, _, q0, q1 = ctx.compute_sets(mu, std, 2.0, -10.0)
        assert ctx.select_subgoal(fr, (2.0, 2.0)) == coracle.get_next_subgoal(fr, g, q0, q1, (2.0, 2.0))


def test_select_rejects_foreign_cells(ctx, pkg):
    ctx.set_grid(0.0, 0.0, 1.0, 1.0, 5, 5)
    ctx.compute_sets(np.zeros(25), np.ones(25), 2.0, -10.0)
    with pytest.raises(pkg.SboError):
        ctx.select_subgoal(np.array([3, 99], dtype=np.int32), (0.0, 0.0))


def test_full_size_properties(ctx):
    """C2-size field: size-independent properties (count == popcount of the mask, idempotence)."""
    m = 1 << 20
    rng = np.random.default_rng(3)
    mu = rng.normal(5, 2, m); std = rng.uniform(0, 1, m)
    a, mask_a, q0, q1 = ctx.compute_sets(mu, std, 2.0, 4.0)
    b, mask_b, *_ = ctx.compute_sets(mu, std, 2.0, 4.0)
    assert np.array_equal(mask_a, mask_b) and a.argmax_ucb == b.argmax_ucb
    assert a.safe_count == int((mask_a == 255).sum()) and set(np.unique(mask_a)) <= {0, 255}
    assert a.argmax_ucb[1] == q1[mask_a == 255].max() and a.argmax_ucb[0] == np.flatnonzero((q1 == a.argmax_ucb[1]) & (mask_a == 255))[0]
