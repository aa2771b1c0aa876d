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


def _formula(mu32, s232, beta, f_min, noise_add=0.0):
    """The reference's arithmetic (:409-416) in numpy fp64 on the fp32 posterior the step left on the device --
    sqrt, *, -, + are IEEE-exact in numpy, so the step's mask and argmax must equal this bit for bit."""
    mu = mu32.astype(np.float64); v = np.maximum(s232.astype(np.float64), 0.0) + noise_add
    conf = beta * np.sqrt(v)
    q0, q1 = mu - conf, mu + conf
    safe = q0 > f_min
    return q0, q1, safe


def _check_step_sets(res, mask, q0, q1, safe):
    assert np.array_equal(mask != 0, safe) and res.safe_count == int(safe.sum())
    if safe.any():
        u = np.where(safe, q1, -np.inf); w = np.where(safe, q1 - q0, -np.inf)
        assert res.argmax_ucb == (int(np.argmax(u)), float(u.max()))           # np.argmax: first (lowest) index of the max
        assert res.argmax_width == (int(np.argmax(w)), float(w.max()))
    else:
        assert res.argmax_ucb[0] == -1 and res.argmax_width[0] == -1


@pytest.mark.parametrize("shape", [(512, 256, 300), (1024, 1024, 200), (333, 77, 150), (4096, 513, 64)])
@pytest.mark.parametrize("noise_in_sigma", [False, True])
def test_step_sets_equal_reference_formula_bit_for_bit(pkg, shape, noise_in_sigma):
    """fp32 step path: the mask, the safe count and both argmax records must equal the reference's fp64 arithmetic
    applied to the fp32 posterior on the device, bit for bit -- also when f_min sits exactly on a cell's LCB, one ulp
    either side of it, at the extremes, and when beta = 0 makes every width tie (lowest index wins)."""
    W, H, n = shape
    p = S.make_problem("C2", n=n, W=W, H=H)
    with make_ctx(pkg, p, precision="fp32", include_noise_in_sigma=noise_in_sigma) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, W, H)
        ctx.set_observations(p.xy, p.y)
        ctx.step(p.beta, p.f_min)
        mu32 = ctx.mu().astype(np.float32); s232 = ctx.sigma2().astype(np.float32)
        assert np.array_equal(mu32.astype(np.float64), ctx.mu())                # the fp64 read-back is the fp32 value
        noise_add = p.noise_var if noise_in_sigma else 0.0
        q0, _, _ = _formula(mu32, s232, p.beta, p.f_min, noise_add)
        on_cell = float(np.sort(q0)[q0.size // 3])
        cases = [(p.beta, p.f_min), (p.beta, on_cell), (p.beta, float(np.nextafter(on_cell, -np.inf))),
                 (p.beta, float(np.nextafter(on_cell, np.inf))), (p.beta, -1e300), (p.beta, 1e300), (0.0, p.f_min),
                 (1e-30, p.f_min), (37.5, float(np.median(q0)) - 80.0)]
        for beta, f_min in cases:
            q0, q1, safe = _formula(mu32, s232, beta, f_min, noise_add)
            res, mask = ctx.step(beta, f_min)
            _check_step_sets(res, mask, q0, q1, safe)
