"""GPU parity: blocked fp64 Cholesky / triangular inverse / alpha and the rank-k append against the
oracle (fp64: 1e-10 relative on L^-1 and alpha -- the two sides order their sums differently)."""
import numpy as np
import pytest

from conftest import make_ctx
from safe_bayesian_optimization_b200 import synthetic as S

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("n", [1, 2, 5, 64, 127, 128, 129, 300, 1000, 1024])
def test_factor_matches_oracle(pkg, oracle, coracle, n):
    p = S.make_problem("C2", n=n, W=16, H=16)
    hp = oracle.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    _, Linv, alpha = coracle.gp_fit(p.xy, p.y, hp)
    with make_ctx(pkg, p, precision="fp64") as ctx:
        ctx.set_observations(p.xy, p.y)
        assert ctx.n_obs == n
        order = ctx.order()
        assert sorted(order.tolist()) == list(range(n))
        _, Linv_sorted, _ = coracle.gp_fit(p.xy[order], p.y[order], hp)       # L^-1 is reported in the internal order
        assert _rel(ctx.linv(), Linv_sorted) < TOL and _rel(ctx.alpha(), alpha) < TOL
        assert np.array_equal(np.triu(ctx.linv(), 1), np.zeros((n, n)))


def test_factor_large_tiles_match_scipy(pkg, oracle):
    """n > 2944 takes the 128x128-tile kernels (factor_full picks 64x64 tiles below that); scipy/LAPACK
    is the checker here, the plain-C oracle being O(n^3) single-threaded."""
    n = 3100
    p = S.make_problem("C3", n=n, W=16, H=16)
    hp = oracle.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    with make_ctx(pkg, p, precision="fp64") as ctx:
        ctx.set_observations(p.xy, p.y)
        order = ctx.order()
        _, Linv_sorted, _ = oracle.gp_fit_scipy(p.xy[order], p.y[order], hp)
        _, _, alpha = oracle.gp_fit_scipy(p.xy, p.y, hp)
        assert _rel(ctx.linv(), Linv_sorted) < 1e-9 and _rel(ctx.alpha(), alpha) < 1e-9


@pytest.mark.parametrize("n0,ks", [(0, [1, 3]), (1, [1]), (100, [1, 1, 16]), (250, [16, 128, 5]), (120, [200])])
def test_append_equals_full_refactor(pkg, oracle, coracle, n0, ks):
    total = n0 + sum(ks)
    p = S.make_problem("C3", n=total, W=16, H=16)
    hp = oracle.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    with make_ctx(pkg, p, precision="fp64", max_obs=total) as ctx:
        ctx.set_observations(p.xy[:n0], p.y[:n0])
        n = n0
        for k in ks:
            ctx.append_observations(p.xy[n:n + k], p.y[n:n + k])
            n += k
            order = ctx.order()
            assert sorted(order.tolist()) == list(range(n)) and np.array_equal(order[n - k:], np.arange(n - k, n))
            _, Linv, _ = coracle.gp_fit(p.xy[order], p.y[order], hp)
            _, _, alpha = coracle.gp_fit(p.xy[:n], p.y[:n], hp)
            assert ctx.n_obs == n
            assert _rel(ctx.linv(), Linv) < 1e-9 and _rel(ctx.alpha(), alpha) < 1e-9


def test_errors_and_state(pkg):
    p = S.make_problem("C2", n=16, W=8, H=8)
    with make_ctx(pkg, p, precision="fp64", max_obs=16) as ctx:
        with pytest.raises(pkg.SboError) as e:                       # capacity
            ctx.set_observations(np.zeros((17, 2)), np.zeros(17))
        assert e.value.status == -5
        bad = p.xy.copy(); bad[3, 1] = np.nan
        with pytest.raises(pkg.SboError) as e:                       # NaN input
            ctx.set_observations(bad, p.y)
        assert e.value.status == -1
        with pytest.raises(pkg.SboError) as e:                       # step before grid
            ctx.step(2.0, 0.0)
        assert e.value.status == -6
        ctx.set_observations(p.xy, p.y)
        with pytest.raises(pkg.SboError) as e:
            ctx.append_observations(p.xy[:1], p.y[:1])                # full
        assert e.value.status == -5 and ctx.n_obs == 16
    # non positive definite: duplicate point, zero noise
    with pkg.SboContext(noise_var=0.0, precision="fp64", max_obs=8, max_cells=64) as ctx:
        xy = np.array([[0.1, 0.1], [0.5, 0.5], [0.1, 0.1]])
        with pytest.raises(pkg.SboError) as e:
            ctx.set_observations(xy, np.zeros(3))
        assert e.value.status == -4 and "positive definite at row" in str(e.value)
        assert ctx.n_obs == 0
        ctx.set_observations(xy[:2], np.array([0.3, -0.2]))           # context still usable
        ctx.set_grid(0.0, 0.0, 0.1, 0.1, 8, 8)
        alpha0, linv0 = ctx.alpha(), ctx.linv()
        before, _ = ctx.step(2.0, -10.0)
        mu0 = ctx.mu()
        with pytest.raises(pkg.SboError) as e:
            ctx.append_observations(xy[2:], np.array([0.3]))
        assert e.value.status == -4 and ctx.n_obs == 2
        # the failed append leaves the last good state: same factor, alpha put back, a step gives the same posterior
        assert np.array_equal(ctx.linv(), linv0) and np.allclose(ctx.alpha(), alpha0, rtol=1e-13, atol=0)
        after, _ = ctx.step(2.0, -10.0)
        assert np.allclose(ctx.mu(), mu0, rtol=1e-13, atol=1e-15) and after.safe_count == before.safe_count
