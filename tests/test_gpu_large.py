"""GPU parity at the sizes of BASELINE.json's large configs (C4: 16 384 observations, fp32, block-sparse + graded walk;
C5: 8 192 observations, fp64) on sampled cells against the fp64 scipy restatement (oracle/parity.py),
the refresh / drift policy of the incremental loop, long incremental chains, and the "last good state" behaviour of
failed fits and appends."""
import numpy as np
import pytest

from conftest import make_ctx
from safe_bayesian_optimization_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def _bands(p, rows=8):
    """bottom edge, the clustered start region at H / 4, mid-domain"""
    return [(0, rows), (p.H // 4 - rows // 2, p.H // 4 - rows // 2 + rows), (p.H // 2, p.H // 2 + rows)]


def _band_step(ctx, p, r0, r1, cells, incremental=False):
    res, mask = ctx.step(p.beta, p.f_min, incremental=incremental)
    loc = cells - r0 * p.W
    return res, ctx.mu()[loc], ctx.sigma2()[loc], mask[loc]


def test_c4_sampled_parity(pkg):
    """C4 geometry at full n (16 384 observations, 8192 x 8192 grid): three bands of 8 grid rows, 4096 sampled cells each.
    The default plan: block-sparse walk with the cut-off from the MEASURED row norm of L^-1, graded split depth, strip-major
    tiles.  Gates: north_star's 1e-5 for mu (relative) and for sigma^2 relative to sigma_f^2, no mask mismatch outside the
    tolerance band; the H1 metric (relative to sigma^2 itself) at the level include/sbo.h states for SBO_FP32.  Also the
    dense, ungraded walk on one band: the two agree far below the tolerance (the skipped and hi.hi-only terms are bounded
    by 1e-7 sigma_f^2)."""
    from oracle import parity as P
    p = S.make_problem("C4")
    ref = P.SampledReference(p)
    rng = np.random.default_rng(11)
    bands = _bands(p)
    cells = P.sample_cells(p, bands, 4096, rng)
    blocks = []
    with make_ctx(pkg, p, max_cells=8 * p.W) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H, *bands[0])
        ctx.set_observations(p.xy, p.y)
        for (r0, r1), c in zip(bands, cells):
            ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H, r0, r1)
            res, mu, s2, mask = _band_step(ctx, p, r0, r1, c)
            plan = ctx.tc_plan_detail()
            assert res.kernel_used == 2 and ctx.tc_plan()[1] is True and plan["linv_row_l1"] > 0 and plan["ring_bits"] < plan["bits"]
            mu_ref, s2_ref = ref.evaluate(c)
            b = P.parity_block(p, mu, s2, mask, mu_ref, s2_ref, 1e-5)
            print(f"C4 rows {r0}-{r1}: {b} plan {plan}")
            blocks.append(b)
            if r0 == bands[1][0]:
                full = (ctx.mu().copy(), ctx.sigma2().copy())
                ctx.set_option("tc_skip", 0)
                ctx.step(p.beta, p.f_min)
                assert np.abs(ctx.sigma2() - full[1]).max() <= 2e-6 and np.abs(ctx.mu() - full[0]).max() <= 2e-5
                ctx.set_option("tc_skip", 1)
    tot = P.merge_blocks(blocks)
    print("C4 parity:", tot)
    assert tot["mu_rel_max"] <= 1e-5 and tot["s2_abs_max_over_sf2"] <= 1e-5 and tot["mask_mismatch_outside_band"] == 0
    assert tot["s2_h1_max"] <= 5e-4 and tot["lcb_abs_max_over_mu_scale"] <= 2e-5


def test_c5_sampled_parity(pkg):
    """C5 geometry at full n (8 192 observations, fp64, 4096 x 4096 grid): the DMMA full refresh on three bands of 4 rows
    at 1e-10, then a chain of appended single observations through the incremental kernels against a refit reference."""
    from oracle import parity as P
    p = S.make_problem("C5")
    k_steps = 3
    xy2, y2 = S.extra_observations(p, k_steps, 0)
    ref0 = P.SampledReference(p)
    ref1 = P.SampledReference(p, np.concatenate([p.xy, xy2]), np.concatenate([p.y, y2]))
    rng = np.random.default_rng(12)
    bands = _bands(p, rows=4)
    cells = P.sample_cells(p, bands, 4096, rng)
    blocks0, blocks1 = [], []
    for (r0, r1), c in zip(bands, cells):
        with make_ctx(pkg, p, max_obs=p.n + k_steps, max_cells=4 * p.W) as ctx:
            ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H, r0, r1)
            ctx.set_observations(p.xy, p.y)
            res, mu, s2, mask = _band_step(ctx, p, r0, r1, c)
            assert res.kernel_used == 4
            blocks0.append(P.parity_block(p, mu, s2, mask, *ref0.evaluate(c), 1e-10))
            for t in range(k_steps):
                ctx.append_observations(xy2[t:t + 1], y2[t:t + 1])
                res, mu, s2, mask = _band_step(ctx, p, r0, r1, c, incremental=True)
                assert res.kernel_used == 3
            blocks1.append(P.parity_block(p, mu, s2, mask, *ref1.evaluate(c), 1e-10))
    for name, blocks in (("full", blocks0), ("after 3 incremental steps", blocks1)):
        tot = P.merge_blocks(blocks)
        print(f"C5 parity ({name}):", tot)
        assert tot["mu_rel_max"] <= 1e-10 and tot["s2_abs_max_over_sf2"] <= 1e-10 and tot["mask_mismatch_outside_band"] == 0
        assert tot["s2_h1_max"] <= 1e-8


@pytest.mark.parametrize("precision", ["fp32", "fp64"])
def test_refresh_and_drift_policy(pkg, oracle, coracle, precision):
    """sbo_step_incremental under the refresh policy: every `refresh_every`-th incremental step is a full
    posterior (refreshed = 1, the full-refresh kernel), every `drift_check_every`-th one re-evaluates sample rows exactly
    and reports the deviation; results meet the oracle throughout."""
    from test_gpu_posterior import reference, check_step
    n0, steps = 300, 12
    p = S.make_problem("C3", n=n0 + steps, W=300, H=19)
    full_kernel = 2 if precision == "fp32" else 4
    with make_ctx(pkg, p, precision=precision, max_obs=n0 + steps,
                  options={"refresh_every": 5, "drift_check_every": 2, "drift_rows": 3}) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy[:n0], p.y[:n0])
        first, _ = ctx.step(p.beta, p.f_min)
        assert first.refreshed == 0 and first.drift == -1.0
        seen_refresh = seen_check = 0
        since = 0
        for t in range(steps):
            ctx.append_observations(p.xy[n0 + t:n0 + t + 1], p.y[n0 + t:n0 + t + 1])
            res, mask = ctx.step(p.beta, p.f_min, incremental=True)
            since += 1
            if res.refreshed:
                assert res.kernel_used == full_kernel and res.drift == -1.0
                seen_refresh += 1
                since = 0
            else:
                assert res.kernel_used == 3
                if since % 2 == 0:
                    assert 0.0 <= res.drift <= (1e-5 if precision == "fp32" else 1e-10)
                    seen_check += 1
                else:
                    assert res.drift == -1.0
        assert seen_refresh >= 2 and seen_check >= 3
        q = S.make_problem("C3", n=n0 + steps, W=300, H=19)
        ref = reference(oracle, coracle, q)
        check_step(q, ref, res, mask, ctx.mu(), ctx.sigma2(), precision, label="policy")


@pytest.mark.parametrize("precision,steps", [("fp32", 320), ("fp64", 320)])
def test_hundreds_of_incremental_steps(pkg, oracle, coracle, precision, steps):
    """Several hundred appended observations, one per step, under the DEFAULT policy (fp32: a full refresh every 256
    incremental steps, a drift check every 16; fp64: drift checks only): the end of the chain meets the oracle of a refit."""
    from test_gpu_posterior import reference, check_step, check_argmax
    n0 = 200
    p = S.make_problem("C3", n=n0 + steps, W=300, H=5)
    with make_ctx(pkg, p, precision=precision, max_obs=n0 + steps) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy[:n0], p.y[:n0])
        ctx.step(p.beta, p.f_min)
        refreshed, worst = 0, 0.0
        for t in range(steps):
            ctx.append_observations(p.xy[n0 + t:n0 + t + 1], p.y[n0 + t:n0 + t + 1])
            res, mask = ctx.step(p.beta, p.f_min, incremental=True)
            refreshed += res.refreshed
            worst = max(worst, res.drift)
        print(f"[{precision}] {steps} incremental steps: {refreshed} refreshed, worst drift estimate {worst:.2e}")
        assert refreshed >= (1 if precision == "fp32" else 0)
        ref = reference(oracle, coracle, p)
        band = check_step(p, ref, res, mask, ctx.mu(), ctx.sigma2(), precision, label=f"chain{steps}")
        check_argmax(res, ref, band, mask, p.f_min)


def _sparse_problem(n, W, H, spread=12.0):
    """Observations several length-scales apart: K stays numerically positive definite with (almost) no noise."""
    p = S.make_problem("C3", n=n, W=W, H=H)
    p.xy = p.xy * spread
    p.rx, p.ry = p.rx * spread, p.ry * spread
    return p


def test_failed_append_leaves_rows_clean_fp32(pkg):
    """a rank-k append that fails (duplicate point, noise ~ 0) used to leave rows [n, n + k) of L^-1 and of its
    fp16 planes written from garbage; with n > 256 and n % 32 != 0 the pair kernel then squared them into |v|^2.  The failed
    append must leave the factor state exactly as it was: a full step afterwards is bit-identical to the one before."""
    n = 300
    p = _sparse_problem(n, 384, 6)
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=1e-20, prior_mean=p.prior_mean, precision="fp32",
                        kernel="tcgen05", max_obs=n + 8, max_cells=p.m) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        before, mask0 = ctx.step(p.beta, p.f_min)
        mu0, s20, linv0 = ctx.mu(), ctx.sigma2(), ctx.linv()
        dup = np.concatenate([p.xy[:1], p.xy[5:6] + 40.0, p.xy[:1]])       # a duplicate of observation 0 (twice)
        with pytest.raises(pkg.SboError) as e:
            ctx.append_observations(dup, np.zeros(3))
        assert e.value.status == -4 and ctx.n_obs == n
        assert np.array_equal(ctx.linv(), linv0)
        after, mask1 = ctx.step(p.beta, p.f_min)
        assert np.array_equal(ctx.mu(), mu0) and np.array_equal(ctx.sigma2(), s20) and np.array_equal(mask0, mask1)
        # ... and a later, smaller, successful append works from clean rows
        good = p.xy[7:8] + np.array([[33.0, 21.0]])
        ctx.append_observations(good, np.array([p.prior_mean]))
        inc, _ = ctx.step(p.beta, p.f_min, incremental=True)
        s2_inc = ctx.sigma2()
        ctx.step(p.beta, p.f_min)
        assert np.abs(ctx.sigma2() - s2_inc).max() <= 1e-5


@pytest.mark.parametrize("precision", ["fp32", "fp64"])
def test_failed_fit_restores_previous_observations(pkg, precision):
    """SURVEY B4 "ctx left in last good state": sbo_set_observations that fails (not positive definite) is followed by a
    refit of the observations held before the call -- same count, same factor, same posterior bit for bit."""
    n = 280
    p = _sparse_problem(n, 256, 5)
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=1e-20, prior_mean=p.prior_mean, precision=precision,
                        max_obs=n + 8, max_cells=p.m) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        ctx.step(p.beta, p.f_min)
        mu0, s20, alpha0, order0 = ctx.mu(), ctx.sigma2(), ctx.alpha(), ctx.order()
        bad = np.concatenate([p.xy, p.xy[:2]])
        with pytest.raises(pkg.SboError) as e:
            ctx.set_observations(bad, np.concatenate([p.y, p.y[:2]]))
        assert e.value.status == -4 and "restored" in str(e.value)
        assert ctx.n_obs == n and np.array_equal(ctx.order(), order0) and np.array_equal(ctx.alpha(), alpha0)
        assert np.array_equal(ctx.mu(), mu0)                               # the stored posterior is still that set's
        ctx.step(p.beta, p.f_min)
        assert np.array_equal(ctx.mu(), mu0) and np.array_equal(ctx.sigma2(), s20)
        with pytest.raises(pkg.SboError):                                  # ingestion of a duplicate: the append fails, state kept
            ctx.ingest(p.xy[:1], p.y[:1])
        assert ctx.n_obs == n
