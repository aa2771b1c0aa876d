"""One context over several GPUs (sbo_group): a group of bands driven from one caller thread must give
the single context's safe mask image, counts, argmax cells and subgoal.  With one GPU the members share device 0 (the
band logic, the one-image mask assembly and the host-side fold are the same); with >= 2 GPUs they sit on distinct devices."""
import os
import subprocess

import numpy as np
import pytest

from conftest import make_ctx
from safe_bayesian_optimization_b200 import synthetic as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _devices(members):
    import torch
    n = torch.cuda.device_count()
    return [i % n for i in range(members)]


@pytest.mark.parametrize("precision,members", [("fp32", 2), ("fp32", 3), ("fp64", 2)])
def test_group_equals_single_context(pkg, precision, members):
    p = S.make_problem("C3", n=1500, W=512, H=96)
    p.xy = p.xy * 3.0; p.rx *= 3.0; p.ry *= 3.0            # a domain wide enough for the block-sparse walk and unequal bands
    with make_ctx(pkg, p, precision=precision, max_obs=p.n + 4) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        one, mask1 = ctx.step(p.beta, p.f_min)
        mu1, s21 = ctx.mu(), ctx.sigma2()
        with pkg.SboGroup(_devices(members), sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var,
                          prior_mean=p.prior_mean, precision=precision, max_obs=p.n + 4, max_cells=p.m) as g:
            g.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
            g.set_observations(p.xy, p.y)
            cuts = g.bands()
            assert cuts[0] == 0 and cuts[-1] == p.H and len(cuts) == members + 1
            res, mask = g.step(p.beta, p.f_min)
            assert np.array_equal(mask, mask1) and res.safe_count == one.safe_count and res.n_cells == p.m
            assert res.argmax_ucb[0] == one.argmax_ucb[0] and res.argmax_width[0] == one.argmax_width[0]
            tol = 2e-6 if precision == "fp32" else 1e-12     # (a band that pairs tiles differently: last-ulp differences in mu)
            assert np.abs(g.mu() - mu1).max() <= tol * np.abs(mu1).max() and np.abs(g.sigma2() - s21).max() <= tol
            fr = np.flatnonzero(mask)[::53].astype(np.int32)  # a "frontier" that crosses every band
            assert g.select_subgoal(fr, p.goal) == ctx.select_subgoal(fr, p.goal)
            assert g.select_subgoal(np.zeros(0, np.int32), p.goal) == -1
            # steady state: one new observation, incremental step on every band
            xy2, y2 = S.extra_observations(p, 1, 0)
            g.append_observations(xy2 * 3.0, y2); ctx.append_observations(xy2 * 3.0, y2)
            ri, mi = g.step(p.beta, p.f_min, incremental=True)
            r1, m1 = ctx.step(p.beta, p.f_min, incremental=True)
            assert np.array_equal(mi, m1) and ri.safe_count == r1.safe_count and ri.n_obs == p.n + 1 and g.n_obs == p.n + 1


@pytest.mark.parametrize("precision,members,n", [("fp64", 2, 700), ("fp32", 3, 1100), ("fp32", 8, 1500), ("fp64", 4, 129),
                                                 ("fp32", 2, 3100), ("fp64", 3, 3300)])
def test_shared_fit_equals_replicated(pkg, precision, members, n):
    """SBO_OPT_DIST_FIT: block columns of the Cholesky / L^-1 work dealt to the members (snake order), panel exchange per
    step, all-gather of L^-1 at the end.  Every member must end with the same L^-1 and alpha as every other member, bit for
    bit, and with the replicated fit's: bit for bit where that runs the same 128 x 128 DMMA tile routine (n > 2944), to
    rounding where it runs the 64 x 64 FMA tiles (small n: a different summation order inside a tile)."""
    p = S.make_problem("C3", n=n, W=256, H=24)
    with make_ctx(pkg, p, precision=precision, max_obs=p.n + 4) as ctx:
        ctx.set_observations(p.xy, p.y)
        linv1, alpha1 = ctx.linv(), ctx.alpha()
        assert ctx.fit_info()[1] == 0
    same_routine = n > 2944

    def check(linv, alpha, who):
        if same_routine:
            assert np.array_equal(linv, linv1), f"{who}: L^-1 differs from the replicated fit"
            assert np.array_equal(alpha, alpha1)
        else:
            assert np.abs(linv - linv1).max() <= 1e-11 * np.abs(linv1).max(), who
            assert np.abs(alpha - alpha1).max() <= 1e-9 * np.abs(alpha1).max(), who

    with pkg.SboGroup(_devices(members), sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var,
                      prior_mean=p.prior_mean, precision=precision, max_obs=p.n + 4, max_cells=p.m) as g:
        g.set_option("dist_fit", 1)
        g.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        for rep in range(2):                                 # twice: the rendezvous and the events are reused
            g.set_observations(p.xy, p.y)
            first = None
            for i in range(members):
                linv, alpha, (ms, shared) = g.member_factor(i)
                assert shared == members and ms > 0
                check(linv, alpha, f"member {i}")
                if first is None:
                    first = (linv, alpha)
                else:
                    assert np.array_equal(linv, first[0]) and np.array_equal(alpha, first[1]), f"member {i} != member 0"
        g.set_option("dist_fit", 0)
        g.set_observations(p.xy, p.y)
        assert g.member_factor(members - 1)[2][1] == 0 and np.array_equal(g.member_factor(0)[0], linv1)


def test_shared_fit_not_positive_definite(pkg):
    """A failing block in the middle of a shared fit: every member reports the same first failing row and goes back to the
    observations it held before."""
    p = S.make_problem("C3", n=600, W=256, H=12)
    p.xy = p.xy * 12.0; p.rx *= 12.0; p.ry *= 12.0
    with pkg.SboGroup(_devices(3), sigma_f=p.sigma_f, length_scale=p.length, noise_var=1e-20, prior_mean=p.prior_mean,
                      precision="fp64", max_obs=p.n + 4, max_cells=p.m) as g:
        g.set_option("dist_fit", 1)
        g.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        g.set_observations(p.xy, p.y)
        a, mask_a = g.step(p.beta, p.f_min)
        linv_a = g.member_factor(1)[0]
        with pytest.raises(pkg.SboError) as e:
            g.set_observations(np.concatenate([p.xy, p.xy[300:302]]), np.concatenate([p.y, p.y[300:302]]))
        assert e.value.status == -4 and g.n_obs == p.n
        assert np.array_equal(g.member_factor(2)[0], linv_a)
        b, mask_b = g.step(p.beta, p.f_min)
        assert np.array_equal(mask_a, mask_b) and a.argmax_ucb == b.argmax_ucb


def test_group_failed_fit_keeps_last_good_state(pkg):
    p = S.make_problem("C3", n=280, W=256, H=12)
    p.xy = p.xy * 12.0; p.rx *= 12.0; p.ry *= 12.0
    with pkg.SboGroup(_devices(2), sigma_f=p.sigma_f, length_scale=p.length, noise_var=1e-20, prior_mean=p.prior_mean,
                      precision="fp64", max_obs=p.n + 4, max_cells=p.m) as g:
        g.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        g.set_observations(p.xy, p.y)
        a, mask_a = g.step(p.beta, p.f_min)
        with pytest.raises(pkg.SboError) as e:
            g.set_observations(np.concatenate([p.xy, p.xy[:2]]), np.concatenate([p.y, p.y[:2]]))   # duplicates, noise ~ 0
        assert e.value.status == -4 and g.n_obs == p.n
        b, mask_b = g.step(p.beta, p.f_min)
        assert np.array_equal(mask_a, mask_b) and a.argmax_ucb == b.argmax_ucb


def test_cpp_group_host(pkg):
    """host_example/group_main.cpp: the single-process C++ host over N members (distinct GPUs when the box has them)."""
    import torch
    pkg.build()
    subprocess.run(["make", "-C", os.path.join(ROOT, "host_example")], check=True, capture_output=True)
    n = torch.cuda.device_count()
    args = [str(max(2, min(n, 8)))] + ([] if n >= 2 else ["--same-device"])
    r = subprocess.run([os.path.join(ROOT, "host_example", "group_example")] + args, capture_output=True, text=True, timeout=600)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and "group_example OK" in r.stdout
