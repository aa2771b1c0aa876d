"""GPU tests of the two "next" rows either side of the hot path: N3 -- from the frontier cell to the
published subgoal point (reference src/safe_bayesian_optimization_node.cpp:650-706) -- and N4 -- SpatialMeasurement
ingestion with the spatial_data_size trigger (:552-566, src/turtlesim_spatial_publisher.py:137-145)."""
import numpy as np
import pytest

from conftest import make_ctx
from safe_bayesian_optimization_b200 import _lib as L
from safe_bayesian_optimization_b200 import synthetic as S
from safe_bayesian_optimization_b200.node import OptimizerCore

pytestmark = pytest.mark.gpu


def _problem(n=160, W=200, H=200):
    p = S.make_problem("C1", n=n, W=W, H=H)
    return p, (p.x0, p.y0, p.rx, p.ry, p.W, p.H)


def test_subgoal_point_three_outcomes(pkg, oracle):
    p, grid = _problem()
    core = OptimizerCore(beta=p.beta, f_min=p.f_min, sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var,
                         prior_mean=p.prior_mean, precision="fp64", max_obs=p.n, max_cells=p.m)
    with core.ctx:
        res = core.process_observations(p.xy, p.y, grid)
        assert 0 < res.safe_count < p.m
        Lx, Ly = p.W * p.rx, p.H * p.ry
        poly = [np.array([(0.2 * Lx, 0.2 * Ly), (0.6 * Lx, 0.25 * Ly), (0.55 * Lx, 0.7 * Ly), (0.15 * Lx, 0.6 * Ly)])]
        # (1) goal inside the polygon: the goal itself
        core.goal_point_callback(0.4 * Lx, 0.4 * Ly)
        assert core.ComputeSubgoal(poly) == (0.4 * Lx, 0.4 * Ly)
        # (2) goal outside: GetNextSubgoal's cell, projected by polydist
        core.goal_point_callback(*p.goal)
        frontier = core.FindSafetyContourIndices()
        assert frontier.size > 0
        g = oracle.Grid(*grid)
        mu, s2 = core.ctx.mu(), core.ctx.sigma2()
        q0, q1 = oracle.confidence_intervals_np(mu, np.sqrt(np.maximum(s2, 0.0)), p.beta)
        cell_ref = oracle.get_next_subgoal_np(frontier, g, q0, q1, p.goal)
        kind, xy, cell, dist = core.ctx.subgoal_point(frontier, p.goal, poly)
        assert kind == L.SBO_SUBGOAL_PROJECTED and cell == cell_ref
        cx, cy = g.cell_xy(np.array([cell_ref]))
        ref_xy, ref_d = oracle.polydist(poly[0].tolist(), (float(cx[0]), float(cy[0])))
        assert xy == ref_xy and dist == ref_d
        assert core.ComputeSubgoal(poly) == ref_xy
        # (3) goal outside and no frontier: nothing to publish
        kind, xy, cell, dist = core.ctx.subgoal_point(np.zeros(0, dtype=np.int32), p.goal, poly)
        assert kind == L.SBO_SUBGOAL_NONE and cell == -1
        # empty polygon: nothing is within it; polydist's "infinite" answer comes back
        kind, xy, cell, dist = core.ctx.subgoal_point(frontier, p.goal, [])
        assert kind == L.SBO_SUBGOAL_PROJECTED and cell == cell_ref and xy == (0.0, 0.0) and dist == 100000000.0


@pytest.mark.parametrize("precision", ["fp32", "fp64"])
def test_ingestion_appends_refits_and_matches_a_fit_on_everything(pkg, oracle, coracle, precision):
    p, grid = _problem(n=420, W=96, H=64)
    tol = 1e-5 if precision == "fp32" else 1e-9
    with make_ctx(pkg, p, precision=precision, max_obs=p.n) as ctx:
        ctx.set_grid(*grid)
        size, refit = ctx.ingest(p.xy[:100], p.y[:100])                 # nothing there yet: full fit
        assert (size, refit) == (100, True)
        for i in range(100, 110):                                       # the node's steady state: one at a time
            size, refit = ctx.ingest(p.xy[i:i + 1], p.y[i:i + 1], uncertainty=[0.3])
            assert (size, refit) == (i + 1, False)
        size, refit = ctx.ingest(p.xy[110:160], p.y[110:160])           # tail 60 <= max(64, 100 / 4): append
        assert (size, refit) == (160, False)
        size, refit = ctx.ingest(p.xy[160:170], p.y[160:170])           # tail 70 > 64: re-sorted refit of all 170
        assert (size, refit) == (170, True)
        size, refit = ctx.ingest(p.xy[170:420], p.y[170:420])           # tail 250 > max(64, 170 / 4)
        assert (size, refit) == (420, True)
        assert ctx.n_obs == 420
        res, mask = ctx.step(p.beta, p.f_min)
        hp = oracle.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
        g = oracle.Grid(*grid)
        _, Linv, alpha = oracle.gp_fit_scipy(p.xy, p.y, hp)
        mu_ref, s2_ref = coracle.posterior(p.xy, hp, Linv, alpha, g)
        assert np.abs(ctx.mu() - mu_ref).max() <= tol * np.abs(mu_ref).max()
        assert np.abs(ctx.sigma2() - s2_ref).max() <= tol * p.sigma_f ** 2
        # a bad record: error, nothing ingested
        with pytest.raises(pkg.SboError):
            ctx.ingest([[0.5, float("nan")]], [1.0])
        assert ctx.n_obs == 420
        with pytest.raises(pkg.SboError):                               # capacity
            ctx.ingest(p.xy[:1], p.y[:1])
        assert ctx.ingest(np.zeros((0, 2)), np.zeros(0)) == (420, False)


def test_spatial_data_size_trigger(pkg):
    p, grid = _problem(n=40, W=64, H=64)
    core = OptimizerCore(beta=p.beta, f_min=p.f_min, sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var,
                         prior_mean=p.prior_mean, precision="fp32", max_obs=64, max_cells=p.m)
    with core.ctx:
        core.ctx.set_grid(*grid)
        assert core.spatial_measurement_callback(p.xy[0, 0], p.xy[0, 1], p.y[0]) is True      # 0 -> 1
        assert core.spatial_data_size_callback(1) is False                                     # unchanged: no request
        assert core.spatial_measurement_callback(p.xy[1, 0], p.xy[1, 1], p.y[1]) is True
        assert core.last_spatial_data_size_ == 2
