"""Small steps through every kernel family (for compute-sanitizer): fp32 tcgen05 full + incremental (thin, k = 1 and
row-block, k = 12), fp64 DMMA full + incremental, sets, frontier, compaction, select."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
for prec in ("fp32", "fp64"):
    p = S.make_problem("C2", n=300, W=200, H=21)
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                        precision=prec, max_obs=400, max_cells=p.m) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        res, mask = ctx.step(p.beta, p.f_min)
        for k in (1, 12):
            xy2, y2 = S.extra_observations(p, k, k)
            ctx.append_observations(xy2, y2)
            res, mask = ctx.step(p.beta, p.f_min, incremental=True)
        img, cpx, rpy = ctx.frontier_image()
        idx, cnt = ctx.compact_cells(border=True)
        cell = ctx.select_subgoal(idx[:50], p.goal) if cnt else -1
        print(prec, res.kernel_used, res.safe_count, cnt, cell)
# fp64, one new row, >= one 128 x 128 tile per SM: the 2-D register-tile kernel; fp32 block-sparse pair kernel (spread-out domain)
p = S.make_problem("C5", n=41, W=1990, H=1250)
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision="fp64", max_obs=64, max_cells=p.m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    ctx.set_observations(p.xy[:40], p.y[:40])
    ctx.step(p.beta, p.f_min, want_mask=False)
    ctx.append_observations(p.xy[40:41], p.y[40:41])
    res, _ = ctx.step(p.beta, p.f_min, incremental=True, want_mask=False)
    print("fp64 2-D tile", res.kernel_used, res.safe_count)
p = S.make_problem("C3", n=1100, W=384, H=6)
p.xy = p.xy * 6.0; p.rx *= 6.0; p.ry *= 6.0   # 6x wider domain: most chunks are out of reach of a tile
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision="fp32", max_obs=1100, max_cells=p.m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    ctx.set_observations(p.xy, p.y)
    res, _ = ctx.step(p.beta, p.f_min, want_mask=False)
    print("fp32 block-sparse", res.kernel_used, ctx.tc_plan(), res.safe_count)
