"""One incremental step through the thin kernel (for ncu): argv = precision n W H k."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
prec, n, W, H, k = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
p = S.make_problem("C4", n=n, W=W, H=H)
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision=prec, max_obs=n + 64, max_cells=p.m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    ctx.set_observations(p.xy, p.y)
    ctx.step(p.beta, p.f_min, want_mask=False)
    for t in range(3):
        xy2, y2 = S.extra_observations(p, k, t)
        ctx.append_observations(xy2, y2)
        inc, _ = ctx.step(p.beta, p.f_min, incremental=True, want_mask=False)
    print(f"{prec} n={n} {W}x{H} k={k}: {inc.ms_posterior:.3f} ms kernel={inc.kernel_used}")
