"""Row-band invariance probe: full grid vs a band computed on its own (bitwise), for shapes with an odd tile count per row."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
n, W, H = [int(x) for x in sys.argv[1:4]] if len(sys.argv) > 3 else (1600, 640, 80)
p = S.make_problem("C2", n=n, W=W, H=H)
tpr = (W + 127) // 128
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision="fp32", max_obs=n, max_cells=p.m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    ctx.set_observations(p.xy, p.y)
    ctx.step(p.beta, p.f_min)
    mu, s2 = ctx.mu(), ctx.sigma2()
    ctx.step(p.beta, p.f_min)
    print("repeat of the full grid: mu differs at", int((ctx.mu() != mu).sum()), "cells")
    for r0, r1 in ((3, H - 2), (4, H), (1, 2)):
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H, r0, r1)
        ctx.step(p.beta, p.f_min)
        bm, bs = ctx.mu(), ctx.sigma2()
        d = (bm != mu[r0 * W:r1 * W]).reshape(r1 - r0, W)
        ds = int((bs != s2[r0 * W:r1 * W]).sum())
        tiles = sorted({(int(r), int(c) // 128) for r, c in zip(*np.nonzero(d))})
        desc = [f"(row {r + r0} t{t}: band idx {r * tpr + t} full idx {(r + r0) * tpr + t}, {int(d[r, t * 128:(t + 1) * 128].sum())} cells)" for r, t in tiles[:8]]
        print(f"band {r0}-{r1}: mu differs at {int(d.sum())} cells in {len(tiles)} tiles, s2 at {ds}", " ".join(desc))
