"""Accuracy probe of the posterior kernels against the fp64 oracle (prints one line per case)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
from oracle import oracle as O

tag = sys.argv[1] if len(sys.argv) > 1 else ""
ns = [int(x) for x in (sys.argv[2].split(",") if len(sys.argv) > 2 else "256,1024,2048".split(","))]
kernels = sys.argv[3].split(",") if len(sys.argv) > 3 else ["tcgen05"]
co = O.COracle()
for n in ns:
    p = S.make_problem("C2", n=n, W=256, H=16)
    if os.environ.get("SORT") == "morton":
        q = np.clip((p.xy / (p.xy.max() + 1e-9) * 65535).astype(np.uint64), 0, 65535)
        def spread(v):
            v = (v | (v << 16)) & 0x0000FFFF0000FFFF; v = (v | (v << 8)) & 0x00FF00FF00FF00FF
            v = (v | (v << 4)) & 0x0F0F0F0F0F0F0F0F; v = (v | (v << 2)) & 0x3333333333333333
            return (v | (v << 1)) & 0x5555555555555555
        order = np.argsort(spread(q[:, 0]) | (spread(q[:, 1]) << 1), kind="stable")
        p.xy, p.y = p.xy[order], p.y[order]
    elif os.environ.get("SORT") == "rows":
        band = np.floor(p.xy[:, 1] / 0.8).astype(np.int64)
        order = np.lexsort((p.xy[:, 0], band))
        p.xy, p.y = p.xy[order], p.y[order]
    hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean); g = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    _, Linv, alpha = O.gp_fit_scipy(p.xy, p.y, hp)
    mu_ref, s2_ref = co.posterior(p.xy, hp, Linv, alpha, g)
    for kernel in kernels:
        with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                            precision="fp32", kernel=kernel, max_obs=n, max_cells=p.m) as ctx:
            ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H); ctx.set_observations(p.xy, p.y)
            res, _ = ctx.step(p.beta, p.f_min)
            mu, s2 = ctx.mu(), ctx.sigma2()
            d = s2 - s2_ref
            print(f"{tag:34s} {kernel:8s} n={n:5d} dmu max={np.abs(mu-mu_ref).max():.2e} rms={np.sqrt(np.mean((mu-mu_ref)**2)):.2e} | "
                  f"ds2 max={np.abs(d).max():.2e} rms={np.sqrt(np.mean(d*d)):.2e} mean={d.mean():+.2e} | ms={res.ms_posterior:.3f}", flush=True)
