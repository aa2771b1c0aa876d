"""fp64 mode (C5 shape: n = 8192) on a cut grid: full step, rank-1 append, incremental step; credited TFLOP/s."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
H = int(sys.argv[2]) if len(sys.argv) > 2 else 37
p = S.make_problem("C5", n=n, W=4096, H=H)
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision="fp64", max_obs=n + 8, max_cells=p.m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    t0 = time.perf_counter(); ctx.set_observations(p.xy, p.y); t_fit = time.perf_counter() - t0
    res, _ = ctx.step(p.beta, p.f_min, want_mask=False)
    res, _ = ctx.step(p.beta, p.f_min, want_mask=False)
    flops = float(p.m) * n * (n + 1) + 2.0 * p.m * n
    print(f"fp64 n={n} m={p.m}: fit {t_fit*1e3:.1f} ms, full step {res.ms_posterior:.1f} ms = {flops/res.ms_posterior/1e9:.2f} TF/s credited, "
          f"{p.m/res.ms_posterior*1e3:.3e} cells/s", flush=True)
    for s in range(3):
        xy2, y2 = S.extra_observations(p, 1, s)
        t0 = time.perf_counter(); ctx.append_observations(xy2, y2); t_app = time.perf_counter() - t0
        inc, _ = ctx.step(p.beta, p.f_min, incremental=True, want_mask=False)
        print(f"   append 1: {t_app*1e3:.2f} ms, incremental step {inc.ms_posterior:.2f} ms ({p.m/inc.ms_posterior*1e3:.3e} cells/s)", flush=True)
