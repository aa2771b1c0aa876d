"""Fit / posterior / append timings at the observation counts of C2..C4 (grid height cut so that a step stays ~1 s).
Prints one line per n; accuracy of the cut-grid posterior against the fp64 SIMT kernel (same library, fp64 mode)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S

ns = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else "1024,4096,8192,16384".split(","))]
opts = [kv.split("=") for kv in (sys.argv[2].split(",") if len(sys.argv) > 2 and sys.argv[2] else [])]   # e.g. tc_skip=0,tc_strip=0
name = {1024: "C2", 4096: "C3", 8192: "C5", 16384: "C4"}
for n in ns:
    p = S.make_problem(name.get(n, "C2"), n=n)
    W = p.W
    H = int(max(8, min(p.H, 6.0e14 / (float(n) * n * W))))   # ~1-1.5 s of credited flops
    p = S.make_problem(name.get(n, "C2"), n=n, W=W, H=H)
    k = 16
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                        precision="fp32", max_obs=n + k, max_cells=p.m) as ctx:
        for k_, v_ in opts:
            ctx.set_option(k_, int(v_))
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        t0 = time.perf_counter(); ctx.set_observations(p.xy, p.y); t_fit = time.perf_counter() - t0
        res, _ = ctx.step(p.beta, p.f_min, want_mask=False)
        res, _ = ctx.step(p.beta, p.f_min, want_mask=False)
        flops = float(p.m) * n * (n + 1) + 2.0 * p.m * n
        xy2, y2 = S.extra_observations(p, k, 0)
        t0 = time.perf_counter(); ctx.append_observations(xy2, y2); t_app = time.perf_counter() - t0
        inc, _ = ctx.step(p.beta, p.f_min, incremental=True, want_mask=False)
        print(f"n={n:6d} grid {p.W}x{p.H} fit={t_fit*1e3:9.2f} ms | posterior={res.ms_posterior:9.2f} ms "
              f"credited={flops/res.ms_posterior/1e9:7.1f} TF/s cells/s={p.m/res.ms_posterior*1e3:.3e} | "
              f"append{k}={t_app*1e3:8.2f} ms incremental step={inc.ms_posterior:8.2f} ms", flush=True)
