"""Where the steady-state step goes at C4 / C3 / C5: sbo_append_observations and sbo_step_incremental timed apart (host
wall clock around each synchronous call; the step's own events for posterior / sets), default mask copy against
SBO_OPT_MASK_DELTA.  Run under `ncu --metrics gpu__time_duration.sum` for the launch list of the same steps."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S

names = [a for a in sys.argv[1:] if "=" not in a] or ["C4"]
opts = dict(a.split("=") for a in sys.argv[1:] if "=" in a)      # e.g. thin_2d=2: sbo_set_option by name
for name in names:
    p = S.make_problem(name)
    k = max(S.CONFIGS[name].get("append", 0), 1)
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                        precision=p.precision, max_obs=p.n + 16 * k, max_cells=p.m) as ctx:
        for key, val in opts.items():
            ctx.set_option(key, int(val))
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        mask = torch.empty(p.m, dtype=torch.uint8).pin_memory().numpy()
        res, _ = ctx.step(p.beta, p.f_min, mask_out=mask)
        print(f"{name} {opts}: full step {res.ms_total:.1f} ms, fit {ctx.fit_info()[0]:.1f} ms", flush=True)
        t = 0
        for delta in (0, 1):
            ctx.set_option("mask_delta", delta)
            for rep in range(4):
                xy2, y2 = S.extra_observations(p, k, t); t += 1
                t0 = time.perf_counter()
                ctx.append_observations(xy2, y2)
                t1 = time.perf_counter()
                res, _ = ctx.step(p.beta, p.f_min, mask_out=mask, incremental=True)
                t2 = time.perf_counter()
                nb, runs = ctx.mask_copy_info()
                print(f"{name} k={k} mask_delta={delta} rep {rep}: append {1e3 * (t1 - t0):.3f} ms | step wall {1e3 * (t2 - t1):.3f} ms "
                      f"(events: posterior {res.ms_posterior:.3f} sets {res.ms_sets:.3f} total {res.ms_total:.3f}) kernel {res.kernel_used} "
                      f"mask bytes {nb} in {runs} copies; host == device mask: {np.array_equal(mask, ctx.mask())}", flush=True)
