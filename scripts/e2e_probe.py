"""End-to-end step (sbo_set_observations + sbo_step, host buffers) at several n; run once per SBO_TC_EAGER setting."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
for n in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "512,1024,1536,2048,4096").split(",")]:
    p = S.make_problem("C2", n=n, W=1024, H=1024)
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                        precision="fp32", max_obs=n, max_cells=p.m) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ts = []
        for rep in range(8):
            t0 = time.perf_counter()
            ctx.set_observations(p.xy, p.y)
            res, mask = ctx.step(p.beta, p.f_min)
            ts.append((time.perf_counter() - t0) * 1e3)
        print(f"n={n:5d} eager={os.environ.get('SBO_TC_EAGER', '1')} plan={ctx.tc_plan()} e2e min {min(ts[2:]):7.3f} ms median {sorted(ts[2:])[3]:7.3f} ms  (posterior part {res.ms_posterior:.3f})", flush=True)
