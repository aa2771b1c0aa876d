"""C2 posterior kernel time under the SBO_TC_DBG timing experiments (results are wrong by design; timing only)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
flags = [int(x) for x in (sys.argv[2].split(",") if len(sys.argv) > 2 else "0,1,2,4,8,9,3,5,7,15".split(","))]
p = S.make_problem("C2", n=n)
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision="fp32", max_obs=n, max_cells=p.m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    ctx.set_observations(p.xy, p.y)
    for f in flags:
        os.environ["SBO_TC_DBG"] = str(f)
        ms = []
        for _ in range(6):
            res, _ = ctx.step(p.beta, p.f_min, want_mask=False)
            ms.append(res.ms_posterior)
        print(f"n={n} dbg={f:2d} posterior min={min(ms[1:]):.3f} ms median={sorted(ms[1:])[2]:.3f} ms", flush=True)
