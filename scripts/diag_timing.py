import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
p = S.make_problem("C2", n=128, W=256, H=16)
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision="fp32", max_obs=128, max_cells=p.m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    for rep in range(3):
        ctx.set_observations(p.xy, p.y)
        t = (C.c_longlong * 16)()
        ctx.lib.sbo_debug_diag_timing(t)
        t = list(t)
        print("rep", rep, {"load->sync": t[1]-t[0], "warp chol+inv kb0": t[3]-t[1], "sync": t[4]-t[3], "panel kb0": t[5]-t[4],
              "trailing kb0": t[6]-t[5], "kb1..3": t[7]-t[6], "full inverse": t[8]-t[7], "store": t[9]-t[8], "total": t[9]-t[0]})
