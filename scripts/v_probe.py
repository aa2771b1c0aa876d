"""Forensics: per-entry error of V = K* L^-T as produced by the tensor core, against fp64."""
import sys, os, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
from oracle import oracle as O

tag = sys.argv[1] if len(sys.argv) > 1 else ""
for n in (256, 1024, 4096):
    p = S.make_problem("C2", n=n, W=256, H=16)
    hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean); g = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    _, Linv, alpha = O.gp_fit_scipy(p.xy, p.y, hp)
    cx, cy = g.cell_xy(np.arange(128))
    Ks = O.rbf_np(np.stack([cx, cy], 1), p.xy, hp)
    V = Ks @ Linv.T
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                        precision="fp32", kernel="tcgen05", max_obs=n, max_cells=p.m) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H); ctx.set_observations(p.xy, p.y)
        ncap = (n + 255) // 256 * 256
        buf = np.zeros((128, ncap), dtype=np.float32)
        fn = ctx.lib.sbo_debug_tc_v; fn.restype = C.c_int32; fn.argtypes = [C.c_void_p, C.c_void_p]
        assert fn(ctx.handle, buf.ctypes.data_as(C.c_void_p)) == 0
        Vg = buf[:, :n].astype(np.float64)
        D = Vg - V
        print(f"{tag:22s} n={n:5d} |V| rms={np.sqrt((V*V).mean()):.3e} max={np.abs(V).max():.3e} | dV mean={D.mean():+.3e} rms={np.sqrt((D*D).mean()):.3e} "
              f"max={np.abs(D).max():.3e} | by column third: {[f'{D[:, a:b].mean():+.2e}' for a, b in ((0, n//3), (n//3, 2*n//3), (2*n//3, n))]} "
              f"| corr(dV,V)={np.corrcoef(D.ravel(), V.ravel())[0,1]:+.3f} | sum dV^2/row={np.mean((D*D).sum(1)):.3e} 2sum(V dV)/row={np.mean(2*(V*D).sum(1)):+.3e}", flush=True)
