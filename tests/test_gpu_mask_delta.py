"""GPU tests of the changed-segment mask copy (SBO_OPT_MASK_DELTA, include/sbo.h): with the option on, a step given the
same host buffer as the step before copies only the 64 KiB mask segments in which a byte changed.  The reference rebuilds
S_ and its image on every step (src/safe_bayesian_optimization_node.cpp:409, :444-456); whatever the copy leaves out, the
host buffer must equal the device mask (sbo_copy_mask) after every step, and the default path must not change."""
import numpy as np
import pytest

from conftest import make_ctx
from safe_bayesian_optimization_b200 import synthetic as S

pytestmark = pytest.mark.gpu

SEG = 1 << 16


def _problem(n=300, W=512, H=520):
    return S.make_problem("C2", n=n, W=W, H=H)    # 266 240 cells = 4 full segments + a partial one


@pytest.mark.parametrize("precision", ["fp32", "fp64"])
def test_delta_copy_keeps_the_host_mask_equal_to_the_device_mask(pkg, precision):
    p = _problem()
    with make_ctx(pkg, p, precision=precision, max_obs=p.n + 16, options={"mask_delta": 1}) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        buf = np.zeros(p.m, dtype=np.uint8)
        res, _ = ctx.step(p.beta, p.f_min, mask_out=buf)
        assert ctx.mask_copy_info() == (p.m, 1)                     # first copy into this buffer: everything
        assert 0 < res.safe_count < p.m and np.array_equal(buf, ctx.mask())
        for step in range(6):
            xy2, y2 = S.extra_observations(p, 1, step)
            ctx.append_observations(xy2, y2)
            res, _ = ctx.step(p.beta, p.f_min, mask_out=buf, incremental=True)
            nbytes, runs = ctx.mask_copy_info()
            assert 0 <= nbytes <= p.m and (nbytes == 0) == (runs == 0)
            assert nbytes % SEG in (0, p.m % SEG)                    # whole segments (the last one is partial)
            assert np.array_equal(buf, ctx.mask()), f"step {step}: host mask differs from the device mask"
            assert int(np.count_nonzero(buf)) == res.safe_count
        # the same state again: nothing changes, nothing is copied
        res, _ = ctx.step(p.beta, p.f_min, mask_out=buf)
        assert ctx.mask_copy_info() == (0, 0) and np.array_equal(buf, ctx.mask())
        # steps without a host buffer accumulate flags; the next copy brings the buffer up to date
        ctx.step_device(p.beta, p.f_min + 0.3)
        ctx.step_device(p.beta, p.f_min - 0.2)
        res, _ = ctx.step(p.beta, p.f_min - 0.2, mask_out=buf)
        assert np.array_equal(buf, ctx.mask()) and int(np.count_nonzero(buf)) == res.safe_count
        # another buffer: a full copy, and from then on that buffer is the mirror
        other = np.zeros(p.m, dtype=np.uint8)
        ctx.step(p.beta, p.f_min, mask_out=other)
        assert ctx.mask_copy_info() == (p.m, 1) and np.array_equal(other, ctx.mask())
        ctx.step(p.beta, p.f_min, mask_out=other)
        assert ctx.mask_copy_info() == (0, 0)
        # sbo_set_grid (another band of the same grid) invalidates the mirror
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H, row0=8, row1=p.H)
        sub = other[:(p.H - 8) * p.W]
        ctx.step(p.beta, p.f_min, mask_out=sub)
        assert ctx.mask_copy_info() == (sub.size, 1) and np.array_equal(sub, ctx.mask())


def test_all_safe_and_all_unsafe_masks_and_option_off(pkg):
    p = _problem(n=120, W=256, H=300)
    with make_ctx(pkg, p, precision="fp32", max_obs=p.n + 4, options={"mask_delta": 1}) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        buf = np.full(p.m, 7, dtype=np.uint8)
        res, _ = ctx.step(p.beta, -1e9, mask_out=buf)               # everything safe
        assert res.safe_count == p.m and (buf == 255).all()
        ctx.step(p.beta, -1e9, mask_out=buf)
        assert ctx.mask_copy_info() == (0, 0)
        res, _ = ctx.step(p.beta, 1e9, mask_out=buf)                # everything unsafe: every segment changes
        assert res.safe_count == 0 and (buf == 0).all() and ctx.mask_copy_info() == (p.m, 1)
        ctx.set_option("mask_delta", 0)                              # default path: the whole mask, every step
        buf[:] = 9
        ctx.step(p.beta, 1e9, mask_out=buf)
        assert (buf == 0).all() and ctx.mask_copy_info() == (p.m, 1)
        ctx.set_option("mask_delta", 1)                              # switched on again: flags were not kept, so a full copy
        buf[:] = 9
        ctx.step(p.beta, 1e9, mask_out=buf)
        assert (buf == 0).all() and ctx.mask_copy_info() == (p.m, 1)


def test_delta_copy_in_a_device_group(pkg):
    """Bands of one image: every member mirrors its own slice of the caller's buffer; the image stays the single
    context's (default copy) through a chain of appends."""
    import torch
    p = _problem(n=200, W=256, H=600)
    devices = [i % torch.cuda.device_count() for i in range(2)]
    with make_ctx(pkg, p, precision="fp32", max_obs=p.n + 8) as ctx, \
         pkg.SboGroup(devices, sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                      precision="fp32", max_obs=p.n + 8, max_cells=p.m) as g:
        g.set_option("mask_delta", 1)
        for c in (ctx, g):
            c.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
            c.set_observations(p.xy, p.y)
        buf = np.zeros(p.m, dtype=np.uint8)
        one, mask1 = ctx.step(p.beta, p.f_min)
        res, _ = g.step(p.beta, p.f_min, mask_out=buf)
        assert np.array_equal(buf, mask1) and res.safe_count == one.safe_count
        for step in range(4):
            xy2, y2 = S.extra_observations(p, 1, step)
            ctx.append_observations(xy2, y2)
            g.append_observations(xy2, y2)
            one, mask1 = ctx.step(p.beta, p.f_min, incremental=True)
            res, _ = g.step(p.beta, p.f_min, mask_out=buf, incremental=True)
            assert res.safe_count == one.safe_count and np.array_equal(buf, mask1), f"step {step}"
