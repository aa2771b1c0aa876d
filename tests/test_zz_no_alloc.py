"""sbo_compute_sets_noalloc (include/sbo.h: no counterpart in the reference):
the CI + mask + argmax pass with its reduction scratch in a fixed __device__ array -- no context, no allocation.
Checked through the C-ABI against the existing entry point sbo_compute_sets_device on the same device-resident inputs:
bit-identical mask bytes, safe count, argmax cells and values (the tolerance of tests/test_gpu_sets.py for this pass is
bit-exactness), on two calls in a row of different sizes (the shared scratch keeps the first call's contents), plus the
capacity error.  Collected last on purpose (-x at the round's end)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(m, seed):
    rng = np.random.default_rng(seed)
    mu = 5.0 + rng.normal(0.0, 1.5, m)
    std = np.abs(rng.normal(0.3, 0.2, m))
    if m > 40:
        mu[7] = np.nan                      # NaN -> unsafe (reference :409)
        mu[11:15] = mu[11]; std[11:15] = std[11]     # ties -> lowest index
    return mu, std


def test_noalloc_entry_point_matches_compute_sets_device(pkg):
    import torch
    from safe_bayesian_optimization_b200 import _lib
    lib = pkg.load()
    torch.cuda.set_device(0)
    beta, f_min = 2.0, 4.2
    stream = torch.cuda.current_stream()
    with pkg.SboContext(precision="fp64", max_obs=8, max_cells=1 << 21) as ctx:
        for call, (m, seed) in enumerate([(1_000_003, 1), (4_099, 2), (1_000_003, 3)]):
            mu, std = _inputs(m, seed)
            d_mu, d_std = torch.from_numpy(mu).cuda(), torch.from_numpy(std).cuda()
            d_mask_a = torch.zeros(m, dtype=torch.uint8, device="cuda")
            d_mask_b = torch.full((m,), 7, dtype=torch.uint8, device="cuda")
            d_rec = torch.zeros(7, dtype=torch.int64, device="cuda")
            torch.cuda.synchronize()
            # existing entry point (context scratch)
            out = _lib.SboStepOut()
            st = lib.sbo_compute_sets_device(ctx.handle, C.c_void_p(d_mu.data_ptr()), C.c_void_p(d_std.data_ptr()), m, beta, f_min,
                                             C.c_void_p(d_mask_a.data_ptr()), C.byref(out))
            assert st == 0, lib.sbo_last_error(ctx.handle)
            # the allocation-free variant: free device memory must not move across the call
            free0 = torch.cuda.mem_get_info()[0]
            st = lib.sbo_compute_sets_noalloc(C.c_void_p(d_mu.data_ptr()), C.c_void_p(d_std.data_ptr()), m, beta, f_min,
                                              C.c_void_p(d_mask_b.data_ptr()), C.c_void_p(d_rec.data_ptr()),
                                              C.c_void_p(stream.cuda_stream))
            assert st == 0, lib.sbo_last_error(None)
            torch.cuda.synchronize()
            if call > 0:                       # (the first call on a device may load the module that holds the array)
                assert torch.cuda.mem_get_info()[0] == free0
            rec = _lib.SboSetsRecord.from_buffer_copy(d_rec.cpu().numpy().tobytes())
            assert torch.equal(d_mask_a, d_mask_b)
            assert rec.safe_count == out.safe_count == int((d_mask_b == 255).sum().item()) and 0 < rec.safe_count < m
            assert (rec.ucb_idx, rec.ucb_value) == (out.argmax_ucb.idx, out.argmax_ucb.value)
            assert (rec.width_idx, rec.width_value) == (out.argmax_width.idx, out.argmax_width.value)
            assert rec.clamped == 0 and rec.low_sigma2 == 0
            # and against numpy directly (fp64, same three operations)
            conf = beta * std
            q0, q1 = mu - conf, mu + conf
            safe = q0 > f_min
            assert rec.safe_count == int(safe.sum())
            assert rec.ucb_idx == int(np.flatnonzero(safe & (q1 == q1[safe].max()))[0])
        # empty input: a record with no safe cell
        d_rec = torch.ones(7, dtype=torch.int64, device="cuda")
        st = lib.sbo_compute_sets_noalloc(None, None, 0, beta, f_min, None, C.c_void_p(d_rec.data_ptr()), C.c_void_p(stream.cuda_stream))
        assert st == 0
        torch.cuda.synchronize()
        rec = _lib.SboSetsRecord.from_buffer_copy(d_rec.cpu().numpy().tobytes())
        assert rec.ucb_idx == -1 and rec.width_idx == -1 and rec.safe_count == 0
        # beyond the fixed capacity: rejected with a clear error, nothing launched
        st = lib.sbo_compute_sets_noalloc(C.c_void_p(d_mu.data_ptr()), C.c_void_p(d_std.data_ptr()), (1 << 27) + 1, beta, f_min,
                                          C.c_void_p(d_mask_b.data_ptr()), C.c_void_p(d_rec.data_ptr()), C.c_void_p(stream.cuda_stream))
        assert st == _lib.SBO_ERR_CAPACITY and b"exceeds the fixed capacity of 134217728 cells" in lib.sbo_last_error(None)
        st = lib.sbo_compute_sets_noalloc(C.c_void_p(d_mu.data_ptr()), C.c_void_p(d_std.data_ptr()), 16, beta, f_min,
                                          C.c_void_p(d_mask_b.data_ptr()), None, None)
        assert st == _lib.SBO_ERR_INVALID_ARG
