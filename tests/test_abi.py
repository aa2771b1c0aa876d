"""CPU tests of the C-ABI surface: the library loads, exports every symbol include/sbo.h declares, the
ctypes mirror of the structs has the C layout, and argument validation that needs no GPU behaves."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "sbo.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sbo_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_what_python_binds(pkg):
    from safe_bayesian_optimization_b200 import _lib
    assert set(header_symbols()) == set(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.load()
    for name in header_symbols():
        assert hasattr(lib, name), name
    assert lib.sbo_abi_version() == 2


def test_struct_layout_matches_c(pkg):
    from safe_bayesian_optimization_b200 import _lib
    assert C.sizeof(_lib.SboConfig) == 4 * 8 + 6 * 4 + 8
    assert C.sizeof(_lib.SboArgmax) == 16
    assert C.sizeof(_lib.SboStepOut) == 8 + 8 + 16 + 16 + 8 + 4 + 4 + 4 * 4 + 8 + 4 + 4
    assert _lib.SboStepOut.argmax_width.offset == 32 and _lib.SboStepOut.ms_posterior.offset == 64
    assert _lib.SboStepOut.low_sigma2.offset == 80 and _lib.SboStepOut.drift.offset == 92


def test_null_and_bad_arguments_are_rejected_without_a_gpu(pkg):
    from safe_bayesian_optimization_b200 import _lib
    lib = pkg.load()
    h = C.c_void_p()
    assert lib.sbo_create(None, C.byref(h)) == _lib.SBO_ERR_INVALID_ARG
    bad = _lib.SboConfig(-1.0, 0.4, 0.1, 0.0, 0, 0, 0, 0, 16, 0, 16)
    assert lib.sbo_create(C.byref(bad), C.byref(h)) == _lib.SBO_ERR_INVALID_ARG
    assert b"sigma_f" in lib.sbo_last_error(None)
    assert lib.sbo_destroy(None) == _lib.SBO_ERR_INVALID_ARG
    assert lib.sbo_status_string(_lib.SBO_ERR_NOT_POSDEF) == b"covariance not positive definite"


def test_create_fails_loudly_without_a_gpu(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.SboError) as ei:
        pkg.SboContext(max_obs=8, max_cells=8)
    assert ei.value.status == -2 and "no CPU fallback" in str(ei.value)


def test_merge_step_out(pkg):
    from safe_bayesian_optimization_b200 import _lib
    lib = pkg.load()
    parts = (_lib.SboStepOut * 3)()
    for i, (ui, uv, wi, wv, sc) in enumerate([(5, 1.0, 7, 0.3, 10), (-1, 0.0, -1, 0.0, 0), (2, 1.0, 9, 0.4, 4)]):
        parts[i].n_cells = 100; parts[i].safe_count = sc
        parts[i].argmax_ucb.idx = ui; parts[i].argmax_ucb.value = uv
        parts[i].argmax_width.idx = wi; parts[i].argmax_width.value = wv
    out = _lib.SboStepOut()
    assert lib.sbo_merge_step_out(parts, 3, C.byref(out)) == 0
    assert (out.argmax_ucb.idx, out.argmax_ucb.value) == (2, 1.0)      # tie on value -> lower index
    assert (out.argmax_width.idx, out.argmax_width.value) == (9, 0.4)
    assert out.safe_count == 14 and out.n_cells == 300


def test_plan_row_bands_is_a_partition_and_balances_work(pkg):
    """sbo_plan_row_bands is host arithmetic: a partition of the rows for every part count; on a domain much wider than
    the kernel's reach (block-sparse walk) the bands carry equal estimated work, not equal rows."""
    from safe_bayesian_optimization_b200 import synthetic as S
    from safe_bayesian_optimization_b200.optimizer import plan_row_bands
    p = S.make_problem("C4", n=4096, W=2048, H=2048)
    for parts in (1, 2, 3, 8):
        cuts, share = plan_row_bands(p, parts)
        assert cuts[0] == 0 and cuts[-1] == p.H and all(b > a for a, b in zip(cuts, cuts[1:]))
        assert abs(sum(share) - 1.0) < 1e-9 and max(share) - min(share) < 0.05
    cuts8, _ = plan_row_bands(p, 8)
    rows = [b - a for a, b in zip(cuts8, cuts8[1:])]
    assert max(rows) > min(rows)                       # equal work is not equal rows here (edge bands see fewer chunks)
    cap, _ = plan_row_bands(p, 8, max_rows=p.H // 8 + 8)
    assert max(b - a for a, b in zip(cap, cap[1:])) <= p.H // 8 + 8 and cap[-1] == p.H
    small = S.make_problem("C2", n=300, W=256, H=100)   # dense walk: equal rows
    c2, _ = plan_row_bands(small, 4)
    assert c2 == [0, 25, 50, 75, 100]


def test_plan_row_bands_cuts_are_the_measured_ones(pkg):
    """The cuts the multi-GPU measurements of round 2 were taken with (profiles/bench_c4_{2,4,8}gpu_r02g.json,
    bench_default_2gpu_r02t.json: identical checksums, 98.8 % at 8 GPUs): a change of the planner that moves them is a
    change of what was measured."""
    from safe_bayesian_optimization_b200 import synthetic as S
    from safe_bayesian_optimization_b200.optimizer import plan_row_bands
    p = S.make_problem("C4")
    assert plan_row_bands(p, 2)[0] == [0, 2778, 8192]
    assert plan_row_bands(p, 8)[0] == [0, 842, 1483, 2102, 2778, 3548, 4471, 5675, 8192]
    assert plan_row_bands(S.make_problem("C3"), 8)[0] == [0, 468, 860, 1216, 1563, 1940, 2393, 2990, 4096]
    assert plan_row_bands(S.make_problem("C5"), 2)[0] == [0, 1510, 4096]
