"""CPU tests of the host-side logic above the C-ABI: row sharding, result merging, the OpenCV frontier
stage (reference :418-497) against the oracle's independent restatement, and the N>1 path over gloo."""
import os
import socket

import numpy as np
import pytest

from safe_bayesian_optimization_b200 import synthetic as S
from safe_bayesian_optimization_b200.optimizer import StepResult, merge_results, shard_rows


def test_shard_rows_partition():
    for H in (1, 7, 64, 8192):
        for world in (1, 2, 3, 4, 8):
            bands = [shard_rows(H, r, world) for r in range(world)]
            assert bands[0][0] == 0 and bands[-1][1] == H
            assert all(bands[i][1] == bands[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in bands]
            assert max(sizes) - min(sizes) <= 1


def _res(ucb, wid, safe, n=10):
    return StepResult(n, safe, ucb, wid, 0, 5, 2, 1.0, 0.1, 0.0, 1.2)


def test_merge_results_order():
    m = merge_results([_res((5, 1.0), (7, 0.3), 3), _res((-1, 0.0), (-1, 0.0), 0), _res((2, 1.0), (9, 0.4), 4)])
    assert m.argmax_ucb == (2, 1.0) and m.argmax_width == (9, 0.4) and m.safe_count == 7 and m.n_cells == 30
    m = merge_results([_res((-1, 0.0), (-1, 0.0), 0)] * 2)
    assert m.argmax_ucb[0] == -1


@pytest.mark.parametrize("shape", [(40, 30, 0.1, 0.1), (64, 48, 0.05, 0.07), (33, 21, 0.31, 0.11)])
def test_frontier_stage_matches_oracle(oracle, pkg, shape):
    W, H, rx, ry = shape
    rng = np.random.default_rng(W * H)
    g = oracle.Grid(0.0, 0.0, rx, ry, W, H)
    yy, xx = np.mgrid[0:H, 0:W]
    blob = ((xx - W * 0.4) ** 2 + (yy - H * 0.5) ** 2 < (0.3 * min(W, H)) ** 2) | (rng.uniform(size=(H, W)) < 0.05)
    blob &= ~((xx - W * 0.4) ** 2 + (yy - H * 0.5) ** 2 < 9)      # a hole: its border is not frontier (RETR_EXTERNAL)
    mask = np.where(blob, 255, 0).astype(np.uint8)
    a = oracle.find_safety_contour_indices(mask.reshape(-1), g)
    b = pkg.find_safety_contour_indices(mask, 0.0, 0.0, rx, ry)
    assert np.array_equal(a, b)
    assert a.size > 0


def test_frontier_degenerate_extent_raises(pkg):
    with pytest.raises(ZeroDivisionError):
        pkg.find_safety_contour_indices(np.full((4, 4), 255, np.uint8), 0.0, 0.0, 0.1, 0.1)   # extent 0.3 < 1 unit


def test_synthetic_problems_are_seeded_and_shaped():
    a, b = S.make_problem("C2", n=64, W=32, H=16), S.make_problem("C2", n=64, W=32, H=16)
    assert np.array_equal(a.xy, b.xy) and np.array_equal(a.y, b.y) and a.f_min == b.f_min
    c1 = S.make_problem("C1")
    assert (c1.W, c1.H, c1.rx, c1.n, c1.precision) == (200, 200, 0.1, 64, "fp64")
    assert S.CONFIGS["C4"]["n"] == 16384 and S.CONFIGS["C4"]["W"] == 8192


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from oracle import oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = S.make_problem("C2", n=80, W=24, H=17)
    hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean); g = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    co = O.COracle()
    _, Linv, alpha = co.gp_fit(p.xy, p.y, hp)
    r0, r1 = shard_rows(p.H, rank, world)
    mu, s2 = co.posterior(p.xy, hp, Linv, alpha, g, r0 * p.W, r1 * p.W)     # this rank's band of rows only
    q0, q1 = co.confidence_intervals(mu, np.sqrt(np.maximum(s2, 0)), p.beta)
    mask, cnt = co.safe_set(q0, p.f_min)
    ui, uv = co.argmax_safe(q1, mask); wi, wv = co.argmax_safe(q1 - q0, mask)
    off = r0 * p.W
    part = StepResult(mask.size, cnt, (ui + off if ui >= 0 else -1, uv), (wi + off if wi >= 0 else -1, wv), 0, p.n, 0,
                      0.0, 0.0, 0.0, 0.0)
    parts = [None] * world
    dist.all_gather_object(parts, part)
    merged = merge_results(parts)
    masks = [None] * world
    dist.all_gather_object(masks, mask)
    if rank == 0:
        q.put((merged, np.concatenate(masks)))
    dist.destroy_process_group()


def test_two_rank_gloo_sharding_equals_single(oracle, coracle):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    merged, mask2 = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    p = S.make_problem("C2", n=80, W=24, H=17)
    hp = oracle.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean); g = oracle.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    _, Linv, alpha = coracle.gp_fit(p.xy, p.y, hp)
    mu, s2 = coracle.posterior(p.xy, hp, Linv, alpha, g)
    q0, q1 = coracle.confidence_intervals(mu, np.sqrt(np.maximum(s2, 0)), p.beta)
    mask, cnt = coracle.safe_set(q0, p.f_min)
    assert np.array_equal(mask, mask2) and merged.safe_count == cnt and merged.n_cells == p.m
    assert merged.argmax_ucb == coracle.argmax_safe(q1, mask)
    assert merged.argmax_width == coracle.argmax_safe(q1 - q0, mask)


def test_compact_cells_known_answers(oracle):
    """Hand-checked cases of the safe / border cell lists (the checker of tests/test_gpu_frontier.py)."""
    g = oracle.Grid(0.0, 0.0, 1.0, 1.0, 5, 4)
    full = np.full(20, 255, dtype=np.uint8)
    assert oracle.compact_cells(full, g).tolist() == list(range(20))
    # all safe: only the frame is border; the interior 3 x 2 block (cells 6,7,8,11,12,13) is not
    assert oracle.compact_cells(full, g, border=True).tolist() == [c for c in range(20) if c not in (6, 7, 8, 11, 12, 13)]
    holed = full.copy(); holed[7] = 0                      # one unsafe interior cell: all its 8 neighbours become border
    assert oracle.compact_cells(holed, g, border=True).tolist() == [c for c in range(20) if c != 7]
    assert oracle.compact_cells(np.zeros(20, dtype=np.uint8), g, border=True).size == 0


def test_safety_image_reference_quirks(oracle):
    """Unit pitch: the reference's int((c - min) / (max - min) * dim) maps cell i of W to pixel int(i * W / (W - 1)), so the
    last column and row fall outside the image and are dropped (:453), and one pixel column/row stays empty."""
    W, H = 6, 5
    g = oracle.Grid(0.0, 0.0, 1.0, 1.0, W, H)
    mask = np.full(W * H, 255, dtype=np.uint8)
    img, lut = oracle.safety_image(mask, g)
    px = [int(i / (W - 1) * W) for i in range(W)]          # 0,1,2,3,4,6 -> the last one is outside
    assert px[-1] == W and sorted(set(px[:-1])) == [0, 1, 2, 3, 4]
    assert (img[:, 5] == 0).all() and (img[4, :] == 0).all() and (img[:4, :5] == 255).all()
    assert (lut[5, :] == -1).all() and lut[2, 3] == 3 * W + 2
