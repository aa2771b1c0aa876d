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


def _shared_fit_worker(rank, world, port, q, n, nb):
    """The schedule of csrc/factor_dist.cuh restated in numpy over gloo: blocked right-looking Cholesky + blocked forward
    substitution for L^-1 with the block columns dealt in snake order; ONE broadcast per step (D_k and panel k from the
    owner of column k), one all-gather of the block columns of L^-1 at the end.  A rank only ever touches the block
    columns it owns plus what the broadcasts delivered -- everything else stays NaN, so a read of somebody else's stale
    data would poison the result."""
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    from safe_bayesian_optimization_b200.optimizer import fit_column_owner, fit_columns_of
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = S.make_problem("C2", n=n, W=8, H=8)
    hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    K = O.rbf_np(p.xy, p.xy, hp) + hp.noise_var * np.eye(n)
    nblk = n // nb
    mine = fit_columns_of(rank, world, nblk)
    blk = lambda i: slice(i * nb, (i + 1) * nb)
    A = np.full((n, n), np.nan)                       # trailing matrix: only my block columns are real
    for c in mine:
        A[:, blk(c)] = K[:, blk(c)]
    X = np.full((n, n), np.nan)                       # L^-1: only my block columns
    for c in mine:
        X[:, blk(c)] = 0.0
    for k in range(nblk):
        own = fit_column_owner(k, world)
        rem = nblk - k - 1
        msg = torch.zeros((1 + rem) * nb, nb, dtype=torch.float64)
        if own == rank:
            Lkk = np.linalg.cholesky(A[blk(k), blk(k)])
            Dk = np.linalg.inv(Lkk)
            panel = A[(k + 1) * nb:, blk(k)] @ Dk.T
            msg[:nb] = torch.from_numpy(Dk); msg[nb:] = torch.from_numpy(panel)
        dist.broadcast(msg, src=own)                  # the one exchange of step k
        Dk, panel = msg[:nb].numpy(), msg[nb:].numpy()
        for b in (c for c in mine if c > k):          # trailing update of MY columns
            rows = slice(b * nb, n)
            A[rows, blk(b)] -= panel[(b - k - 1) * nb:] @ panel[(b - k - 1) * nb:(b - k) * nb].T
        for c in (c for c in mine if c <= k):         # row block k of L^-1, my columns; then the running right-hand side
            X[blk(k), blk(c)] = Dk if c == k else Dk @ X[blk(k), blk(c)]
            if rem:
                X[(k + 1) * nb:, blk(c)] -= panel @ X[blk(k), blk(c)]
    cols = [None] * world
    dist.all_gather_object(cols, {c: X[:, blk(c)] for c in mine})   # every rank's block columns to every rank
    Linv = np.zeros((n, n))
    for part in cols:
        for c, v in part.items():
            Linv[:, blk(c)] = v
    q.put((rank, Linv))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n,nb", [(2, 96, 8), (3, 80, 8)])
def test_shared_fit_schedule_over_gloo(oracle, coracle, world, n, nb):
    """N > 1 host logic of the shared fit, on CPU: every rank ends with the oracle's L^-1."""
    import torch.multiprocessing as mp
    from safe_bayesian_optimization_b200.optimizer import fit_column_owner, fit_columns_of
    assert [fit_column_owner(c, 3) for c in range(8)] == [0, 1, 2, 2, 1, 0, 0, 1]
    assert fit_columns_of(1, 4, 20) == [1, 6, 9, 14, 17] and sorted(sum((fit_columns_of(r, 4, 20) for r in range(4)), [])) == list(range(20))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_shared_fit_worker, args=(r, world, port, q, n, nb)) for r in range(world)]
    [p.start() for p in procs]
    got = dict(q.get(timeout=90) for _ in range(world))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    p = S.make_problem("C2", n=n, W=8, H=8)
    hp = oracle.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    _, Linv, _ = coracle.gp_fit(p.xy, p.y, hp)
    for r in range(world):
        assert np.isfinite(got[r]).all() and np.abs(got[r] - Linv).max() <= 1e-10 * np.abs(Linv).max()
        assert np.array_equal(got[r], got[0])


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
