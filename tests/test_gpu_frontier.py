"""GPU parity for the stages right after the mask: the safety image of FindSafetyContourIndices
with the reference's pixel map (:431-456), the pixel -> cell tables that replace its hash map (:468-475), the contour
index list they yield with cv2.findContours, and the compacted safe / border cell lists (:944-949).  Integer / byte
work: bit-exact against the oracle."""
import numpy as np
import pytest

from conftest import make_ctx
from safe_bayesian_optimization_b200 import synthetic as S
from safe_bayesian_optimization_b200.node import contour_indices_from_image

pytestmark = pytest.mark.gpu

# (W, H, rx, ry, x0, y0): unit pitch (identity map except the dropped last row/column), fractional pitch with a
# non-integer extent (rows beyond int(max) dropped, several cells per pixel), offset origin, tall/ragged grids
SHAPES = [(40, 30, 0.1, 0.1, 0.0, 0.0), (64, 48, 0.05, 0.07, 0.0, 0.0), (33, 21, 0.31, 0.11, 0.0, 0.0),
          (301, 121, 1.0 / 300, 2.5 / 120, 0.0, 0.0), (200, 200, 0.1, 0.1, -3.3, 1.7), (1000, 517, 0.013, 0.021, 0.4, 0.2)]


def _step(pkg, W, H, rx, ry, x0, y0, n=200):
    p = S.make_problem("C2", n=n, W=W, H=H)
    p.x0, p.y0, p.rx, p.ry = x0, y0, rx, ry
    p.xy = p.xy / max(p.xy.max(), 1e-9) * np.array([W * rx, H * ry]) + np.array([x0, y0])
    ctx = make_ctx(pkg, p)
    ctx.set_grid(x0, y0, rx, ry, W, H)
    ctx.set_observations(p.xy, p.y)
    res, mask = ctx.step(p.beta, float(np.median(p.y)))
    return p, ctx, res, mask


@pytest.mark.parametrize("shape", SHAPES)
def test_frontier_image_and_contours_match_oracle(pkg, oracle, shape):
    W, H, rx, ry, x0, y0 = shape
    p, ctx, res, mask = _step(pkg, *shape)
    with ctx:
        assert 0 < res.safe_count < W * H
        g = oracle.Grid(x0, y0, rx, ry, W, H)
        img_ref, lut = oracle.safety_image(mask, g)
        img, col_of_px, row_of_py = ctx.frontier_image()
        assert np.array_equal(img, img_ref)
        # the separable tables reproduce the reference's per-cell hash map
        lut_gpu = np.where((col_of_px[:, None] >= 0) & (row_of_py[None, :] >= 0),
                           row_of_py[None, :].astype(np.int64) * W + col_of_px[:, None], -1)
        assert np.array_equal(lut_gpu, lut)
        assert np.array_equal(contour_indices_from_image(img, col_of_px, row_of_py), oracle.find_safety_contour_indices(mask, g))


@pytest.mark.parametrize("shape", SHAPES + [(4096, 1031, 0.01, 0.01, 0.0, 0.0)])
def test_compacted_cell_lists_match_oracle(pkg, oracle, shape):
    W, H, rx, ry, x0, y0 = shape
    p, ctx, res, mask = _step(pkg, *shape)
    with ctx:
        g = oracle.Grid(x0, y0, rx, ry, W, H)
        for border in (False, True):
            ref = oracle.compact_cells(mask, g, border)
            idx, count = ctx.compact_cells(border=border)
            assert count == ref.size and np.array_equal(idx, ref)
        assert ctx.compact_cells(border=False)[1] == res.safe_count
        few, count = ctx.compact_cells(border=True, cap=7)          # count-only / truncated read-back
        assert few.size == min(7, count) and np.array_equal(few, oracle.compact_cells(mask, g, True)[:few.size])


def test_shard_compaction_concatenates_to_single(pkg, oracle):
    """Row-band shards: the safe lists concatenate to the single-shard list; the border lists are supersets that only
    add cells on the shard seams (rows outside a shard are unknown to it)."""
    W, H = 256, 90
    p = S.make_problem("C2", n=150, W=W, H=H)
    g = oracle.Grid(p.x0, p.y0, p.rx, p.ry, W, H)
    parts, bparts, masks = [], [], []
    for r0, r1 in ((0, 31), (31, 64), (64, 90)):
        with make_ctx(pkg, p) as ctx:
            ctx.set_grid(p.x0, p.y0, p.rx, p.ry, W, H, r0, r1)
            ctx.set_observations(p.xy, p.y)
            _, mask = ctx.step(p.beta, p.f_min)
            masks.append(mask)
            parts.append(ctx.compact_cells()[0]); bparts.append(ctx.compact_cells(border=True)[0])
    mask = np.concatenate(masks)
    assert np.array_equal(np.concatenate(parts), oracle.compact_cells(mask, g, False))
    ref_b = set(oracle.compact_cells(mask, g, True).tolist())
    got_b = set(np.concatenate(bparts).tolist())
    extra = got_b - ref_b
    assert ref_b <= got_b and all((c // W) in (30, 31, 63, 64) for c in extra)


def test_degenerate_extent_is_an_error(pkg):
    p = S.make_problem("C2", n=20, W=16, H=16)
    with make_ctx(pkg, p) as ctx:
        ctx.set_grid(0.0, 0.0, 0.01, 0.01, 16, 16)      # 0.15 units wide: int(max) == int(min), the reference divides by 0
        ctx.set_observations(p.xy, p.y)
        ctx.step(p.beta, p.f_min)
        with pytest.raises(pkg.SboError):
            ctx.frontier_image()
