"""BASELINE.json configs at full size through the C ABI vs the C (fp64) oracle,
plus size-independent properties (idempotence, mirror symmetry, RGBA == RGBA-of-'both')."""
import numpy as np
import pytest

from tests import util as U

pytestmark = pytest.mark.gpu


def _imgs(dims, seed0=0x5717C4, pattern=0):
    return [U.synth_image(w, h, seed0 + i, pattern) for i, (w, h) in enumerate(dims)]


def _check(ctx, dims, bitmaps, direction, mode, platform, output="rgba", gap=0):
    prof = U.profile_of(platform)
    lay, want = U.oracle_render(dims, bitmaps, direction, mode, gap, prof, yuv=(output != "rgba"))
    got = ctx.run(U.page_images(dims, bitmaps), direction, mode=mode, gap=gap, platform=platform, output=output)
    assert (got.width, got.height) == (int(lay.canvasW), int(lay.canvasH))
    assert got.layout["rects"] == [list(r) for r in lay.rects]
    assert got.progress == lay.progress
    if output in ("rgba", "both"):
        U.check_pixels(got.rgba[:, :, :3], want["rgba"][:, :, :3], f"{direction}/{mode}/{platform}")
    if output in ("yuv420", "both"):
        for k in ("y", "cb", "cr"):
            U.check_pixels(getattr(got, k), want[k], f"{direction}/{mode}/{platform} {k}")
    return got, lay


def test_cfg1_two_1080x1920_vertical(ctx):
    """configs[0]: 2x 1080x1920 vertical to common width 1080; iOS supersamples 2.2x -> 2376x8448."""
    dims = [(1080, 1920)] * 2
    bm = _imgs(dims)
    got, lay = _check(ctx, dims, bm, "vertical", "min", "ios")
    assert (got.width, got.height) == (2376, 8448) and lay.superSample == 2.2
    got, lay = _check(ctx, dims, bm, "vertical", "min", "android")
    assert (got.width, got.height) == (1152, 4096)


@pytest.mark.parametrize("platform,canvas", [("ios", (1820, 12288)), ("android", (606, 4096))])
def test_cfg2_nine_12mp_vertical(ctx, platform, canvas):
    """configs[1]: 9x 4032x3024 vertical with the canvas-size/memory clamp (heavy downscale)."""
    dims = [(4032, 3024)] * 9
    bm = _imgs(dims)
    got, lay = _check(ctx, dims, bm, "vertical", "min", platform)
    assert (got.width, got.height) == canvas
    # white tail strip below the last image (sum of rounded heights < clamped canvas height)
    end = int(lay.rects[-1][1] + lay.rects[-1][3])
    assert end < got.height and (got.rgba[end:, :, :3] == 255).all()


@pytest.mark.parametrize("mode", ["max", "min"])
def test_cfg3_mixed_horizontal(ctx, mode):
    """configs[2]: 9 mixed-aspect images horizontal to common height (upscale + downscale mix in 'max')."""
    bm = _imgs(U.MIXED)
    _check(ctx, U.MIXED, bm, "horizontal", mode, "ios")


def test_cfg4_yuv420_epilogue(ctx):
    """configs[3]: 9x 12MP vertical with the fused RGBA->YUV420 export epilogue."""
    dims = [(4032, 3024)] * 9
    bm = _imgs(dims)
    got, _ = _check(ctx, dims, bm, "vertical", "min", "ios", output="both")
    only = ctx.run(U.page_images(dims, bm), "vertical", platform="ios", output="yuv420")
    assert (only.y == got.y).all() and (only.cb == got.cb).all() and (only.cr == got.cr).all()
    rgba = ctx.run(U.page_images(dims, bm), "vertical", platform="ios", output="rgba")
    # all three outputs run the streaming units (31- vs 30-group bands): same arithmetic per pixel
    assert (got.path_info & 0xFF) >= 9 and (only.path_info & 0xFF) >= 9 and not (only.path_info & (1 << 24))
    d = np.abs(rgba.rgba.astype(np.int16) - got.rgba.astype(np.int16))
    assert d.max() <= 1 and (d != 0).mean() < 0.02


def test_mixed_original_gap(ctx):
    """Fractional cursors (gap*scaleDown), dx=-1 / dw=outW+1 overflow (SURVEY App. B last row)."""
    bm = _imgs(U.MIXED)
    got, lay = _check(ctx, U.MIXED, bm, "vertical", "original", "ios", gap=7)
    assert lay.rects[7][0] == -1.0 and lay.rects[7][2] == 2401.0


def test_properties_full_size(ctx):
    """Idempotence and mirror symmetry at full size (no oracle involved)."""
    dims = [(4032, 3024)] * 9
    bm = _imgs(dims, pattern=1)
    a = ctx.run(U.page_images(dims, bm), "vertical", platform="ios")
    b = ctx.run(U.page_images(dims, bm), "vertical", platform="ios")
    assert (a.rgba == b.rgba).all()
    # horizontally mirrored bitmaps drawn with EXIF 2 (mirror) reproduce the same canvas
    mirrored = [np.ascontiguousarray(x[:, ::-1]) for x in bm]
    c = ctx.run(U.page_images(dims, mirrored, orientations=[2] * 9), "vertical", platform="ios")
    d = np.abs(a.rgba.astype(np.int16) - c.rgba.astype(np.int16))
    assert d.max() <= 1 and (d != 0).mean() < 1e-3
