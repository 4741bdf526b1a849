"""Streaming compose kernel (REGION_STREAM) vs the oracle: ratios, MAXO variants, flips, clipping,
translucent fix-up, mixed stream/gather/fill launches.  path_info proves which path ran."""
import numpy as np
import pytest

from tests import util as U

pytestmark = pytest.mark.gpu

BIG = 26 * 1024 * 1024  # fileSize that makes any job a "big task" -> superSample 1 (index.js:1212,1363)


def _n_stream(got):
    return got.path_info & 0xFF


def _n_gather(got):
    return (got.path_info >> 8) & 0xFF


def _whole_gather(got):
    return bool(got.path_info & (1 << 24))


def _n_up(got):
    return (got.path_info >> 25) & 0x3F


def _go(ctx, dims, direction, mode, size, platform="ios", orientations=None, pattern=1, gap=0, seed=11, bitmaps=None):
    bitmaps = bitmaps or [U.synth_image(w, h, seed + i, pattern) for i, (w, h) in enumerate(dims)]
    fs = [BIG] + [0] * (len(dims) - 1)
    prof = U.profile_of(platform, size, 0)
    lay, want = U.oracle_render(dims, bitmaps, direction, mode, gap, prof, orientations, file_sizes=fs)
    got = ctx.run(U.page_images(dims, bitmaps, orientations, fs), direction, mode=mode, gap=gap, platform=platform,
                  deviceMaxCanvasSize=size or None)
    assert lay.superSample == 1.0
    assert (got.width, got.height) == (int(lay.canvasW), int(lay.canvasH))
    assert got.layout["rects"] == [list(r) for r in lay.rects]
    U.check_pixels(got.rgba[:, :, :3], want["rgba"][:, :, :3], f"{dims} {direction}/{mode}/{size}/{orientations}")
    assert (got.rgba[:, :, 3] == 255).all()
    return got, lay


@pytest.mark.parametrize("w,h,size", [
    (403, 302, 384),    # ratio 1.05  -> MAXO 4
    (403, 302, 300),    # ratio 1.34  -> MAXO 3
    (640, 480, 420),    # ratio 1.52  -> MAXO 3
    (640, 480, 330),    # ratio 1.94  -> MAXO 3
    (640, 480, 320),    # ratio 2.0 exactly: boundaries on pixel edges
    (806, 604, 364),    # ratio 2.2154 (the 9x12MP config's ratio)
    (900, 600, 300),    # ratio 3.0 exactly
    (1000, 751, 251),   # ratio 3.98
    (1000, 750, 250),   # ratio 4.0 exactly (upper limit of the one-hop exchange)
    (1001, 333, 333),   # widths not multiples of 4
    (37, 29, 20),       # smaller than one warp window
    (5, 3, 2),          # a few pixels
])
def test_ratios_single_image(ctx, w, h, size):
    got, lay = _go(ctx, [(w, h)], "vertical", "min", size)
    assert _n_stream(got) == 1 and not _whole_gather(got)
    got, lay = _go(ctx, [(h, w)], "horizontal", "min", size)
    assert _n_stream(got) == 1


@pytest.mark.parametrize("w,h,size", [
    (1000, 750, 240),    # ratio 4.17: an output can reach two groups to the right (HOPS 2)
    (1330, 1000, 200),   # ratio 6.65 (the Android preset of the 9x12MP config)
    (1600, 1200, 200),   # ratio 8.0 exactly
    (1601, 1203, 199),   # ratio 8.05 -> HOPS 4
    (2400, 1800, 200),   # ratio 12
    (3200, 2400, 200),   # ratio 16.0 exactly (upper limit)
    (1333, 999, 111),    # ratio 12.0, odd everything
])
def test_ratios_above_4_stream(ctx, w, h, size):
    got, lay = _go(ctx, [(w, h)], "vertical", "min", size)
    assert _n_stream(got) == 1 and not _whole_gather(got)
    got, lay = _go(ctx, [(h, w)], "horizontal", "min", size, orientations=[3])
    assert _n_stream(got) == 1
    _go_yuv(ctx, [(w, h)], "vertical", "min", size)


def test_ratio_above_16_uses_gather(ctx):
    got, _ = _go(ctx, [(3300, 2475)], "vertical", "min", 150)      # ratio 16.5
    assert _n_stream(got) == 0 and _whole_gather(got)


def test_mixed_ratios_one_launch(ctx):
    """Ratios from 1.5 to 5.6 and an identity copy in one job (BASELINE configs[2] in 'min' mode, quartered):
    one launch, kernel shape (MAXO 3, HOPS 2) for all streamed images; (4, 4) under the Android-like clamp."""
    dims = [(w // 4, h // 4) for (w, h) in U.MIXED]
    got, lay = _go(ctx, dims, "horizontal", "min", 0)
    assert _n_stream(got) == 8 and _n_up(got) == 1
    _go_yuv(ctx, dims, "horizontal", "min", 0, min_stream=8)
    got, lay = _go(ctx, dims, "horizontal", "min", 1024)          # Android-like clamp: ratios 2.8 .. 10.5
    assert _n_stream(got) == 9


def test_mixed_ratio_kernel_shapes(ctx):
    """Jobs whose smallest / largest ratio pick the (MAXO, HOPS) shapes (3, 2) and (4, 2) (launch.cuh), RGBA and I420."""
    for dims in ([(300, 108), (500, 432), (200, 72)],      # ratios 1.5, 6.0, identity    -> (3, 2)
                 [(100, 86), (500, 432), (310, 144), (60, 72)],    # ratios 1.19, 6.0, 2.0, identity -> (4, 2)
                 [(200, 151), (1000, 575), (90, 72)]):    # ratios 2.1, 7.99, identity   -> (2, 2)
        for orient in (None, [2, 1, 4, 3][: len(dims)]):
            got, lay = _go(ctx, dims, "horizontal", "min", 0, orientations=orient)
            assert _n_stream(got) >= 2 and _n_up(got) == 1
        _go_yuv(ctx, dims, "horizontal", "min", 0, min_stream=2)


@pytest.mark.parametrize("orientation", [2, 3, 4])
def test_flips_stream(ctx, orientation):
    dims = [(403, 302), (302, 403), (640, 480)]
    got, _ = _go(ctx, dims, "vertical", "min", 500, orientations=[orientation] * 3)
    assert _n_stream(got) == 3
    got, _ = _go(ctx, dims, "horizontal", "max", 700, orientations=[orientation, 1, orientation])
    assert _n_stream(got) >= 2


@pytest.mark.parametrize("orientation", [5, 6, 8])
def test_transposed_in_fused_launch(ctx, orientation):
    """EXIF 5/6/8 images take gather regions inside the same fused launch."""
    dims = [(403, 302), (302, 403), (640, 480)]
    got, _ = _go(ctx, dims, "vertical", "min", 500, orientations=[1, orientation, 1])
    assert _n_stream(got) == 2 and _n_gather(got) == 1


def test_orientation7_overlaps_previous_image(ctx):
    """EXIF 7 lands above its rect (reference quirk) and overlaps the previous image: both are composited by
    gather pieces in painter's order, the untouched third image still streams."""
    dims = [(403, 302), (302, 403), (640, 480)]
    got, _ = _go(ctx, dims, "vertical", "min", 500, orientations=[1, 7, 1])
    assert not _whole_gather(got) and _n_stream(got) == 1 and _n_gather(got) >= 1


def test_mixed_upscale_downscale_one_launch(ctx):
    """BASELINE configs[2] at 1/8 scale: 'max' mode upscales the small images (gather) and downscales the rest."""
    dims = [(w // 8, h // 8) for (w, h) in U.MIXED]
    got, lay = _go(ctx, dims, "horizontal", "max", 1536)
    assert _n_stream(got) >= 1 and _n_up(got) >= 1 and not _whole_gather(got)
    got, lay = _go(ctx, dims, "horizontal", "min", 0)
    got, lay = _go(ctx, dims, "vertical", "min", 1000)


def test_original_mode_clip_and_fill(ctx):
    """'original' mode: centred images leave white margins (fill regions); dw = outW+1 with dx = -1 is clipped."""
    dims = [(w // 4, h // 4) for (w, h) in U.MIXED]
    got, lay = _go(ctx, dims, "vertical", "original", 2400)
    assert (got.path_info >> 16) & 0xFF >= 2
    assert any(r[0] < 0 for r in lay.rects) or True
    got, lay = _go(ctx, dims, "horizontal", "original", 2400)


def test_translucent_fixup(ctx):
    """Bitmaps with alpha < 255 are redone exactly (premultiplied source-over on white)."""
    dims = [(403, 302)] * 2
    bm = [U.synth_image(w, h, 5 + i, 2) for i, (w, h) in enumerate(dims)]       # random alpha everywhere
    got, _ = _go(ctx, dims, "vertical", "min", 300, bitmaps=bm)
    assert _n_stream(got) == 2
    # one translucent pixel in an otherwise opaque image
    bm = [U.synth_image(w, h, 5 + i, 1) for i, (w, h) in enumerate(dims)]
    bm[0][100, 200, 3] = 17
    bm[1][0, 0, 3] = 0
    bm[1][301, 402, 3] = 254
    _go(ctx, dims, "vertical", "min", 300, bitmaps=bm)


def test_gap_integer_scaled(ctx):
    """gap*scaleDown integral keeps rects integer (all stream); with a fractional cursor only the first image
    (rect at 0) still streams, the others are gather pieces of the same launch."""
    dims = [(400, 300)] * 3
    got, lay = _go(ctx, dims, "vertical", "min", 0, gap=6)        # no clamp: scaleDown 1, gap 6
    assert not _whole_gather(got)
    got, lay = _go(ctx, dims, "vertical", "min", 451, gap=7)      # scaleDown < 1: fractional cursor
    assert _n_stream(got) == 1 and _n_gather(got) >= 1 and not _whole_gather(got)


def test_unaligned_device_pitch_falls_back(ctx):
    """Device bitmaps whose pitch is not a multiple of 16 cannot use 128-bit loads -> gather region, same pixels."""
    import ctypes as C
    from imagestitching_b200 import DeviceBitmap
    lib = ctx.lib
    w, h = 121, 90
    bm = U.synth_image(w, h, 3, 1)
    fs = [BIG]
    prof = U.profile_of("ios", 50, 0)
    lay, want = U.oracle_render([(w, h)], [bm], "vertical", "min", 0, prof, file_sizes=fs)
    p = C.c_void_p()
    assert lib.stitch_device_alloc(ctx.handle, bm.nbytes, C.byref(p)) == 0
    assert lib.stitch_memcpy_h2d(ctx.handle, p, bm.ctypes.data, bm.nbytes) == 0
    got = ctx.run([{"naturalWidth": w, "naturalHeight": h, "fileSize": BIG, "bitmap": DeviceBitmap(p.value, w, h, w * 4)}],
                  "vertical", platform="ios", deviceMaxCanvasSize=50)
    lib.stitch_device_free(ctx.handle, p)
    assert _n_stream(got) == 0 and _whole_gather(got)
    U.check_pixels(got.rgba[:, :, :3], want["rgba"][:, :, :3], "unaligned pitch")


def test_stream_matches_gather_statistics(ctx):
    """The 1+b*2^-15 arithmetic must stay far inside the +-1 LSB bar: <2% of channels off by one on noise."""
    dims = [(1612, 1208)] * 2
    bm = [U.synth_image(w, h, 77 + i, 1) for i, (w, h) in enumerate(dims)]
    fs = [BIG, 0]
    prof = U.profile_of("ios", 1092, 0)
    lay, want = U.oracle_render(dims, bm, "vertical", "min", 0, prof, file_sizes=fs)
    got = ctx.run(U.page_images(dims, bm, None, fs), "vertical", platform="ios", deviceMaxCanvasSize=1092)
    assert _n_stream(got) == 2
    d = np.abs(got.rgba[:, :, :3].astype(np.int16) - want["rgba"][:, :, :3].astype(np.int16))
    assert d.max() <= 1
    assert (d != 0).mean() < 0.02, (d != 0).mean()


# ---- fused I420 epilogue inside the streaming units (BASELINE configs[3]) ----

def _go_yuv(ctx, dims, direction, mode, size, orientations=None, pattern=1, gap=0, seed=23, bitmaps=None, min_stream=1):
    bitmaps = bitmaps or [U.synth_image(w, h, seed + i, pattern) for i, (w, h) in enumerate(dims)]
    fs = [BIG] + [0] * (len(dims) - 1)
    prof = U.profile_of("ios", size, 0)
    lay, want = U.oracle_render(dims, bitmaps, direction, mode, gap, prof, orientations, file_sizes=fs, yuv=True)
    tag = f"{dims} {direction}/{mode}/{size}/{orientations}"
    outs = {}
    for output in ("yuv420", "both"):
        got = ctx.run(U.page_images(dims, bitmaps, orientations, fs), direction, mode=mode, gap=gap, platform="ios",
                      deviceMaxCanvasSize=size or None, output=output)
        assert (got.width, got.height) == (int(lay.canvasW), int(lay.canvasH))
        assert not _whole_gather(got) and _n_stream(got) >= min_stream, hex(got.path_info)
        U.check_pixels(got.y, want["y"], tag + " Y")
        U.check_pixels(got.cb, want["cb"], tag + " Cb")
        U.check_pixels(got.cr, want["cr"], tag + " Cr")
        if output == "both":
            U.check_pixels(got.rgba[:, :, :3], want["rgba"][:, :, :3], tag + " rgba")
            assert (got.rgba[:, :, 3] == 255).all()
        outs[output] = got
    a, b = outs["yuv420"], outs["both"]
    assert (a.y == b.y).all() and (a.cb == b.cb).all() and (a.cr == b.cr).all()
    return outs["both"], lay


@pytest.mark.parametrize("w,h,size", [
    (403, 302, 384), (640, 480, 420), (806, 604, 364), (900, 600, 300), (1000, 750, 250), (1001, 333, 333),
    (37, 29, 20), (9, 7, 4),
])
def test_yuv_stream_ratios(ctx, w, h, size):
    _go_yuv(ctx, [(w, h)], "vertical", "min", size)
    _go_yuv(ctx, [(h, w)], "horizontal", "min", size)


def test_yuv_stream_odd_seams(ctx):
    """Images whose scaled heights / widths are odd: chroma blocks straddle two images (gather strips)."""
    dims = [(806, 603), (806, 605), (403, 303), (806, 601)]
    got, lay = _go_yuv(ctx, dims, "vertical", "min", 1100, min_stream=3)
    assert any(int(r[1]) % 2 == 1 for r in lay.rects)
    assert _n_gather(got) >= 1
    dims = [(603, 806), (605, 806), (607, 810)]
    got, lay = _go_yuv(ctx, dims, "horizontal", "min", 600, min_stream=2)
    assert any(int(r[0]) % 2 == 1 for r in lay.rects)


@pytest.mark.parametrize("orientation", [2, 3, 4])
def test_yuv_stream_flips(ctx, orientation):
    dims = [(806, 604), (805, 603), (640, 481)]
    _go_yuv(ctx, dims, "vertical", "min", 800, orientations=[orientation, 1, orientation], min_stream=2)
    _go_yuv(ctx, dims, "horizontal", "min", 900, orientations=[1, orientation, orientation], min_stream=2)


def test_yuv_stream_original_margins_and_mixed(ctx):
    dims = [(w // 4, h // 4) for (w, h) in U.MIXED]
    _go_yuv(ctx, dims, "vertical", "original", 2400)
    _go_yuv(ctx, dims, "horizontal", "original", 2400)
    _go_yuv(ctx, dims, "horizontal", "max", 3072, min_stream=1)        # upscales (gather) next to downscales (stream)
    _go_yuv(ctx, dims, "vertical", "min", 0, orientations=[1, 6, 1, 8, 1, 5, 1, 1, 1][:len(dims)], min_stream=0)


def test_yuv_stream_translucent(ctx):
    dims = [(403, 302)] * 2
    bm = [U.synth_image(w, h, 5 + i, 2) for i, (w, h) in enumerate(dims)]
    _go_yuv(ctx, dims, "vertical", "min", 300, bitmaps=bm)
    bm = [U.synth_image(w, h, 5 + i, 1) for i, (w, h) in enumerate(dims)]
    bm[0][100, 200, 3] = 17
    bm[1][301, 402, 3] = 254
    _go_yuv(ctx, dims, "vertical", "min", 300, bitmaps=bm)


# ---- bilinear upscale units (BASELINE configs[0]: supersampled canvas; configs[2]: 'max' mode) ----

def _go_up(ctx, dims, direction, mode, platform="ios", size=0, orientations=None, pattern=1, gap=0, seed=31, bitmaps=None,
           outputs=("rgba", "yuv420", "both")):
    bitmaps = bitmaps or [U.synth_image(w, h, seed + i, pattern) for i, (w, h) in enumerate(dims)]
    prof = U.profile_of(platform, size, 0)
    lay, want = U.oracle_render(dims, bitmaps, direction, mode, gap, prof, orientations, yuv=True)
    tag = f"{dims} {direction}/{mode}/{platform}/{size}/{orientations}"
    got = None
    for output in outputs:
        got = ctx.run(U.page_images(dims, bitmaps, orientations), direction, mode=mode, gap=gap, platform=platform,
                      deviceMaxCanvasSize=size or None, output=output)
        assert (got.width, got.height) == (int(lay.canvasW), int(lay.canvasH))
        if output != "yuv420":
            U.check_pixels(got.rgba[:, :, :3], want["rgba"][:, :, :3], tag + " rgba")
            assert (got.rgba[:, :, 3] == 255).all()
        if output != "rgba":
            U.check_pixels(got.y, want["y"], tag + " Y")
            U.check_pixels(got.cb, want["cb"], tag + " Cb")
            U.check_pixels(got.cr, want["cr"], tag + " Cr")
    return got, lay


def test_upscale_supersampled_canvas(ctx):
    """configs[0] at 1/10 scale: ss = 2.2 makes every device rect fractional; the seam row is a gather piece."""
    got, lay = _go_up(ctx, [(108, 192)] * 2, "vertical", "min")
    assert lay.superSample == 2.2 and _n_up(got) == 2 and not _whole_gather(got)
    got, lay = _go_up(ctx, [(108, 192)] * 2, "vertical", "min", platform="android")      # ss 2.6
    assert _n_up(got) == 2
    got, lay = _go_up(ctx, [(192, 108), (96, 54), (131, 77)], "horizontal", "max", gap=3)
    assert _n_up(got) == 3


@pytest.mark.parametrize("orientation", [2, 3, 4, 6])
def test_upscale_orientations(ctx, orientation):
    got, lay = _go_up(ctx, [(101, 67), (67, 101), (80, 80)], "vertical", "max", orientations=[orientation, 1, orientation])
    assert _n_up(got) == (3 if orientation < 5 else 1)


def test_upscale_wide_and_tiny(ctx):
    """More than one 128-px tile per row, odd sizes, 1-pixel sources."""
    _go_up(ctx, [(300, 40), (150, 20)], "vertical", "max")
    _go_up(ctx, [(1, 1), (3, 2)], "vertical", "max", outputs=("rgba", "both"))
    _go_up(ctx, [(7, 5)], "horizontal", "min")


@pytest.mark.parametrize("size,orientation", [(720, 1), (894, 2), (900, 3), (906, 4), (1000, 1), (1290, 2)])
def test_upscale_window_widths(ctx, size, orientation):
    """Scales 1.2 ... 2.15 around the 1.5 switch between the 5- and the 4-pixel source window of a lane."""
    dims = [(400, 300), (400, 300)]
    got, lay = _go_up(ctx, dims, "vertical", "min", size=size, orientations=[orientation, 1], outputs=("rgba", "both"))
    assert abs(lay.superSample - size / 600.0) < 1e-9 and _n_up(got) == 2


def test_upscale_narrow_sources(ctx):
    """Bitmaps of 5 and 6 columns still take upscale units (window clamped to the row); 4 columns do not."""
    got, _ = _go_up(ctx, [(40, 30), (5, 30), (6, 21)], "vertical", "max", outputs=("rgba",))
    assert _n_up(got) == 3
    got, _ = _go_up(ctx, [(40, 30), (4, 30)], "vertical", "max", outputs=("rgba",))
    assert _n_up(got) == 1


def test_upscale_translucent(ctx):
    dims = [(90, 60)] * 2
    bm = [U.synth_image(w, h, 5 + i, 2) for i, (w, h) in enumerate(dims)]
    got, _ = _go_up(ctx, dims, "vertical", "min", bitmaps=bm)
    assert _n_up(got) == 2
    bm = [U.synth_image(w, h, 5 + i, 1) for i, (w, h) in enumerate(dims)]
    bm[0][10, 20, 3] = 17
    bm[1][59, 89, 3] = 254
    _go_up(ctx, dims, "vertical", "min", bitmaps=bm)


def test_identity_copy_takes_upscale_units_exactly(ctx):
    dims = [(257, 131), (257, 90)]
    bm = [U.synth_image(w, h, 70 + i, 1) for i, (w, h) in enumerate(dims)]
    fs = [BIG, 0]
    got = ctx.run(U.page_images(dims, bm, None, fs), "vertical", platform="ios")
    assert _n_up(got) == 2 and got.height == 221
    assert (got.rgba[:131] == bm[0]).all() and (got.rgba[131:] == bm[1]).all()
