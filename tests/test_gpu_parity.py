"""GPU parity: CUDA path through the C ABI vs the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): canvas size and rects bit-exact; pixels within
+-1 LSB per channel and PSNR >= 50 dB of the stated AB1 filter (fp64 oracle).
"""
import itertools

import os

import numpy as np
import pytest

from imagestitching_b200 import StitchError, StitchPage
from imagestitching_b200 import _lib as L
from oracle import layout_ref as LR
from tests import util as U

pytestmark = pytest.mark.gpu

SMALL = [(40, 30), (31, 47), (64, 64)]
BIG_FILE = 26 * 1024 * 1024   # makes a job a 'big task' (P:1212): superSample stays 1


def _noise(dims, seed, alpha=False):
    rng = np.random.default_rng(seed)
    out = []
    for (w, h) in dims:
        a = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        if not alpha:
            a[:, :, 3] = 255
        out.append(a)
    return out


def _run_and_check(ctx, dims, bitmaps, direction, mode, gap, platform="ios", size=0, pixels=0, orientations=None,
                   output="rgba", use_numpy=False, file_sizes=None):
    prof = U.profile_of(platform, size, pixels)
    lay, want = U.oracle_render(dims, bitmaps, direction, mode, gap, prof, orientations, yuv=(output != "rgba"),
                                use_numpy=use_numpy, file_sizes=file_sizes)
    got = ctx.run(U.page_images(dims, bitmaps, orientations, file_sizes), direction, mode=mode, gap=gap,
                  platform=platform, deviceMaxCanvasSize=size or None, deviceMaxCanvasPixels=pixels or None,
                  output=output)
    # layout bit-exact
    assert (got.width, got.height) == (int(lay.canvasW), int(lay.canvasH))
    assert got.layout["rects"] == [list(r) for r in lay.rects]
    assert got.layout["superSample"] == lay.superSample and got.layout["scaleDown"] == lay.scaleDown
    assert got.progress == lay.progress
    what = f"{direction}/{mode}/gap{gap}/{platform}/{size}/{orientations}"
    if output in ("rgba", "both"):
        U.check_pixels(got.rgba[:, :, :3], want["rgba"][:, :, :3], what)
        assert (got.rgba[:, :, 3] == 255).all()
    if output in ("yuv420", "both"):
        U.check_pixels(got.y, want["y"], what + " Y")
        U.check_pixels(got.cb, want["cb"], what + " Cb")
        U.check_pixels(got.cr, want["cr"], what + " Cr")
    return got, lay


@pytest.mark.parametrize("direction", ["vertical", "horizontal"])
@pytest.mark.parametrize("mode", ["min", "max", "original"])
@pytest.mark.parametrize("gap", [0, 7])
def test_small_all_modes(ctx, direction, mode, gap):
    """Every direction x mode x gap, three limit profiles (supersample 2.2, heavy clamp, tiny side)."""
    bitmaps = _noise(SMALL, 1)
    for (platform, size, pixels) in (("ios", 0, 0), ("ios", 100, 3000), ("other", 64, 0), ("android", 0, 0)):
        _run_and_check(ctx, SMALL, bitmaps, direction, mode, gap, platform, size, pixels, use_numpy=True)


@pytest.mark.parametrize("orientation", [1, 2, 3, 4, 5, 6, 7, 8, 0, 9])
def test_orientations(ctx, orientation):
    """EXIF 1-8 (utils/canvas.js:153-202) incl. the orientation-7 quirk; 0/9 draw plain."""
    bitmaps = _noise(SMALL, 2)
    for direction in ("vertical", "horizontal"):
        for (platform, size) in (("ios", 0), ("ios", 100)):
            _run_and_check(ctx, SMALL, bitmaps, direction, "min", 3, platform, size, 0,
                           orientations=[orientation, 1, orientation], use_numpy=True)


def test_alpha_source_over(ctx):
    """Straight-alpha bitmaps composite source-over on white (index.js:1423-1424)."""
    bitmaps = _noise(SMALL, 3, alpha=True)
    for direction, mode in itertools.product(("vertical", "horizontal"), ("min", "original")):
        _run_and_check(ctx, SMALL, bitmaps, direction, mode, 5, "ios", 90, 0, use_numpy=True)
        _run_and_check(ctx, SMALL, bitmaps, direction, mode, 0, "ios", 0, 0, orientations=[6, 3, 8], use_numpy=True)


def test_constant_colour_exact(ctx):
    """A constant image must come out exactly (weights sum to 1)."""
    dims = [(100, 80), (50, 120)]
    bitmaps = []
    for (w, h), col in zip(dims, ((10, 200, 77, 255), (255, 0, 128, 255))):
        a = np.empty((h, w, 4), np.uint8)
        a[:] = col
        bitmaps.append(a)
    got = ctx.run(U.page_images(dims, bitmaps), "vertical", mode="min", gap=0, platform="ios",
                  deviceMaxCanvasSize=60)
    lay = got.layout
    y0 = int(lay["rects"][1][1])
    assert (got.rgba[: y0, :, :3] == np.array([10, 200, 77])).all()
    h1 = int(lay["rects"][1][3])
    assert (got.rgba[y0: y0 + h1, :, :3] == np.array([255, 0, 128])).all()


def test_identity_copy_exact(ctx):
    """scale 1, integer offsets: drawImage is a plain copy."""
    dims = [(64, 48), (64, 32)]
    bitmaps = _noise(dims, 4)
    # 9 images would make it a big task (ss=1); with 2 images force ss=1 through a pixel cap == base pixels
    got = ctx.run(U.page_images(dims, bitmaps), "vertical", mode="min", gap=0, platform="ios",
                  deviceMaxCanvasSize=80, deviceMaxCanvasPixels=64 * 80)
    assert (got.width, got.height) == (64, 80)
    assert (got.rgba[:48] == bitmaps[0]).all() and (got.rgba[48:] == bitmaps[1]).all()


@pytest.mark.parametrize("pattern", [0, 1, 3, 4])
def test_patterns_downscale(ctx, pattern):
    """Natural-ish, noise, 1-px checkerboard and ramp content through a 2.2x / 6.6x downscale."""
    dims = [(403, 302)] * 3
    bitmaps = [U.synth_image(w, h, 100 + i, pattern) for i, (w, h) in enumerate(dims)]
    _run_and_check(ctx, dims, bitmaps, "vertical", "min", 0, "ios", 400, 0)
    _run_and_check(ctx, dims, bitmaps, "vertical", "min", 0, "android", 130, 0)
    _run_and_check(ctx, dims, bitmaps, "horizontal", "min", 0, "ios", 500, 0)


@pytest.mark.parametrize("output", ["yuv420", "both"])
def test_yuv420_epilogue(ctx, output):
    """Fused RGBA->I420 (JFIF BT.601) incl. odd canvas sizes and seams at odd rows."""
    bitmaps = _noise(SMALL, 5)
    for (direction, size) in (("vertical", 0), ("vertical", 101), ("horizontal", 77)):
        _run_and_check(ctx, SMALL, bitmaps, direction, "min", 0, "ios", size, 0, output=output, use_numpy=True)


def test_ragged_and_extreme(ctx):
    """1x1 images, 1-pixel-wide strips, single image, nine images."""
    cases = [
        [(1, 1)], [(1, 1), (1, 1)], [(1, 300), (200, 1)], [(300, 1), (2, 2), (1, 7)],
        [(17, 13)] * 9, [(5, 400), (400, 5)],
    ]
    for dims in cases:
        bitmaps = _noise(dims, 6)
        for direction in ("vertical", "horizontal"):
            for mode in ("min", "max", "original"):
                _run_and_check(ctx, dims, bitmaps, direction, mode, 2, "ios", 0, 0, use_numpy=True)


def test_big_task_by_file_size(ctx):
    """>= 25 MiB total fileSize disables supersampling (index.js:1212,1363)."""
    dims = [(40, 30), (40, 30)]
    bitmaps = _noise(dims, 7)
    got, lay = _run_and_check(ctx, dims, bitmaps, "vertical", "min", 0, file_sizes=[13 * 1024 * 1024] * 2, use_numpy=True)
    assert lay.superSample == 1.0 and lay.bigTask
    got, lay = _run_and_check(ctx, dims, bitmaps, "vertical", "min", 0, file_sizes=[12 * 1024 * 1024] * 2, use_numpy=True)
    assert lay.superSample == 2.2 and not lay.bigTask


def test_errors(ctx):
    """Error behaviour mirrors the reference's throw sites."""
    assert ctx.run([], "vertical") is None                               # index.js:1189
    bm = _noise([(8, 8)], 8)[0]
    with pytest.raises(StitchError) as e:
        ctx.run([{"naturalWidth": 8, "naturalHeight": 8, "bitmap": None}], "vertical")
    assert e.value.status == L.E_BAD_IMAGE and "图片0解码异常" in e.value.message   # index.js:1513
    with pytest.raises(StitchError) as e:
        ctx.run([{"naturalWidth": 8, "naturalHeight": 8, "bitmap": bm}] * 10, "vertical")
    assert e.value.status == L.E_TOO_MANY
    # missing natural size falls back to width/height, then to 1 (index.js:1243-1244)
    got = ctx.run([{"width": 8, "height": 8, "bitmap": bm}], "vertical", platform="ios", deviceMaxCanvasSize=8,
                  deviceMaxCanvasPixels=64)
    assert (got.width, got.height) == (8, 8) and (got.rgba == bm).all()


def test_page_object(ctx):
    """StitchPage keeps the page's field names and progress/flag behaviour (index.js:14-21,1611,1618-1624)."""
    page = StitchPage(0, "ios")
    page.onStitchVertical()
    assert page.lastToast == "请先添加图片" and not page.data["isStitching"]
    dims = SMALL
    bitmaps = _noise(dims, 9)
    page.setData({"images": U.page_images(dims, bitmaps), "gap": 4, "horizontalStitchMode": "max"})
    page.onStitchHorizontal()
    assert page.data["direction"] == "horizontal" and page.data["isStitching"] is False
    assert page.data["stitchProgress"] == 100 and page.data["stitchedTempPath"]
    lay, want = U.oracle_render(dims, bitmaps, "horizontal", "max", 4, U.profile_of("ios"), use_numpy=True)
    assert page.progressLog == lay.progress
    U.check_pixels(page.result.rgba[:, :, :3], want["rgba"][:, :, :3], "page")
    # failure path: toast + progress reset
    page.setData({"images": [{"naturalWidth": 5, "naturalHeight": 5, "bitmap": None}]})
    page.onStitchVertical()
    assert page.lastToast.startswith("拼图失败：") and page.data["stitchProgress"] == 0
    assert page.data["isStitching"] is False


def test_device_buffers_and_async(ctx):
    """Device-resident inputs/outputs + submit/wait give the same bytes as the host path."""
    import ctypes as C
    lib = ctx.lib
    dims = [(120, 90)] * 3
    bitmaps = [U.synth_image(w, h, 40 + i, 0) for i, (w, h) in enumerate(dims)]
    ref = ctx.run(U.page_images(dims, bitmaps), "vertical", platform="ios", deviceMaxCanvasSize=200)
    from imagestitching_b200 import DeviceBitmap
    from imagestitching_b200.stitch import _profile_struct
    dptrs = []
    imgs = []
    for (w, h), b in zip(dims, bitmaps):
        p = C.c_void_p()
        assert lib.stitch_device_alloc(ctx.handle, b.nbytes, C.byref(p)) == 0
        assert lib.stitch_memcpy_h2d(ctx.handle, p, b.ctypes.data, b.nbytes) == 0
        dptrs.append(p)
        imgs.append({"naturalWidth": w, "naturalHeight": h, "bitmap": DeviceBitmap(p.value, w, h, w * 4)})
    out = C.c_void_p()
    nbytes = ref.rgba.nbytes
    assert lib.stitch_device_alloc(ctx.handle, nbytes, C.byref(out)) == 0
    job, keep = ctx.build_job(imgs, "vertical", "min", 0, _profile_struct("ios", 200, None), "rgba", True, out.value, 0,
                              nbytes, 0)
    res = L.StitchResult()
    tickets = []
    for _ in range(3):
        t = C.c_uint64()
        assert lib.stitch_submit(ctx.handle, C.byref(job), C.byref(res), C.byref(t)) == 0
        tickets.append(t.value)
    assert lib.stitch_wait(ctx.handle, tickets[-1]) == 0
    host = np.empty_like(ref.rgba)
    assert lib.stitch_memcpy_d2h(ctx.handle, host.ctypes.data, out, nbytes) == 0
    assert (host == ref.rgba).all()
    for p in dptrs + [out]:
        lib.stitch_device_free(ctx.handle, p)


# STITCH_FUZZ_SEEDS / STITCH_FUZZ_BASE widen the sweep for a one-off fuzz run (e.g. 64 seeds from 100)
_FUZZ_BASE = int(os.environ.get("STITCH_FUZZ_BASE", "0"))


@pytest.mark.parametrize("seed", range(_FUZZ_BASE, _FUZZ_BASE + int(os.environ.get("STITCH_FUZZ_SEEDS", "8"))))
def test_randomized_jobs(ctx, seed):
    """Random jobs (sizes, modes, gaps, clamps, EXIF, alpha, output formats): whatever mixture of stream / upscale /
    gather / fill units the plan chooses, layout is bit-exact and every plane is within the parity bar."""
    rng = np.random.default_rng(1000 + seed)
    for case in range(10):
        n = int(rng.integers(1, 10))
        dims = [(int(rng.integers(3, 420)), int(rng.integers(3, 420))) for _ in range(n)]
        direction = ["vertical", "horizontal"][int(rng.integers(0, 2))]
        mode = ["min", "max", "original"][int(rng.integers(0, 3))]
        gap = [0, 0, 3, 7.5, 20][int(rng.integers(0, 5))]
        platform = ["ios", "android", "other"][int(rng.integers(0, 3))]
        size = [0, 0, int(rng.integers(60, 900))][int(rng.integers(0, 3))]
        orientations = [int(rng.integers(1, 9)) if rng.random() < 0.25 else 1 for _ in range(n)]
        fs = [BIG_FILE if (i == 0 and rng.random() < 0.6) else 0 for i in range(n)]
        output = ["rgba", "yuv420", "both"][int(rng.integers(0, 3))]
        bitmaps = []
        for i, (w, h) in enumerate(dims):
            pat = 2 if rng.random() < 0.1 else int(rng.integers(0, 2))      # 10 % translucent noise
            bitmaps.append(U.synth_image(w, h, 5000 + 100 * seed + 10 * case + i, pat))
        _run_and_check(ctx, dims, bitmaps, direction, mode, gap, platform, size, 0, orientations, output=output,
                       file_sizes=fs)


def test_job_lanes_batch(ctx):
    """A batch of different host-buffer jobs rotated over 3 lanes (stitch_ctx_set_lanes): every job's export
    equals the one-at-a-time result; tickets retire in order; lanes cannot change while jobs are in flight."""
    import ctypes as C
    from imagestitching_b200.stitch import _profile_struct
    lib = ctx.lib
    cases = []
    for k in range(7):
        dims = [(160 + 8 * k, 120 + k)] * (2 + k % 3)
        bitmaps = [U.synth_image(w, h, 900 + 10 * k + i, 0) for i, (w, h) in enumerate(dims)]
        ref = ctx.run(U.page_images(dims, bitmaps), "vertical", platform="ios", deviceMaxCanvasSize=170)
        cases.append((dims, bitmaps, ref))
    ctx.set_lanes(3)
    try:
        outs, jobs, ress, tickets = [], [], [], []
        for dims, bitmaps, ref in cases:
            out = np.zeros_like(ref.rgba)
            job, keep = ctx.build_job(U.page_images(dims, bitmaps), "vertical", "min", 0, _profile_struct("ios", 170, None),
                                      "rgba", False, out.ctypes.data, 0, out.nbytes, 0)
            res = L.StitchResult()
            t = C.c_uint64()
            assert lib.stitch_submit(ctx.handle, C.byref(job), C.byref(res), C.byref(t)) == 0
            outs.append(out); jobs.append((job, keep)); ress.append(res); tickets.append(t.value)
        assert tickets == sorted(tickets)
        assert lib.stitch_ctx_set_lanes(ctx.handle, 2) == L.E_BUSY        # jobs in flight
        assert lib.stitch_wait(ctx.handle, tickets[-1]) == 0
        for out, (dims, bitmaps, ref) in zip(outs, cases):
            assert (out == ref.rgba).all()
        assert lib.stitch_ctx_join(ctx.handle) == 0
    finally:
        ctx.set_lanes(1)
    assert lib.stitch_ctx_set_lanes(ctx.handle, 0) == L.E_BAD_ARG
    assert lib.stitch_ctx_set_lanes(ctx.handle, 99) == L.E_BAD_ARG


def test_generator_matches_cpu_twin(ctx):
    import ctypes as C
    lib = ctx.lib
    for pattern in (0, 1, 2, 3, 4):
        w, h = 133, 57
        p = C.c_void_p()
        assert lib.stitch_device_alloc(ctx.handle, w * h * 4, C.byref(p)) == 0
        assert lib.stitch_generate_image(ctx.handle, p, w, h, w * 4, 0x5717C4 + pattern, pattern) == 0
        host = np.empty((h, w, 4), np.uint8)
        assert lib.stitch_memcpy_d2h(ctx.handle, host.ctypes.data, p, host.nbytes) == 0
        lib.stitch_device_free(ctx.handle, p)
        assert (host == U.synth_image(w, h, 0x5717C4 + pattern, pattern)).all(), pattern
