"""Device bitmaps in allocations of EXACTLY pitch*(h-1) + 4*w bytes (GPU): nothing may be read past the last pixel.

Round 1 staged source rows with 16-byte per-lane copies, which read up to 12 bytes past the end of a bitmap whose width
is no multiple of 4.  The TMA ring addresses the bitmap as a tensor of {bmp_width, bmp_height} u32 elements,
so the hardware zero-fills what lies beyond the last column instead of fetching it.  Run under
`compute-sanitizer --tool memcheck` (profiles/sanitizer.sh) this file is the proof; without it it still checks the pixels
of the right-edge groups that the zero fill touches.  Every draw path is covered: streaming (downscale), upscale, gather
(EXIF 6, and a 16-byte-misaligned base), and the canvas itself is an exact-size device allocation too."""
import ctypes as C

import numpy as np
import pytest

from tests import util as U


def _device_bitmap(ctx, a, pitch, offset=0):
    """Uploads a [h, w, 4] array into a device allocation of exactly offset + pitch*(h-1) + 4*w bytes."""
    from imagestitching_b200 import DeviceBitmap
    h, w = a.shape[:2]
    size = pitch * (h - 1) + 4 * w
    host = np.zeros(pitch * h, np.uint8)
    host.reshape(h, pitch)[:, : 4 * w] = a.reshape(h, 4 * w)
    p = C.c_void_p()
    assert ctx.lib.stitch_device_alloc(ctx.handle, offset + size, C.byref(p)) == 0
    assert ctx.lib.stitch_memcpy_h2d(ctx.handle, p.value + offset, host.ctypes.data, size) == 0
    return DeviceBitmap(p.value + offset, w, h, pitch, keepalive=p), p


def _run_device(ctx, dims, bitmaps, pitches, direction, orientations=None, offsets=None, **kw):
    """Job with exact-size device bitmaps AND an exact-size device canvas; returns (layout, rgba)."""
    from imagestitching_b200 import _lib as L
    from imagestitching_b200.stitch import _profile_struct
    dev, ptrs = [], []
    for i, a in enumerate(bitmaps):
        d, p = _device_bitmap(ctx, a, pitches[i], (offsets or [0] * len(bitmaps))[i])
        dev.append(d)
        ptrs.append(p)
    imgs = U.page_images(dims, dev, orientations)
    prof = _profile_struct(kw.get("platform", "ios"), kw.get("deviceMaxCanvasSize"), kw.get("deviceMaxCanvasPixels"))
    from imagestitching_b200 import compute_layout
    lay = compute_layout(imgs, direction, kw.get("mode", "min"), kw.get("gap", 0), kw.get("platform", "ios"),
                         kw.get("deviceMaxCanvasSize"), kw.get("deviceMaxCanvasPixels"))
    cw, ch = int(lay["canvasW"]), int(lay["canvasH"])
    out = C.c_void_p()
    assert ctx.lib.stitch_device_alloc(ctx.handle, cw * ch * 4, C.byref(out)) == 0
    job, keep = ctx.build_job(imgs, direction, kw.get("mode", "min"), kw.get("gap", 0), prof, "rgba", True, out.value, 0, cw * ch * 4, 0)
    res = L.StitchResult()
    st = ctx.lib.stitch_run(ctx.handle, C.byref(job), C.byref(res))
    assert st == 0, ctx.lib.stitch_last_error(ctx.handle)
    rgba = np.empty((ch, cw, 4), np.uint8)
    assert ctx.lib.stitch_memcpy_d2h(ctx.handle, rgba.ctypes.data, out.value, rgba.nbytes) == 0
    for p in ptrs + [out]:
        ctx.lib.stitch_device_free(ctx.handle, p)
    return res, rgba


@pytest.mark.gpu
@pytest.mark.parametrize("wmod", [0, 1, 2, 3])
def test_streaming_units_on_tight_device_bitmaps(wmod):
    from imagestitching_b200 import StitchContext
    ctx = StitchContext(0)
    # widths = 4k + wmod and 128k + small remainders: the right-edge unit's box reaches past the bitmap
    dims = [(516 + wmod, 301), (640 + wmod, 207), (132 + wmod, 97)]
    bitmaps = [U.synth_image(w, h, 11 * wmod + i, 1) for i, (w, h) in enumerate(dims)]
    pitches = [(4 * w + 15) & ~15 for (w, _) in dims]
    prof = U.profile_of("ios", 120, 0)
    lay, want = U.oracle_render(dims, bitmaps, "vertical", "min", 0, prof)
    res, rgba = _run_device(ctx, dims, bitmaps, pitches, "vertical", platform="ios", deviceMaxCanvasSize=120)
    ctx.close()
    assert (res.path_info & 0xFF) == 3, hex(res.path_info)      # three streaming regions
    U.check_pixels(rgba[:, :, :3], want["rgba"][:, :, :3], f"tight stream w%4={wmod}")


@pytest.mark.gpu
def test_upscale_and_gather_units_on_tight_device_bitmaps():
    from imagestitching_b200 import StitchContext
    ctx = StitchContext(0)
    # supersampled canvas (bilinear upscale units) from bitmaps of odd widths
    dims = [(109, 61), (110, 75), (111, 40)]
    bitmaps = [U.synth_image(w, h, 70 + i, 0) for i, (w, h) in enumerate(dims)]
    pitches = [(4 * w + 15) & ~15 for (w, _) in dims]
    lay, want = U.oracle_render(dims, bitmaps, "vertical", "max", 1, U.profile_of("ios"))
    res, rgba = _run_device(ctx, dims, bitmaps, pitches, "vertical", mode="max", gap=1, platform="ios")
    assert (res.path_info >> 25) & 63, hex(res.path_info)       # upscale regions were used
    U.check_pixels(rgba[:, :, :3], want["rgba"][:, :, :3], "tight upscale")
    # EXIF 6 / 8 (transposed: gather path) and a bitmap whose base is only 4-byte aligned with a tight 4*w pitch
    dims = [(301, 203), (203, 301), (257, 131)]
    orient = [6, 8, 1]
    bitmaps = [U.synth_image(w, h, 90 + i, 1) for i, (w, h) in enumerate(dims)]
    pitches = [4 * w for (w, _) in dims]
    prof = U.profile_of("ios", 150, 0)
    lay, want = U.oracle_render(dims, bitmaps, "horizontal", "min", 0, prof, orientations=orient)
    res, rgba = _run_device(ctx, dims, bitmaps, pitches, "horizontal", orientations=orient, offsets=[0, 4, 12],
                            platform="ios", deviceMaxCanvasSize=150)
    ctx.close()
    U.check_pixels(rgba[:, :, :3], want["rgba"][:, :, :3], "tight gather")
