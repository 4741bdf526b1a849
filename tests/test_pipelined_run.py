"""stitch_run's pipelined form (GPU): per-image upload -> per-image launch -> per-image download on a second stream.

The reference only ever runs one job at a time (isStitching, index.js:772,781), so stitch_run's latency is the page's
latency.  The per-image launches run SUBSETS of the fused plan's regions, so the canvas must equal the fused launch's
(stitch_submit never pipelines) byte for byte; the 30..90 progress values (index.js:1556-1557) keep their order."""
import numpy as np
import pytest

from tests import util as U

CASES = [
    # dims, direction, kwargs -- vertical full-width bands, horizontal column bands, centred 'original' rects with gaps
    ([(806, 604), (604, 806), (1210, 908), (808, 600)], "vertical", dict(mode="min", gap=0, platform="ios", deviceMaxCanvasSize=700)),
    ([(806, 604), (604, 806), (1210, 908)], "horizontal", dict(mode="min", gap=3, platform="ios", deviceMaxCanvasSize=900)),
    ([(806, 604), (604, 806), (300, 200)], "vertical", dict(mode="original", gap=5, platform="ios", deviceMaxCanvasSize=1000)),
    ([(800, 600), (600, 800)], "vertical", dict(mode="max", gap=1, platform="ios")),                       # supersampled: upscale units
    ([(806, 604), (604, 806), (806, 604)], "horizontal", dict(mode="max", gap=0, platform="android")),
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", range(len(CASES)))
def test_pipelined_run_equals_fused_launch(case):
    from imagestitching_b200 import StitchContext
    dims, direction, kw = CASES[case]
    orient = [1, 3, 1, 2][: len(dims)] if case == 0 else ([6, 1, 1][: len(dims)] if case == 2 else None)
    bitmaps = [U.synth_image(w, h, 13 * case + i, i % 2) for i, (w, h) in enumerate(dims)]
    imgs = U.page_images(dims, bitmaps, orient)
    ctx = StitchContext(0)
    seen = []
    piped = ctx.run(imgs, direction, on_progress=seen.append, **kw)          # stitch_run: pipelined (>= 1 MB of bitmaps)
    fused = ctx.wait(ctx.submit(imgs, direction, **kw))                       # stitch_submit: one fused launch
    ctx.close()
    assert fused.kernel_launches == 1
    if fused.path_info & (1 << 24):      # no fast region at all (case 2: EXIF 6 + fractional rects): nothing to pipeline
        assert piped.kernel_launches == 1
    else:
        assert piped.kernel_launches >= 2
    assert (piped.width, piped.height) == (fused.width, fused.height)
    assert np.array_equal(piped.rgba, fused.rgba)
    from oracle import layout_ref as LR
    assert seen == LR.progress_sequence(len(dims))


@pytest.mark.gpu
def test_pipelined_run_at_full_size_against_oracle():
    """BASELINE configs[1] shape at a quarter of the linear size, against the fp64 oracle."""
    from imagestitching_b200 import StitchContext
    dims = [(1008, 756)] * 9
    bitmaps = [U.synth_image(w, h, 200 + i, 0) for i, (w, h) in enumerate(dims)]
    prof = U.profile_of("ios", 3072, 0)
    lay, want = U.oracle_render(dims, bitmaps, "vertical", "min", 0, prof)
    ctx = StitchContext(0)
    got = ctx.run(U.page_images(dims, bitmaps), "vertical", platform="ios", deviceMaxCanvasSize=3072)
    ctx.close()
    assert got.kernel_launches == 10            # 9 images + the white tail rows
    U.check_pixels(got.rgba[:, :, :3], want["rgba"][:, :, :3], "pipelined cfg2/4")
