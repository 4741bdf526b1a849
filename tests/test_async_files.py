"""File-in / file-out jobs through stitch_submit + stitch_wait (GPU): nothing waits for the GPU at submit, and the results
equal the synchronous job's byte for byte.

index.js:1441-1463 (loadImageFrom inside the stitch loop) and 1577-1579 (the PNG export) are the steps a submitted job
now carries without host round trips: a fixed number of entropy-decoder synchronisation rounds gated on a device flag,
file sizes fetched when the job is waited for."""
import io

import numpy as np
import pytest

from tests import util as U

PIL = pytest.importorskip("PIL.Image")


def _jpeg(w, h, seed, quality=90, subsampling=2, **kw):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([(128 + 100 * np.sin(xx / (7.0 + seed % 5) + yy / 13.0)), (xx * 255.0 / max(1, w - 1)), (yy * 255.0 / max(1, h - 1))], -1)
    img = np.clip(img + rng.integers(-20, 20, img.shape), 0, 255).astype(np.uint8)
    b = io.BytesIO()
    PIL.fromarray(img, "RGB").save(b, "JPEG", quality=quality, subsampling=subsampling, **kw)
    return b.getvalue()


DIMS = [(806, 604), (604, 806), (1210, 908), (333, 200)]


@pytest.mark.gpu
@pytest.mark.parametrize("output", ["rgba", "png", "jpeg", "rgba+png", "yuv420+jpeg"])
def test_submitted_file_job_equals_synchronous_job(output):
    from imagestitching_b200 import StitchContext, prepare_image
    files = [_jpeg(w, h, seed=40 + i, subsampling=(2, 0, 1, 2)[i], restart_marker_blocks=(0, 4, 0, 0)[i]) for i, (w, h) in enumerate(DIMS)]
    imgs = [dict(prepare_image(f), file=f) for f in files]
    ctx = StitchContext(0)
    kw = dict(mode="min", gap=2, platform="ios", output=output, jpeg_quality=88)
    want = ctx.run(imgs, "vertical", **kw)
    ctx.set_lanes(3)
    # several jobs in flight, more jobs than lanes: lanes are reused while earlier jobs are still unfinalised
    handles = [ctx.submit(imgs if k % 2 == 0 else imgs[::-1], "vertical", **kw) for k in range(7)]
    outs = [ctx.wait(h) for h in handles]
    rev = ctx.run(imgs[::-1], "vertical", **kw)
    ctx.close()
    for k, got in enumerate(outs):
        ref = want if k % 2 == 0 else rev
        assert (got.width, got.height) == (ref.width, ref.height)
        if ref.rgba is not None:
            assert np.array_equal(got.rgba, ref.rgba)
        if ref.y is not None:
            assert np.array_equal(got.y, ref.y) and np.array_equal(got.cb, ref.cb) and np.array_equal(got.cr, ref.cr)
        assert got.png == ref.png and got.jpeg == ref.jpeg
        if ref.png is not None:
            assert len(got.png) > 1000
        if ref.jpeg is not None:
            assert len(got.jpeg) > 1000


@pytest.mark.gpu
def test_submitted_bitmap_job_with_file_output_and_errors_at_wait():
    from imagestitching_b200 import StitchContext, StitchError
    dims = [(806, 604), (604, 806)]
    bitmaps = [U.synth_image(w, h, 5 + i, 0) for i, (w, h) in enumerate(dims)]
    imgs = U.page_images(dims, bitmaps)
    ctx = StitchContext(0)
    ctx.set_lanes(2)
    want = ctx.run(imgs, "vertical", platform="ios", deviceMaxCanvasSize=700, output="png")
    ok = ctx.submit(imgs, "vertical", platform="ios", deviceMaxCanvasSize=700, output="png")
    small = ctx.submit(imgs, "vertical", platform="ios", deviceMaxCanvasSize=700, output="png", png_capacity=4096)
    assert ctx.wait(ok).png == want.png
    with pytest.raises(StitchError) as e:   # the size is only known on the device: the overflow surfaces at wait
        ctx.wait(small)
    assert "too small" in str(e.value) and str(len(want.png)) in str(e.value)
    again = ctx.submit(imgs, "vertical", platform="ios", deviceMaxCanvasSize=700, output="png")
    assert ctx.wait(again).png == want.png
    ctx.close()


@pytest.mark.gpu
def test_file_record_without_natural_dims_is_probed_from_the_header():
    """P:1227-1231: a record that lacks naturalWidth / naturalHeight is re-probed before the layout."""
    from imagestitching_b200 import StitchContext, prepare_image
    files = [_jpeg(w, h, seed=3 + i) for i, (w, h) in enumerate(DIMS[:2])]
    ctx = StitchContext(0)
    full = ctx.run([dict(prepare_image(f), file=f) for f in files], "horizontal", platform="ios")
    bare = ctx.run([{"file": f} for f in files], "horizontal", platform="ios")
    ctx.close()
    assert bare.layout == full.layout and np.array_equal(bare.rgba, full.rgba)
