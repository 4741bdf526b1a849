"""Preview fit-scale (index.js:1593-1603, utils/canvas.js:17-24).

CPU: the C-ABI layout twin against the Python restatement (bit-exact doubles), numpy oracle against its C twin.
GPU: stitch_preview against the oracle -- straight-alpha RGBA8, +-1 LSB, PSNR >= 50 dB like the stitch path.
"""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from imagestitching_b200 import StitchError, preview_layout
from imagestitching_b200 import _lib as L
from oracle import layout_ref as LR
from oracle import raster_c as RC
from oracle import raster_ref as RR
from tests import util as U


def _same_layout(got, want):
    assert (got["canvasW"], got["canvasH"], got["pvW"], got["pvH"]) == (want.canvasW, want.canvasH, want.pvW, want.pvH)
    assert (got["scaleFit"], got["dx"], got["dy"], got["dpr"]) == (want.scaleFit, want.dx, want.dy, want.dpr)


def test_preview_layout_hand_checked():
    # cfg2 export 1820 x 12288 on a 311 x 420.5 CSS px canvas at dpr 3: height-bound fit
    pl = preview_layout(1820, 12288, 311, 420.5, 3)
    assert (pl["canvasW"], pl["canvasH"]) == (933, 1261)                 # floor(311*3), floor(420.5*3)
    assert pl["scaleFit"] == 420.5 / 12288
    assert (pl["pvW"], pl["pvH"]) == (62, 421)                           # round(62.28), round(420.5) (half up)
    assert (pl["dx"], pl["dy"]) == (124.5, -0.25)                        # the image overhangs the canvas by 0.25 px
    _same_layout(pl, LR.preview_layout(1820, 12288, 311, 420.5, 3))
    # pixelRatio || 1
    assert preview_layout(100, 50, 200, 200, 0)["dpr"] == 1.0
    assert preview_layout(100, 50, 200, 200, 0)["canvasW"] == 200


@settings(max_examples=300, deadline=None)
@given(ew=st.integers(1, 12288), eh=st.integers(1, 12288), vw=st.floats(1, 1400), vh=st.floats(1, 2800),
       dpr=st.sampled_from([1, 1.5, 2, 2.625, 2.75, 3, 3.5]))
def test_preview_layout_twin_equals_restatement(ew, eh, vw, vh, dpr):
    _same_layout(preview_layout(ew, eh, vw, vh, dpr), LR.preview_layout(ew, eh, vw, vh, dpr))


def test_preview_layout_errors():
    with pytest.raises(StitchError) as e:
        preview_layout(0, 10, 100, 100, 2)
    assert e.value.status == L.E_BAD_DIMS
    with pytest.raises(StitchError) as e:
        preview_layout(10, 10, 0.2, 100, 2)
    assert e.value.status == L.E_BAD_OUT


CASES = [  # export w, h, view w, h, dpr, translucent source
    (182, 1228, 311.0, 420.5, 3, False),     # downscale, overhang above / below
    (640, 200, 343.0, 180.0, 2.75, True),    # fractional dpr, translucent export
    (50, 40, 343.0, 600.0, 2, False),        # upscale
    (300, 300, 300.0, 300.0, 1, False),      # identity
    (333, 77, 97.5, 211.25, 1.5, True),
]


@pytest.mark.parametrize("case", CASES, ids=[f"{c[0]}x{c[1]}" for c in CASES])
def test_preview_oracle_numpy_vs_c(case):
    ew, eh, vw, vh, dpr, translucent = case
    bmp = U.synth_image(ew, eh, 0xBEEF + ew, 0)
    if translucent:
        bmp[..., 3] = (np.arange(ew)[None, :] * 7 + np.arange(eh)[:, None] * 3) % 256
    pl = LR.preview_layout(ew, eh, vw, vh, dpr)
    a, b = RR.render_preview(bmp, pl), RC.render_preview(bmp, pl)
    assert a.shape == (pl.canvasH, pl.canvasW, 4)
    assert int(np.abs(a.astype(np.int16) - b.astype(np.int16)).max()) <= 1     # summation order only
    assert np.count_nonzero(a != b) <= a.size // 500
    # transparent outside the fit rect (clearRect, index.js:1598)
    p = LR.preview_placement(pl)
    x0, x1 = int(np.floor(p.X0)), int(np.ceil(p.X0 + p.W))
    if x0 > 0:
        assert not a[:, :x0].any()
    if x1 < pl.canvasW:
        assert not a[:, x1:].any()
    if not translucent:                       # opaque inside
        xi0, xi1 = int(np.ceil(p.X0)), int(np.floor(p.X0 + p.W))
        yi0, yi1 = max(0, int(np.ceil(p.Y0))), min(pl.canvasH, int(np.floor(p.Y0 + p.H)))
        assert (a[yi0:yi1, xi0:xi1, 3] == 255).all()


GPU_CASES = CASES + [
    (1820, 12288, 311.0, 420.5, 3, False),   # the cfg2 export on a phone-sized canvas
    (2376, 8448, 343.0, 1219.6, 3, False),   # the cfg1 export, width-bound
    (4096, 606, 343.0, 180.0, 2, True),
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", GPU_CASES, ids=[f"{c[0]}x{c[1]}" for c in GPU_CASES])
def test_preview_gpu_vs_oracle(case):
    from imagestitching_b200 import StitchContext
    ew, eh, vw, vh, dpr, translucent = case
    bmp = U.synth_image(ew, eh, 0xBEEF + ew, 0)
    if translucent:
        bmp[..., 3] = (np.arange(ew)[None, :] * 7 + np.arange(eh)[:, None] * 3) % 256
    pl = LR.preview_layout(ew, eh, vw, vh, dpr)
    want = RC.render_preview(bmp, pl)
    ctx = StitchContext(0)
    got = ctx.preview(bmp, vw, vh, dpr)
    ctx.close()
    assert got.shape == want.shape
    if translucent:
        # un-premultiplying amplifies the filter's fp32 rounding by 255/alpha: hold colour only where alpha >= 16
        assert int(np.abs(got[..., 3].astype(np.int16) - want[..., 3].astype(np.int16)).max()) <= 1
        m = want[..., 3] >= 16
        d = np.abs(got[..., :3].astype(np.int16) - want[..., :3].astype(np.int16))[m]
        assert int(d.max()) <= 1
    else:
        U.check_pixels(got, want, f"preview {case}")
        assert (got[..., 3] == want[..., 3]).mean() > 0.99     # half-covered border columns are rounding ties


@pytest.mark.gpu
def test_preview_of_a_stitch_result():
    """The page's own sequence: stitch, export, then preview the export (index.js:1577-1603)."""
    from imagestitching_b200 import StitchContext
    dims = [(806, 604), (604, 806), (806, 604)]
    bitmaps = [U.synth_image(w, h, 77 + i, 0) for i, (w, h) in enumerate(dims)]
    ctx = StitchContext(0)
    out = ctx.run(U.page_images(dims, bitmaps), "vertical", mode="min", gap=4, platform="ios")
    got = ctx.preview(out.rgba, 343.0, 520.0, 3)
    ctx.close()
    pl = LR.preview_layout(out.width, out.height, 343.0, 520.0, 3)
    U.check_pixels(got, RC.render_preview(out.rgba, pl), "preview of stitch")
