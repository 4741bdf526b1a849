"""The pixel oracle checked against itself (numpy fp64 vs C fp64 vs C fp32) and against cv2 (no GPU).

The reference has no pixel fixtures (its resampler is WeChat's closed Canvas2D): "parity unpinned".
cv2.INTER_AREA / INTER_LINEAR are the independent anchors for the stated AB1 filter.
"""
import cv2
import numpy as np
import pytest

from oracle import layout_ref as LR
from oracle import raster_c as RC
from oracle import raster_ref as RR
from tests import util as U

SMALL = [(40, 30), (31, 47), (64, 64)]


def _noise(dims, seed, alpha=True):
    rng = np.random.default_rng(seed)
    out = []
    for (w, h) in dims:
        a = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        if not alpha:
            a[:, :, 3] = 255
        out.append(a)
    return out


@pytest.mark.parametrize("direction", ["vertical", "horizontal"])
@pytest.mark.parametrize("mode", ["min", "max", "original"])
def test_numpy_vs_c(direction, mode):
    bm = _noise(SMALL, 1)
    for gap in (0, 7):
        for prof in (LR.preset("ios"), LR.Profile("ios", 100, 3000), LR.Profile("other", 64, 0)):
            for orients in ([1, 1, 1], [2, 6, 7], [3, 5, 8], [4, 1, 6]):
                lay = LR.compute_layout(U.recs(SMALL, orients), direction, mode, gap, prof)
                a = RR.render_float(lay, bm, orients)
                c = RC.render(lay, bm, orients, want_yuv=True, return_float=True)
                assert np.abs(a - c["float"]).max() < 1e-9
                y, cb, cr = RR.yuv420_from_float(a)
                for p, q in ((RR.quantise(a), c["rgba"][:, :, :3]), (y, c["y"]), (cb, c["cb"]), (cr, c["cr"])):
                    d = np.abs(p.astype(int) - q.astype(int))
                    assert d.max() <= 1 and d.sum() <= 2          # exact ties may round either way
                assert (c["rgba"][:, :, 3] == 255).all()


def test_f32_baseline_matches_f64():
    """The timed CPU baseline (fp32) is the same filter: within 1 LSB, PSNR >= 60 dB of the fp64 checker."""
    dims = [(403, 302)] * 3
    bm = [U.synth_image(w, h, 9 + i, 1) for i, (w, h) in enumerate(dims)]
    lay = LR.compute_layout(U.recs(dims), "vertical", "min", 0, LR.Profile("ios", 400, 0).with_fallback())
    a = RC.render(lay, bm, [1] * 3, precision="f64")["rgba"]
    b = RC.render(lay, bm, [1] * 3, precision="f32")["rgba"]
    U.check_pixels(b, a, "f32 vs f64", min_psnr=60.0)


@pytest.mark.parametrize("W,H", [(504, 378), (555, 400), (252, 189), (300, 200)])
def test_upscale_is_cv2_inter_linear(W, H):
    rng = np.random.default_rng(0)
    src = rng.integers(0, 256, (189, 252, 4), dtype=np.uint8)
    src[:, :, 3] = 255
    lay = LR.Layout(n=1, canvasW=W, canvasH=H, superSample=1.0, rects=[(0.0, 0.0, float(W), float(H))])
    got = RC.render(lay, [src], [1])["rgba"]
    ref = cv2.resize(src, (W, H), interpolation=cv2.INTER_LINEAR)
    d = np.abs(got.astype(int) - ref.astype(int))
    assert d.max() <= 1 and RR.psnr(got, ref) >= 55


@pytest.mark.parametrize("W,H", [(126, 63), (100, 77), (113, 85), (252, 189), (84, 63)])
def test_downscale_is_cv2_inter_area(W, H):
    rng = np.random.default_rng(1)
    src = rng.integers(0, 256, (189, 252, 4), dtype=np.uint8)
    src[:, :, 3] = 255
    lay = LR.Layout(n=1, canvasW=W, canvasH=H, superSample=1.0, rects=[(0.0, 0.0, float(W), float(H))])
    got = RC.render(lay, [src], [1])["rgba"]
    ref = cv2.resize(src, (W, H), interpolation=cv2.INTER_AREA)
    d = np.abs(got.astype(int) - ref.astype(int))
    assert d.max() <= 1 and RR.psnr(got, ref) >= 55


def test_orientation_table_matches_ctm_composition():
    """placement_for() (closed form) == composing the CTMs of utils/canvas.js:160-200."""
    import math

    def ctm_for(o, dx, dy, dw, dh):
        # matrix as canvas: x' = a x + c y + e ; y' = b x + d y + f
        M = [1.0, 0.0, 0.0, 1.0, 0.0, 0.0]

        def mul(m, n):
            a, b, c, d, e, f = m
            a2, b2, c2, d2, e2, f2 = n
            return [a * a2 + c * b2, b * a2 + d * b2, a * c2 + c * d2, b * c2 + d * d2, a * e2 + c * f2 + e, b * e2 + d * f2 + f]

        def translate(x, y):
            nonlocal M
            M = mul(M, [1, 0, 0, 1, x, y])

        def rotate(t):
            nonlocal M
            c, s = round(math.cos(t)), round(math.sin(t))
            M = mul(M, [c, s, -s, c, 0, 0])

        def scale(x, y):
            nonlocal M
            M = mul(M, [x, 0, 0, y, 0, 0])

        if o == 2:
            translate(dx + dw, dy); scale(-1, 1); local = (dw, dh)
        elif o == 3:
            translate(dx + dw, dy + dh); rotate(math.pi); local = (dw, dh)
        elif o == 4:
            translate(dx, dy + dh); scale(1, -1); local = (dw, dh)
        elif o == 5:
            translate(dx, dy); rotate(0.5 * math.pi); scale(1, -1); local = (dh, dw)
        elif o == 6:
            translate(dx + dw, dy); rotate(0.5 * math.pi); local = (dh, dw)
        elif o == 7:
            translate(dx + dw, dy); rotate(0.5 * math.pi); scale(-1, 1); local = (dh, dw)
        elif o == 8:
            translate(dx, dy + dh); rotate(-0.5 * math.pi); local = (dh, dw)
        else:
            translate(dx, dy); local = (dw, dh)
        return M, local

    dx, dy, dw, dh = 10.0, 20.0, 30.0, 50.0
    for o in range(0, 10):
        M, (lw, lh) = ctm_for(o, dx, dy, dw, dh)
        a, b, c, d, e, f = M
        # device position of the bitmap's unit corners (u,v in [0,1]) through the CTM
        def dev(u, v):
            x, y = u * lw, v * lh
            return (a * x + c * y + e, b * x + d * y + f)
        p = LR.placement_for((dx, dy, dw, dh), o, 1.0)

        def mine(u, v):
            if p.transpose:
                uu, vv = (1 - u if p.flip_u else u), (1 - v if p.flip_v else v)
                return (p.X0 + vv * p.W, p.Y0 + uu * p.H)
            uu, vv = (1 - u if p.flip_u else u), (1 - v if p.flip_v else v)
            return (p.X0 + uu * p.W, p.Y0 + vv * p.H)
        for (u, v) in ((0, 0), (1, 0), (0, 1), (1, 1), (0.25, 0.75)):
            gx, gy = dev(u, v)
            mx, my = mine(u, v)
            assert abs(gx - mx) < 1e-9 and abs(gy - my) < 1e-9, (o, u, v, (gx, gy), (mx, my))


def test_synth_image_patterns():
    for pattern in range(5):
        a = U.synth_image(33, 17, 5, pattern)
        assert a.shape == (17, 33, 4) and a.dtype == np.uint8
        if pattern != 2:
            assert (a[:, :, 3] == 255).all()
