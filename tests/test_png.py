"""PNG export (index.js:1577-1579 -> canvasToTempFilePath fileType 'png', utils/canvas.js:205-242).

CPU: the oracle writer (oracle/png_ref.c) is pinned by decoding its files with libpng/zlib (PIL): lossless.
GPU: STITCH_OUT_PNG files are byte-identical to the oracle writer's and decode to the RGBA canvas exactly.
"""
import io
import zlib

import numpy as np
import pytest
from PIL import Image

from oracle import png_c as P
from tests import util as U

SIZES = [(1, 1), (3, 2), (17, 9), (64, 48), (123, 77), (333, 200), (700, 9)]


def _back(data):
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGBA"))


@pytest.mark.parametrize("size", SIZES, ids=[f"{w}x{h}" for w, h in SIZES])
def test_oracle_png_roundtrip(size):
    w, h = size
    rng = np.random.default_rng(w * 1000 + h)
    for a in (rng.integers(0, 256, (h, w, 4), dtype=np.uint8), np.full((h, w, 4), 255, np.uint8), U.synth_image(w, h, 5, 0)):
        data = P.encode_rgba(a)
        assert data[:8] == b"\x89PNG\r\n\x1a\n" and data[-12:] == b"\0\0\0\0IEND\xaeB`\x82"
        assert np.array_equal(_back(data), a)


def test_oracle_crc_and_length_limit():
    for s in (b"", b"IEND", b"IDAT" + bytes(range(256)) * 7):
        assert P.crc32(s) == zlib.crc32(s)
    f = np.zeros(286, np.uint32)
    f[:40] = [int(1.6 ** i) + 1 for i in range(40)]      # Fibonacci-like skew: plain Huffman would need > 15 bits
    f[256] = 1
    lens = P.code_lengths(f)
    assert lens.max() == 15 and abs(sum(2.0 ** -int(x) for x in lens if x) - 1.0) < 1e-12
    assert (lens[f == 0] == 0).all() and (lens[f > 0] > 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize("case", [
    ([(806, 604), (604, 806), (806, 604)], "vertical", "min", 700, 0, "rgba+png"),
    ([(320, 200), (100, 260)], "horizontal", "original", 0, 7, "png"),
    ([(64, 48)], "vertical", "min", 0, 0, "rgba+png"),                   # one band
    ([(1080, 1920), (1080, 1920)], "vertical", "min", 0, 0, "png"),      # cfg1: 2376 x 8448 canvas
], ids=["three", "original-gap", "tiny", "cfg1"])
def test_gpu_png_equals_oracle_and_is_lossless(case):
    from imagestitching_b200 import StitchContext
    dims, direction, mode, size, gap, output = case
    bitmaps = [U.synth_image(w, h, 300 + i, 0) for i, (w, h) in enumerate(dims)]
    ctx = StitchContext(0)
    ref = ctx.run(U.page_images(dims, bitmaps), direction, mode=mode, gap=gap, platform="ios", deviceMaxCanvasSize=size or None)
    got = ctx.run(U.page_images(dims, bitmaps), direction, mode=mode, gap=gap, platform="ios", deviceMaxCanvasSize=size or None,
                  output=output)
    ctx.close()
    assert got.png is not None and got.kernel_launches == 5
    back = _back(got.png)
    assert np.array_equal(back, ref.rgba), "PNG does not decode to the canvas"
    if got.rgba is not None:
        assert np.array_equal(got.rgba, ref.rgba)
    want = P.encode_rgba(ref.rgba)
    assert len(got.png) == len(want) and got.png == want, "GPU PNG differs from the oracle writer"


@pytest.mark.gpu
def test_gpu_png_randomized_and_flat_canvases():
    """Random job shapes (band counts, widths, gaps -> white areas) and a canvas that is almost all white."""
    from imagestitching_b200 import StitchContext
    rng = np.random.default_rng(77)
    ctx = StitchContext(0)
    for k in range(8):
        n = int(rng.integers(1, 5))
        dims = [(int(rng.integers(8, 500)), int(rng.integers(8, 400))) for _ in range(n)]
        bitmaps = [U.synth_image(w, h, 40 * k + i, int(rng.integers(0, 2))) for i, (w, h) in enumerate(dims)]
        direction = "vertical" if rng.random() < 0.5 else "horizontal"
        mode = ["min", "max", "original"][int(rng.integers(0, 3))]
        got = ctx.run(U.page_images(dims, bitmaps), direction, mode=mode, gap=int(rng.integers(0, 21)), platform="ios",
                      output="rgba+png")
        assert np.array_equal(_back(got.png), got.rgba)
        assert got.png == P.encode_rgba(got.rgba), f"job {k}: GPU PNG differs from the oracle writer"
    # 'original' mode with one tiny and one wide image: the canvas is mostly background
    dims = [(16, 16), (900, 20)]
    bitmaps = [np.full((h, w, 4), 255, np.uint8) for w, h in dims]
    got = ctx.run(U.page_images(dims, bitmaps), "vertical", mode="original", gap=20, platform="ios", output="rgba+png")
    assert got.png == P.encode_rgba(got.rgba) and len(got.png) < got.rgba.nbytes // 5
    ctx.close()


def _screenshot_like(w, h, seed):
    """Flat colour blocks with a few noisy stripes: long runs of equal filtered bytes, ends at arbitrary offsets."""
    rng = np.random.default_rng(seed)
    img = np.empty((h, w, 4), np.uint8)
    img[:] = rng.integers(0, 256, 4, dtype=np.uint8)
    for _ in range(6):
        x0, y0 = int(rng.integers(0, w)), int(rng.integers(0, h))
        x1, y1 = int(rng.integers(x0, w)) + 1, int(rng.integers(y0, h)) + 1
        img[y0:y1, x0:x1] = rng.integers(0, 256, 4, dtype=np.uint8)
    if h > 4:
        y = int(rng.integers(0, h - 2))
        img[y:y + 2] = rng.integers(0, 256, (2, w, 4), dtype=np.uint8)
    img[..., 3] = 255
    return img


def test_oracle_runs_roundtrip_and_shrink():
    """Run-length tokens: flat content decodes exactly through libpng and is far smaller than literal coding."""
    for k, (w, h) in enumerate([(1, 1), (1, 40), (2, 17), (3, 9), (129, 8), (512, 8), (513, 9), (700, 33), (1031, 20)]):
        img = _screenshot_like(w, h, k)
        data = P.encode_rgba(img)
        assert np.array_equal(_back(data), img), (w, h)
    white = np.full((200, 333, 4), 255, np.uint8)
    assert len(P.encode_rgba(white)) < white.nbytes // 40


@pytest.mark.gpu
def test_gpu_png_run_tokens():
    """Screenshot-like bitmaps (flat blocks -> long runs, ends anywhere in a round): the GPU file equals the writer's."""
    from imagestitching_b200 import StitchContext
    ctx = StitchContext(0)
    for k, (w, h) in enumerate([(1, 1), (1, 40), (2, 17), (3, 9), (129, 8), (512, 8), (513, 9), (700, 33), (1031, 20), (1820, 64)]):
        img = _screenshot_like(w, h, 100 + k)
        got = ctx.run(U.page_images([(w, h)], [img]), "vertical", mode="original", gap=0, platform="ios", output="rgba+png")
        assert got.png == P.encode_rgba(got.rgba), f"{w}x{h}: GPU PNG differs from the oracle writer"
        assert np.array_equal(_back(got.png), got.rgba)
    ctx.close()


@pytest.mark.gpu
def test_gpu_png_capacity_error():
    from imagestitching_b200 import StitchContext, StitchError
    from imagestitching_b200 import _lib as L
    dims = [(320, 200), (320, 200)]
    bitmaps = [U.synth_image(w, h, 1, 0) for w, h in dims]
    ctx = StitchContext(0)
    with pytest.raises(StitchError) as e:
        ctx.run(U.page_images(dims, bitmaps), "vertical", output="png", png_capacity=4096)
    assert e.value.status == L.E_NO_SURFACE and "bytes needed" in e.value.message
    ctx.close()
