"""JPEG export (utils/canvas.js:205-242 with fileType 'jpg').

CPU part: the sequential oracle encoder (oracle/jpeg_ref.c) is pinned by an independent decoder (libjpeg inside
PIL): its files decode, to the planes it was given, within the quantisation error of the quality setting.
GPU part: the CUDA encoder's file is byte-identical to the oracle's for the same I420 planes."""
import io
import os
import re

import numpy as np
import pytest

from oracle import jpeg_c as J
from tests import util as U

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _planes(w, h, seed, noise=3.0):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    y = 128 + 90 * np.sin(xx / 17.0) * np.cos(yy / 23.0) + rng.normal(0, noise, (h, w))
    ch, cw = (h + 1) // 2, (w + 1) // 2
    yy, xx = np.mgrid[0:ch, 0:cw]
    cb = 128 + 60 * np.cos(xx / 9.0) + rng.normal(0, noise, (ch, cw))
    cr = 128 + 60 * np.sin((xx + yy) / 13.0) + rng.normal(0, noise, (ch, cw))
    return [np.clip(np.floor(p + 0.5), 0, 255).astype(np.uint8) for p in (y, cb, cr)]


def _decode_ycc(data):
    from PIL import Image
    im = Image.open(io.BytesIO(data))
    assert im.format == "JPEG" and im.mode == "YCbCr" or im.mode == "RGB"
    im.draft("YCbCr", im.size)
    return im.size, np.asarray(im.convert("YCbCr") if im.mode != "YCbCr" else im)


def _psnr(a, b):
    mse = ((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean()
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


@pytest.mark.parametrize("w,h,q,ri,min_psnr", [
    (211, 123, 100, 8, 44.0), (211, 123, 90, 8, 38.0), (64, 48, 75, 0, 34.0), (17, 9, 50, 3, 28.0),
    (8, 8, 100, 1, 44.0), (1, 1, 100, 8, 30.0), (333, 200, 95, 5, 40.0),
])
def test_oracle_files_decode_with_libjpeg(w, h, q, ri, min_psnr):
    y, cb, cr = _planes(w, h, w * 1000 + h)
    data = J.encode_i420(y, cb, cr, q, ri)
    assert data[:2] == b"\xff\xd8" and data[-2:] == b"\xff\xd9"
    size, ycc = _decode_ycc(data)
    assert size == (w, h)
    assert _psnr(ycc[:, :, 0], y) >= min_psnr, _psnr(ycc[:, :, 0], y)
    # restart markers appear in order
    marks = [m.start() for m in re.finditer(b"\xff[\xd0-\xd7]", data)]
    n_mcu = ((w + 15) // 16) * ((h + 15) // 16)
    if ri:
        assert len(marks) >= (n_mcu + ri - 1) // ri - 1


def test_quality_scaling_and_tables():
    assert (J.quant_table(0, 100) == 1).all() and (J.quant_table(1, 100) == 1).all()
    q50 = J.quant_table(0, 50)
    assert q50[0] == 16 and q50[63] == 99 and q50[1] == 11
    assert J.quant_table(0, 1).max() == 255
    code, length = J.huff_tables()
    # DC luma category 0 is the 2-bit code 00, EOB of AC luma is 1010 (T.81 tables K.3, K.5)
    assert (code[0, 0], length[0, 0]) == (0, 2)
    assert (code[2, 0x00], length[2, 0x00]) == (0b1010, 4)
    assert (code[2, 0xF0], length[2, 0xF0]) == (0b11111111001, 11)
    assert (code[3, 0x00], length[3, 0x00]) == (0b00, 2)
    for t in range(4):   # prefix-free: Kraft sum of the used codes is <= 1
        used = length[t][length[t] > 0].astype(np.float64)
        assert (2.0 ** -used).sum() <= 1.0


def test_product_tables_equal_oracle_tables():
    """imagestitching_b200/csrc/jpeg_tables.h carries its own copy of the Annex K tables."""
    def nums(path, name):
        src = open(os.path.join(ROOT, path), encoding="utf-8").read()
        m = re.search(name + r"\[\d+\]\s*=\s*\{([^}]*)\}", src)
        return [int(x, 0) for x in m.group(1).replace("\n", " ").split(",") if x.strip()]
    for name in ("JQ_LUMA", "JQ_CHROMA", "JZIGZAG", "JDC_LUMA_BITS", "JDC_LUMA_VAL", "JDC_CHROMA_BITS", "JDC_CHROMA_VAL",
                 "JAC_LUMA_BITS", "JAC_LUMA_VAL", "JAC_CHROMA_BITS", "JAC_CHROMA_VAL"):
        assert nums("oracle/jpeg_tables.inc", name) == nums("imagestitching_b200/csrc/jpeg_tables.h", name), name


# ---------------------------------------------------------------- GPU
BIG = 26 * 1024 * 1024


def _gpu_case(ctx, dims, direction, size, q, ri, output="both+jpeg", mode="min", pattern=0):
    bitmaps = [U.synth_image(w, h, 321 + i, pattern) for i, (w, h) in enumerate(dims)]
    fs = [BIG] + [0] * (len(dims) - 1)
    got = ctx.run(U.page_images(dims, bitmaps, None, fs), direction, mode=mode, platform="ios",
                  deviceMaxCanvasSize=size or None, output=output, jpeg_quality=q, jpeg_restart_interval=ri)
    return got


@pytest.mark.gpu
@pytest.mark.parametrize("dims,direction,size,q,ri", [
    ([(403, 302)] * 2, "vertical", 300, 100, 8),
    ([(403, 302)] * 2, "vertical", 300, 90, 8),
    ([(640, 480), (333, 211), (500, 500)], "horizontal", 700, 75, 3),
    ([(37, 29)], "vertical", 20, 50, 1),
    ([(1612, 1208)] * 3, "vertical", 1638, 95, 8),
    ([(200, 150)] * 2, "vertical", 0, 100, 65535),
])
def test_gpu_file_is_byte_identical_to_oracle(ctx, dims, direction, size, q, ri):
    got = _gpu_case(ctx, dims, direction, size, q, ri)
    want = J.encode_i420(got.y, got.cb, got.cr, q or 100, ri or 8)
    assert len(got.jpeg) == len(want), (len(got.jpeg), len(want))
    assert got.jpeg == want
    size_, ycc = _decode_ycc(got.jpeg)
    assert size_ == (got.width, got.height)
    assert _psnr(ycc[:, :, 0], got.y) >= (40.0 if q >= 95 else 36.0 if q >= 90 else 28.0)


@pytest.mark.gpu
@pytest.mark.parametrize("dims,q,ri,pattern", [
    ([(256, 160)], 100, 8, 1),       # uniform noise copied 1:1, all-ones tables: ~60 long codes per block, so a
    ([(256, 160)], 100, 64, 1),      # packer round takes fewer than 32 blocks and a segment needs many rounds
    ([(333, 77)], 100, 65535, 1),    # one segment for the whole file
    ([(320, 240)], 90, 8, 3),        # 1-px checkerboard: one extreme coefficient behind a long zero run (ZRL)
    ([(64, 64)], 100, 1, 2),         # noise with translucent alpha over white, a segment per MCU
])
def test_gpu_worst_case_blocks(ctx, dims, q, ri, pattern):
    """Blocks near the 1665-bit worst case: the block coder's word drain and the packer's window rounds."""
    got = _gpu_case(ctx, dims, "vertical", 0, q, ri, pattern=pattern)
    assert (got.width, got.height) == dims[0]
    want = J.encode_i420(got.y, got.cb, got.cr, q, ri)
    assert got.jpeg == want
    size_, ycc = _decode_ycc(got.jpeg)
    assert size_ == (got.width, got.height)
    assert _psnr(ycc[:, :, 0], got.y) >= 36.0
    if pattern == 1 and q == 100:
        assert len(got.jpeg) > got.width * got.height   # > 8 bits per pixel: these blocks really are long


@pytest.mark.gpu
def test_gpu_jpeg_only_and_defaults(ctx):
    """JPEG alone (planes stay inside the library), default quality 100 / restart interval 8."""
    dims = [(403, 302)] * 2
    both = _gpu_case(ctx, dims, "vertical", 300, 0, 0)
    only = _gpu_case(ctx, dims, "vertical", 300, 0, 0, output="jpeg")
    assert only.rgba is None and only.y is None and only.jpeg == both.jpeg
    assert only.kernel_launches == 6
    withrgba = _gpu_case(ctx, dims, "vertical", 300, 0, 0, output="rgba+jpeg")
    assert withrgba.jpeg == both.jpeg and (withrgba.rgba == both.rgba).all()


@pytest.mark.gpu
def test_gpu_jpeg_capacity_error(ctx):
    from imagestitching_b200 import StitchError
    dims = [(403, 302)] * 2
    bitmaps = [U.synth_image(w, h, 321 + i, 1) for i, (w, h) in enumerate(dims)]
    with pytest.raises(StitchError) as e:
        ctx.run(U.page_images(dims, bitmaps, None, [BIG, 0]), "vertical", platform="ios", deviceMaxCanvasSize=300,
                output="jpeg", jpeg_capacity=4096)
    assert "bytes needed" in str(e.value)
