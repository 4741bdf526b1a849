"""BMP inputs (P:4 SUPPORTED_IMAGE_TYPES 'bmp'; utils/canvas.js:27-121 loadImageFrom).

CPU: the oracle (oracle/bmp_ref.py) is pinned to PIL's BmpImagePlugin on PIL-written and hand-made files.
GPU: stitch_decode_bmp / stitch_decode_image and BMP job inputs (STITCH_MEM_HOST_BMP, STITCH_MEM_HOST_FILE) equal the oracle.
"""
import io
import struct

import numpy as np
import pytest
from PIL import Image

from imagestitching_b200 import _lib as L
from oracle import bmp_ref as B


def _picture(w, h, c, seed=0):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    base = (np.sin(x / 9.0) + np.cos(y / 13.0)) * 60 + 128
    return np.stack([(base + 25 * k + rng.normal(0, 8, (h, w))).clip(0, 255) for k in range(c)], -1).astype(np.uint8)


def _pil_bmp(arr, mode):
    b = io.BytesIO()
    Image.fromarray(arr, mode).save(b, "BMP")
    return b.getvalue()


def _pil_rgba(data):
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGBA"))


def _raw_bmp(w, h, bpp, rows, palette=None, compression=0, masks=(), top_down=False, header=40, core=False):
    """A BMP file from packed pixel rows (top row first); rows are padded to 4 bytes here."""
    stride = (w * bpp + 31) // 32 * 4
    body = b"".join(bytes(r).ljust(stride, b"\0") for r in (rows if top_down else rows[::-1]))
    pal = b""
    if palette is not None:
        pal = b"".join(bytes([b_, g, r]) + (b"" if core else b"\0") for r, g, b_ in palette)
    if core:
        dib = struct.pack("<IHHHH", 12, w, h, 1, bpp)
    else:
        dib = struct.pack("<IiiHHIIiiII", header, w, -h if top_down else h, 1, bpp, compression, len(body), 2835, 2835,
                          len(palette) if palette is not None else 0, 0)
        m = b"".join(struct.pack("<I", v) for v in masks)
        if header > 40:
            dib += m.ljust(header - 40, b"\0")
            m = b""
        dib += m
    off = 14 + len(dib) + len(pal)
    return b"BM" + struct.pack("<IHHI", off + len(body), 0, 0, off) + dib + pal + body


def files():
    rng = np.random.default_rng(3)
    out = [("pil_rgb", _pil_bmp(_picture(131, 77, 3), "RGB")), ("pil_gray", _pil_bmp(_picture(130, 51, 1)[..., 0], "L")),
           ("pil_rgba", _pil_bmp(_picture(64, 33, 4, seed=2), "RGBA")), ("pil_1x1", _pil_bmp(np.array([[[1, 2, 3]]], np.uint8), "RGB"))]
    b = io.BytesIO()
    Image.fromarray(_picture(97, 40, 3, seed=4), "RGB").quantize(50).save(b, "BMP")
    out.append(("pil_palette", b.getvalue()))
    b = io.BytesIO()
    Image.fromarray((_picture(70, 30, 1)[..., 0] > 128)).save(b, "BMP")
    out.append(("pil_1bit", b.getvalue()))
    w, h = 37, 19
    for bpp in (1, 4, 8):
        n = 1 << bpp
        pal = [tuple(int(v) for v in rng.integers(0, 256, 3)) for _ in range(min(n, 200))]
        rb = (w * bpp + 7) // 8
        rows = [rng.integers(0, 256, rb, dtype=np.uint8) for _ in range(h)]
        if bpp == 8:
            rows = [r % len(pal) for r in rows]
        out.append((f"pal{bpp}", _raw_bmp(w, h, bpp, rows, palette=pal)))
        out.append((f"pal{bpp}_topdown", _raw_bmp(w, h, bpp, rows, palette=pal, top_down=True)))
        out.append((f"pal{bpp}_os2", _raw_bmp(w, h, bpp, rows, palette=pal[: n], core=True) if len(pal) == n else out[-1][1]))
    rows16 = [rng.integers(0, 256, w * 2, dtype=np.uint8) for _ in range(h)]
    out.append(("rgb555", _raw_bmp(w, h, 16, rows16)))
    out.append(("rgb565", _raw_bmp(w, h, 16, rows16, compression=3, masks=(0xF800, 0x7E0, 0x1F))))
    out.append(("rgb555_fields", _raw_bmp(w, h, 16, rows16, compression=3, masks=(0x7C00, 0x3E0, 0x1F))))
    rows24 = [rng.integers(0, 256, w * 3, dtype=np.uint8) for _ in range(h)]
    out.append(("rgb24_topdown", _raw_bmp(w, h, 24, rows24, top_down=True)))
    rows32 = [rng.integers(0, 256, w * 4, dtype=np.uint8) for _ in range(h)]
    out.append(("rgb32_plain", _raw_bmp(w, h, 32, rows32)))                                   # opaque: the fourth byte is unused
    out.append(("rgb32_fields_bgrx", _raw_bmp(w, h, 32, rows32, compression=3, masks=(0xFF0000, 0xFF00, 0xFF))))
    out.append(("rgb32_v4_bgra", _raw_bmp(w, h, 32, rows32, compression=3, masks=(0xFF0000, 0xFF00, 0xFF, 0xFF000000), header=108)))
    out.append(("rgb32_v4_rgba", _raw_bmp(w, h, 32, rows32, compression=3, masks=(0xFF, 0xFF00, 0xFF0000, 0xFF000000), header=108)))
    out.append(("rgb32_v5_xbgr", _raw_bmp(w, h, 32, rows32, compression=3, masks=(0xFF000000, 0xFF0000, 0xFF00, 0), header=124)))
    return out


def test_oracle_equals_pil():
    for name, data in files():
        assert np.array_equal(B.decode(data), _pil_rgba(data)), name


def test_oracle_rejects():
    good = _pil_bmp(_picture(20, 10, 3), "RGB")
    for bad in (b"GIF89a" + bytes(40), good[:40], good[:30] + struct.pack("<I", 1) + good[34:]):   # not BMP, cut, RLE8
        with pytest.raises(B.BmpError):
            B.decode(bad)


@pytest.fixture()
def ctx():
    from imagestitching_b200 import StitchContext
    c = StitchContext(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_gpu_bmp_equals_oracle(ctx):
    for name, data in files():
        want = B.decode(data)
        assert np.array_equal(ctx.decode_bmp(data), want), name
        assert np.array_equal(ctx.decode_image(data), want), name
    big = _pil_bmp(_picture(2000, 1500, 3, seed=9), "RGB")
    assert np.array_equal(ctx.decode_bmp(big), B.decode(big))


@pytest.mark.gpu
def test_gpu_decode_image_dispatches_and_refuses(ctx):
    from imagestitching_b200 import StitchError
    from oracle import jpeg_c as J
    from oracle import png_dec_c as P
    a = _picture(120, 80, 3, seed=5)
    jb, pb = io.BytesIO(), io.BytesIO()
    Image.fromarray(a).save(jb, "JPEG", quality=90)
    Image.fromarray(a).save(pb, "PNG")
    assert np.array_equal(ctx.decode_image(jb.getvalue()), J.decode(jb.getvalue()))
    assert np.array_equal(ctx.decode_image(pb.getvalue()), P.decode(pb.getvalue()))
    webp = b"RIFF" + struct.pack("<I", 30) + b"WEBPVP8 " + struct.pack("<I", 18) + bytes(18)     # probed, never decoded
    good = _pil_bmp(a, "RGB")
    rle = good[:30] + struct.pack("<I", 1) + good[34:]
    for bad in (webp, rle, good[:60]):
        with pytest.raises(StitchError) as e:
            ctx.decode_image(bad)
        assert e.value.status == L.E_BAD_IMAGE


@pytest.mark.gpu
def test_gpu_job_with_bmp_jpeg_and_png_files(ctx):
    """Files of three types in one job (the library looks at their signatures) == the job over the decoded bitmaps;
    a WebP file in the list is the page's decode error."""
    from imagestitching_b200 import StitchError
    from oracle import jpeg_c as J
    from oracle import png_dec_c as P
    from tests import util as U
    dims = [(131, 77), (120, 80), (97, 40)]
    bmp = _pil_bmp(_picture(131, 77, 3), "RGB")
    jb, pb = io.BytesIO(), io.BytesIO()
    Image.fromarray(_picture(120, 80, 3, seed=5)).save(jb, "JPEG", quality=90)
    Image.fromarray(_picture(97, 40, 4, seed=6), "RGBA").save(pb, "PNG")
    files_ = [bmp, jb.getvalue(), pb.getvalue()]
    bitmaps = [B.decode(bmp), J.decode(jb.getvalue()), P.decode(pb.getvalue())]
    want = ctx.run(U.page_images(dims, bitmaps), "vertical", platform="ios")
    got = ctx.run([{"file": f} for f in files_], "vertical", platform="ios")
    assert np.array_equal(got.rgba, want.rgba)
    ctx.set_lanes(2)
    tickets = [ctx.submit([{"file": f} for f in files_], "vertical", platform="ios") for _ in range(3)]
    for t in tickets:
        assert np.array_equal(ctx.wait(t).rgba, want.rgba)
    webp = b"RIFF" + struct.pack("<I", 30) + b"WEBPVP8 " + struct.pack("<I", 18) + bytes(18)
    with pytest.raises(StitchError) as e:
        ctx.run([{"file": bmp}, {"file": webp, "naturalWidth": 50, "naturalHeight": 40, "width": 50, "height": 40}], "vertical", platform="ios")
    assert e.value.status == L.E_BAD_IMAGE and "图片1解码异常" in e.value.message
