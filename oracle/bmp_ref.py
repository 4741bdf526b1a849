"""TEST INFRASTRUCTURE (oracle): a numpy restatement of BMP decoding, the checker of imagestitching_b200/csrc/bmp_decode.cuh.

The reference decodes its inputs inside the closed WeChat runtime (utils/canvas.js:27-121 loadImageFrom; 'bmp' is in
SUPPORTED_IMAGE_TYPES, pages/index/index.js:4); the stated pin is PIL's BmpImagePlugin (tests/test_bmp_decode.py):
rows bottom-up unless the height is negative, 1 / 4 / 8 bit indices through the BGR(X) palette, 16 bit 5-5-5 unless bit
fields say otherwise, 24 bit B G R, 32 bit B G R X -- opaque unless a V3+ header names an alpha mask --, n-bit fields
widened by v * 255 // (2^n - 1).  Only tests/, smoke() and bench.py's CPU legs may import this.
"""
from __future__ import annotations

import struct

import numpy as np


class BmpError(ValueError):
    pass


def _field(mask: int):
    if not mask:
        return 0, 0
    shift = (mask & -mask).bit_length() - 1
    bits = (mask >> shift).bit_length()
    if (mask >> shift) != (1 << bits) - 1:
        raise BmpError("bit field is not contiguous")
    return shift, bits


def decode(data: bytes) -> np.ndarray:
    """RGBA8 [h, w, 4]."""
    if len(data) < 26 or data[:2] != b"BM":
        raise BmpError("not a BMP file")
    off_bits, hs = struct.unpack_from("<II", data, 10)
    if len(data) < 14 + hs:
        raise BmpError("header cut short")
    compression = colours = 0
    pal_entry = 4
    top_down = False
    if hs == 12:
        w, h, _, bpp = struct.unpack_from("<HHHH", data, 18)
        pal_entry = 3
    elif hs in (40, 52, 56, 64, 108, 124):
        w, h, _, bpp, compression = struct.unpack_from("<iiHHI", data, 18)
        colours = struct.unpack_from("<I", data, 46)[0]
        if h < 0:
            top_down, h = True, -h
    else:
        raise BmpError("unsupported header")
    if w <= 0 or h <= 0 or bpp not in (1, 4, 8, 16, 24, 32) or compression not in (0, 3):
        raise BmpError("unsupported variant")
    masks = [0, 0, 0, 0]
    if compression == 3:
        if bpp not in (16, 32):
            raise BmpError("bit fields need 16 or 32 bit pixels")
        nm = 4 if hs >= 56 else 3
        masks[:nm] = struct.unpack_from("<" + "I" * nm, data, 54)
        if bpp == 32 and not any(masks):
            masks = [0xFF0000, 0xFF00, 0xFF, 0xFF000000]
    elif bpp == 16:
        masks = [0x7C00, 0x3E0, 0x1F, 0]
    elif bpp == 32:
        masks = [0xFF0000, 0xFF00, 0xFF, 0]
    stride = (w * bpp + 31) // 32 * 4
    if off_bits + stride * h > len(data):
        raise BmpError("pixel array cut short")
    rows = np.frombuffer(data, np.uint8, stride * h, off_bits).reshape(h, stride)
    if not top_down:
        rows = rows[::-1]
    out = np.empty((h, w, 4), np.uint8)
    out[..., 3] = 255
    if bpp <= 8:
        if not colours or colours > (1 << bpp):
            colours = 1 << bpp
        pal = np.zeros((256, 4), np.uint8)
        pal[:, 3] = 255
        raw = np.frombuffer(data, np.uint8, colours * pal_entry, 14 + hs).reshape(colours, pal_entry)
        pal[:colours, 0], pal[:colours, 1], pal[:colours, 2] = raw[:, 2], raw[:, 1], raw[:, 0]
        bits = np.unpackbits(rows, axis=1)[:, : w * bpp].reshape(h, w, bpp)
        idx = (bits * (1 << np.arange(bpp - 1, -1, -1))).sum(2)
        out[:] = pal[idx]
    elif bpp == 24:
        px = rows[:, : w * 3].reshape(h, w, 3)
        out[..., 0], out[..., 1], out[..., 2] = px[..., 2], px[..., 1], px[..., 0]
    else:
        nb = bpp // 8
        b = rows[:, : w * nb].reshape(h, w, nb).astype(np.uint32)
        px = b[..., 0] | b[..., 1] << 8
        if nb == 4:
            px |= b[..., 2] << 16 | b[..., 3] << 24
        for k in range(4):
            shift, nbits = _field(masks[k])
            if not nbits:
                out[..., k] = 255 if k == 3 else 0
                continue
            v = (px >> shift) & ((1 << nbits) - 1)
            out[..., k] = v if nbits == 8 else (v >> (nbits - 8) if nbits > 8 else v * 255 // ((1 << nbits) - 1))
    return out
