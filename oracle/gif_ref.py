"""TEST INFRASTRUCTURE (oracle): a plain-Python restatement of GIF decoding (first frame), the checker of
imagestitching_b200/csrc/gif_decode.cuh.

The reference decodes its inputs inside the closed WeChat runtime (utils/canvas.js:27-121 loadImageFrom; 'gif' is in
SUPPORTED_IMAGE_TYPES, pages/index/index.js:4) and paints the first frame; the stated pin is PIL's GifImagePlugin, frame
0, convert("RGBA") (tests/test_gif_decode.py).  GIF89a: logical screen, global / local colour table, graphic control
extension (transparent index), variable-width LZW codes packed LSB first, 4-pass row interlace.  Only tests/, smoke()
and bench.py's CPU legs may import this.
"""
from __future__ import annotations

import numpy as np


class GifError(ValueError):
    pass


def lzw(stream: bytes, min_code: int, n_out: int) -> np.ndarray:
    """The index stream of a GIF image: classic table-of-strings LZW (the GPU keeps (start, length) pairs instead)."""
    clear, eoi = 1 << min_code, (1 << min_code) + 1
    out = bytearray()
    table = None
    size = min_code + 1
    acc = cnt = 0
    prev = None
    it = iter(stream)
    while len(out) < n_out:
        while cnt < size:
            try:
                acc |= next(it) << cnt
            except StopIteration:
                return np.frombuffer(bytes(out).ljust(n_out, b"\0"), np.uint8)
            cnt += 8
        code = acc & ((1 << size) - 1)
        acc >>= size
        cnt -= size
        if code == clear:
            table = [bytes([i]) for i in range(clear)] + [b"", b""]
            size = min_code + 1
            prev = None
            continue
        if code == eoi:
            break
        if table is None:
            table = [bytes([i]) for i in range(clear)] + [b"", b""]
        if code < len(table) and code not in (clear, eoi):
            s = table[code]
        elif code == len(table) and prev is not None:
            s = prev + prev[:1]
        else:
            raise GifError("code outside the dictionary")
        out += s
        if prev is not None and len(table) < 4096:
            table.append(prev + s[:1])
            if len(table) == (1 << size) and size < 12:
                size += 1
        prev = s
    return np.frombuffer(bytes(out[:n_out]).ljust(n_out, b"\0"), np.uint8)


def decode(data: bytes, lzw_fn=None) -> np.ndarray:
    """RGBA8 [h, w, 4] of the first frame.  lzw_fn: another LZW decoder to put in (tests restate the kernel's form)."""
    if len(data) < 13 or data[:6] not in (b"GIF87a", b"GIF89a"):
        raise GifError("not a GIF file")
    sw, sh = data[6] | data[7] << 8, data[8] | data[9] << 8
    o = 13
    table = b""
    if data[10] & 0x80:
        n = 2 << (data[10] & 7)
        table = data[o:o + 3 * n]
        o += 3 * n
    transparent = -1
    while o < len(data):
        b = data[o]
        o += 1
        if b == 0x21:
            label = data[o]
            o += 1
            first = True
            while True:
                n = data[o]
                o += 1
                if not n:
                    break
                if label == 0xF9 and first and n >= 4:
                    transparent = data[o + 3] if data[o] & 1 else -1
                first = False
                o += n
            continue
        if b != 0x2C:
            raise GifError("no image")
        left, top = data[o] | data[o + 1] << 8, data[o + 2] | data[o + 3] << 8
        w, h = data[o + 4] | data[o + 5] << 8, data[o + 6] | data[o + 7] << 8
        packed = data[o + 8]
        o += 9
        if (left, top, w, h) != (0, 0, sw, sh):
            raise GifError("frame smaller than the screen")
        if packed & 0x80:
            n = 2 << (packed & 7)
            table = data[o:o + 3 * n]
            o += 3 * n
        min_code = data[o]
        o += 1
        stream = bytearray()
        while o < len(data):
            n = data[o]
            o += 1
            if not n:
                break
            stream += data[o:o + n]
            o += n
        idx = (lzw_fn or lzw)(bytes(stream), min_code, w * h).reshape(h, w)
        if packed & 0x40:
            order = list(range(0, h, 8)) + list(range(4, h, 8)) + list(range(2, h, 4)) + list(range(1, h, 2))
            full = np.empty_like(idx)
            full[order] = idx
            idx = full
        pal = np.zeros((256, 4), np.uint8)
        pal[:, 3] = 255
        t = np.frombuffer(table, np.uint8).reshape(-1, 3)
        pal[: len(t), :3] = t
        if transparent >= 0:
            pal[transparent, 3] = 0
        return pal[idx]
    raise GifError("no image")
