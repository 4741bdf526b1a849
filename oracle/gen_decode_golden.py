"""Regenerates tests/golden/decode_golden.json: SHA-256 of the bitmaps the oracle decoders produce for files that are built
here from seeded data by deterministic writers (no PIL, nothing whose bytes depend on a library version that the decoded
pixels would depend on): progressive JPEG files under four scan scripts, Adam7 PNG files, BMP files of every depth, GIF
files written by the small LZW encoder below.

    python -m oracle.gen_decode_golden

The reference decodes inside the closed WeChat runtime (utils/canvas.js:27-121) and holds no vectors; the decoders are
pinned to libjpeg-turbo / libpng / PIL in tests/test_jpeg_progressive.py, test_png_decode.py, test_bmp_decode.py and
test_gif_decode.py.  This fixture freezes what they return, so a change to an oracle decoder -- or, through
tests/test_decode_golden.py, to a GPU decoder -- shows up as a diff of this file.
"""
import hashlib
import json
import os
import struct

import numpy as np

from . import bmp_ref, gif_ref, jpeg_c, png_dec_c
from .gen_export_golden import planes


def gif_file(idx: np.ndarray, palette: np.ndarray, interlaced: bool, transparent: int = -1) -> bytes:
    """A GIF89a file of one frame: global colour table, optional graphic control extension, LZW with a growing table."""
    h, w = idx.shape
    nbits = max(2, int(np.ceil(np.log2(max(2, len(palette))))))
    table = np.zeros((1 << nbits, 3), np.uint8)
    table[: len(palette)] = palette
    out = bytearray(b"GIF89a" + struct.pack("<HHBBB", w, h, 0x80 | (nbits - 1), 0, 0) + table.tobytes())
    if transparent >= 0:
        out += b"\x21\xf9\x04\x01\x00\x00" + bytes([transparent]) + b"\x00"
    out += b"\x2c" + struct.pack("<HHHHB", 0, 0, w, h, 0x40 if interlaced else 0)
    rows = idx
    if interlaced:
        order = list(range(0, h, 8)) + list(range(4, h, 8)) + list(range(2, h, 4)) + list(range(1, h, 2))
        rows = idx[order]
    data = rows.reshape(-1).tolist()
    clear, eoi = 1 << nbits, (1 << nbits) + 1
    codes = [clear]
    table_d = {(i,): i for i in range(clear)}
    nxt, cur = clear + 2, ()
    sizes = []
    size = nbits + 1
    emitted = []                                   # (code, size)
    emitted.append((clear, size))
    for px in data:
        cand = cur + (px,)
        if cand in table_d:
            cur = cand
            continue
        emitted.append((table_d[cur], size))
        if nxt < 4096:
            table_d[cand] = nxt
            nxt += 1
            if nxt > (1 << size) and size < 12:
                size += 1
        else:                                       # table full: start over
            emitted.append((clear, size))
            table_d = {(i,): i for i in range(clear)}
            nxt, size = clear + 2, nbits + 1
        cur = (px,)
    if cur:
        emitted.append((table_d[cur], size))
    emitted.append((eoi, size))
    acc = cnt = 0
    stream = bytearray()
    for code, sz in emitted:
        acc |= code << cnt
        cnt += sz
        while cnt >= 8:
            stream.append(acc & 255)
            acc >>= 8
            cnt -= 8
    if cnt:
        stream.append(acc & 255)
    out.append(nbits)
    for o in range(0, len(stream), 255):
        out.append(min(255, len(stream) - o))
        out += stream[o:o + 255]
    out += b"\x00\x3b"
    del codes, sizes
    return bytes(out)


def files():
    """(name, kind, file bytes); kind names the oracle decoder."""
    from tests import jprog_encode as JE
    from tests.test_bmp_decode import _raw_bmp
    from tests.test_png_decode import _adam7_png
    out = []
    for (w, h, q, seed) in ((64, 48, 90, 1), (333, 200, 75, 2), (17, 9, 100, 3)):
        y, cb, cr = planes(w, h, seed)
        base = jpeg_c.encode_i420(y, cb, cr, q, 8)
        for script in sorted(JE.SCRIPTS):
            for ri in (0, 5):
                out.append((f"jpeg_prog_{w}x{h}_q{q}_{script}_ri{ri}", "jpeg", JE.encode_progressive(base, JE.SCRIPTS[script](3), ri)))
    rng = np.random.default_rng(11)
    for (w, h) in ((67, 41), (5, 3), (8, 8)):
        for depth, ctype in ((1, 0), (4, 3), (8, 2), (8, 6), (16, 6)):
            ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
            rb = (w * ch * depth + 7) // 8
            rows = [rng.integers(0, 256, rb, dtype=np.uint8) for _ in range(h)]
            extra = []
            if ctype == 3:
                extra = [(b"PLTE", rng.integers(0, 256, 16 * 3, dtype=np.uint8).tobytes()), (b"tRNS", rng.integers(0, 256, 8, dtype=np.uint8).tobytes())]
            out.append((f"png_adam7_{w}x{h}_ct{ctype}_d{depth}", "png", _adam7_png(w, h, depth, ctype, rows, extra)))
    w, h = 37, 19
    for bpp in (1, 4, 8):
        pal = [tuple(int(v) for v in rng.integers(0, 256, 3)) for _ in range(1 << min(bpp, 7))]
        rows = [rng.integers(0, len(pal) if bpp == 8 else 256, (w * bpp + 7) // 8, dtype=np.uint8) for _ in range(h)]
        out.append((f"bmp_pal{bpp}", "bmp", _raw_bmp(w, h, bpp, rows, palette=pal)))
    rows16 = [rng.integers(0, 256, w * 2, dtype=np.uint8) for _ in range(h)]
    rows32 = [rng.integers(0, 256, w * 4, dtype=np.uint8) for _ in range(h)]
    out.append(("bmp_555", "bmp", _raw_bmp(w, h, 16, rows16)))
    out.append(("bmp_565", "bmp", _raw_bmp(w, h, 16, rows16, compression=3, masks=(0xF800, 0x7E0, 0x1F))))
    out.append(("bmp_24_topdown", "bmp", _raw_bmp(w, h, 24, [rng.integers(0, 256, w * 3, dtype=np.uint8) for _ in range(h)], top_down=True)))
    out.append(("bmp_32_plain", "bmp", _raw_bmp(w, h, 32, rows32)))
    out.append(("bmp_32_v4_bgra", "bmp", _raw_bmp(w, h, 32, rows32, compression=3, masks=(0xFF0000, 0xFF00, 0xFF, 0xFF000000), header=108)))
    for (w, h, colours, laced, tr) in ((1, 1, 2, False, -1), (64, 48, 16, True, 3), (123, 77, 200, False, -1), (250, 131, 256, True, 0), (400, 300, 5, False, -1)):
        yy, xx = np.mgrid[0:h, 0:w]
        idx = (((xx // 7 + yy // 5) * 3 + rng.integers(0, 2, (h, w))) % colours).astype(np.uint8)
        if colours == 5:
            idx[:] = 0
            idx[40:200, 60:300] = 1                 # flat areas: long strings, the table fills and clears
            idx[::23] = 2
        pal = rng.integers(0, 256, (colours, 3), dtype=np.uint8)
        out.append((f"gif_{w}x{h}_{colours}{'_laced' if laced else ''}{'_tr' if tr >= 0 else ''}", "gif", gif_file(idx, pal, laced, tr)))
    return out


DECODERS = {"jpeg": jpeg_c.decode, "png": png_dec_c.decode, "bmp": bmp_ref.decode, "gif": gif_ref.decode}


def build():
    out = []
    for name, kind, f in files():
        px = DECODERS[kind](f)
        out.append({"name": name, "kind": kind, "width": int(px.shape[1]), "height": int(px.shape[0]),
                    "decoded_sha256": hashlib.sha256(np.ascontiguousarray(px).tobytes()).hexdigest()})
    return out


if __name__ == "__main__":
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "decode_golden.json")
    with open(path, "w") as fh:
        json.dump(build(), fh, indent=1)
        fh.write("\n")
    print("wrote", path)
