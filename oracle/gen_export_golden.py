"""Regenerates tests/golden/export_golden.json: SHA-256 of the files the sequential CPU writers (oracle/jpeg_ref.c,
oracle/png_ref.c) produce for seeded inputs, and of the bitmaps oracle/jpeg_dec_ref.c decodes from those JPEG files.

    python -m oracle.gen_export_golden

The reference exports through the closed WeChat runtime (utils/canvas.js:205-242), so it holds no file vectors; the
writers are pinned by decoding their files with libjpeg / libpng (tests/test_jpeg.py, tests/test_png.py).  This
fixture freezes their exact bytes, so a change to an encoder choice (tables, band structure, run tokens) in the oracle
or -- through the byte-identity tests -- in the GPU encoders shows up as a diff of this file, not silently.
"""
import hashlib
import json
import os

import numpy as np

from . import jpeg_c, png_c


def planes(w, h, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    y = (128 + 90 * np.sin(xx / 9.0 + seed) * np.cos(yy / 7.0) + rng.integers(-20, 21, (h, w))).clip(0, 255).astype(np.uint8)
    cw, ch = (w + 1) // 2, (h + 1) // 2
    cb = rng.integers(0, 256, (ch, cw), dtype=np.uint8)
    cr = ((np.arange(cw)[None, :] * 3 + np.arange(ch)[:, None] * 5 + seed) % 256).astype(np.uint8)
    return y, cb, cr


def rgba(w, h, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    a[: h // 2, :, :3] = (np.arange(w)[None, :, None] // 16 * 16) % 256   # flat steps: run tokens
    a[h // 3: h // 2, : w // 2] = 255                                      # a white block: long runs
    a[:, :, 3] = 255
    return a


JPEG_CASES = [(64, 48, 90, 8, 1), (333, 200, 75, 3, 2), (17, 9, 100, 1, 3), (128, 128, 50, 65535, 4)]
PNG_CASES = [(64, 48, 1), (333, 77, 2), (1, 1, 3), (700, 19, 4)]


def build():
    out = {"jpeg": [], "png": []}
    for (w, h, q, ri, seed) in JPEG_CASES:
        y, cb, cr = planes(w, h, seed)
        f = jpeg_c.encode_i420(y, cb, cr, q, ri)
        dec = jpeg_c.decode(f)
        out["jpeg"].append({"w": w, "h": h, "quality": q, "restart_interval": ri, "seed": seed, "bytes": len(f),
                            "sha256": hashlib.sha256(f).hexdigest(),
                            "decoded_sha256": hashlib.sha256(np.ascontiguousarray(dec).tobytes()).hexdigest()})
    for (w, h, seed) in PNG_CASES:
        f = png_c.encode_rgba(rgba(w, h, seed))
        out["png"].append({"w": w, "h": h, "seed": seed, "bytes": len(f), "sha256": hashlib.sha256(f).hexdigest()})
    return out


if __name__ == "__main__":
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "export_golden.json")
    with open(path, "w") as fh:
        json.dump(build(), fh, indent=1)
        fh.write("\n")
    print("wrote", path)
