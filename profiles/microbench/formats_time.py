"""GIF / BMP inputs: wall clock of stitch_decode_image for 1600 x 1200 files (includes the 7.7 MB bitmap download)."""
import io
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from PIL import Image  # noqa: E402

from imagestitching_b200 import StitchContext  # noqa: E402

y, x = np.mgrid[0:1200, 0:1600]
a = np.stack([128 + 100 * np.sin(x / 17 + y / 29), 128 + 90 * np.cos(x / 23 - y / 13), (x * 3 + y * 5) % 256], -1)
a = np.clip(a + np.random.default_rng(1).integers(-4, 5, a.shape), 0, 255).astype(np.uint8)
ctx = StitchContext(0)
for name, im, fmt in (("gif, photo, 256 colours", Image.fromarray(a).quantize(256), "GIF"),
                      ("gif, graphic, 16 colours", Image.fromarray((a // 64) * 64).quantize(16), "GIF"),
                      ("bmp, 24 bit", Image.fromarray(a), "BMP")):
    b = io.BytesIO()
    im.save(b, fmt)
    f = b.getvalue()
    ctx.decode_image(f)
    t0 = time.perf_counter()
    ctx.decode_image(f)
    dt = time.perf_counter() - t0
    print(f"{name}: 1600x1200, {len(f) / 1e6:.2f} MB file, decode_image {dt * 1e3:.1f} ms")
ctx.close()
