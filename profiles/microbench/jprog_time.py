"""Progressive-JPEG inputs: wall-clock of file jobs through the Python binding (9 files -> stitched JPEG file), beside
the same pictures as baseline files.  Usage: python profiles/microbench/jprog_time.py [width height]"""
import io
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from PIL import Image, ImageFile  # noqa: E402

from imagestitching_b200 import StitchContext  # noqa: E402

ImageFile.MAXBLOCK = 1 << 26
w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4032, 3024)


def picture(seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    a = np.stack([128 + 100 * np.sin(x / 17 + y / 29), 128 + 90 * np.cos(x / 23 - y / 13), (x * 3 + y * 5) % 256], -1)
    a = a + rng.integers(-12, 13, a.shape)
    return Image.fromarray(np.clip(a, 0, 255).astype(np.uint8))


def files(progressive):
    out = []
    for i in range(9):
        b = io.BytesIO()
        picture(i).save(b, "JPEG", quality=90, subsampling=2, progressive=progressive)
        out.append(b.getvalue())
    return out


ctx = StitchContext(0)
for name, prog in (("baseline", False), ("progressive", True)):
    fs = files(prog)
    for env in ((None,) if not prog else (None, "1")):
        if env:
            os.environ["STITCH_JPROG_SEQ"] = env
        imgs = [{"file": f} for f in fs]
        ctx.run(imgs, "vertical", platform="ios", output="jpeg")
        t0 = time.perf_counter()
        n = 3
        for _ in range(n):
            ctx.run(imgs, "vertical", platform="ios", output="jpeg")
        dt = (time.perf_counter() - t0) / n
        one = [{"file": fs[0]}, {"file": fs[1]}]
        ctx.run(one, "vertical", platform="ios", output="jpeg")
        t0 = time.perf_counter()
        for _ in range(n):
            ctx.run(one, "vertical", platform="ios", output="jpeg")
        dt2 = (time.perf_counter() - t0) / n
        print(f"{name}{' (one thread per scan)' if env else ''}: 9 x {w}x{h} files ({sum(map(len, fs)) / 1e6:.1f} MB) -> JPEG: "
              f"{dt * 1e3:.1f} ms/job; 2 files: {dt2 * 1e3:.1f} ms/job")
    os.environ.pop("STITCH_JPROG_SEQ", None)
ctx.close()
