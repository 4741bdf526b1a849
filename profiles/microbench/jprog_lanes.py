"""Progressive-JPEG jobs in flight: N submitted jobs of 9 files (1600 x 1200) on N lanes, wall clock per job.
usage: jprog_lanes.py [lanes ...]   (with a -DJPROG_TRACE build every refinement CTA prints its SM and its start / end)"""
import io
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from PIL import Image  # noqa: E402

from imagestitching_b200 import StitchContext  # noqa: E402

y, x = np.mgrid[0:1200, 0:1600]
files = []
for i in range(3):
    a = np.stack([128 + 100 * np.sin(x / 17 + y / 29), 128 + 90 * np.cos(x / 23 - y / 13), (x * 3 + y * 5) % 256], -1)
    a = np.clip(a + np.random.default_rng(i).integers(-4, 5, a.shape), 0, 255).astype(np.uint8)
    b = io.BytesIO()
    Image.fromarray(a).save(b, "JPEG", quality=90, subsampling=2, progressive=True)
    files.append(b.getvalue())
imgs = [{"file": files[i % 3]} for i in range(9)]
ctx = StitchContext(0)
for lanes in [int(v) for v in sys.argv[1:]] or [1, 2, 4, 8]:
    ctx.set_lanes(lanes)
    for t in [ctx.submit(imgs, "vertical", platform="ios", output="jpeg") for _ in range(lanes)]:
        ctx.wait(t)
    t0 = time.perf_counter()
    n = 3 * lanes
    tickets = []
    for i in range(n):
        tickets.append(ctx.submit(imgs, "vertical", platform="ios", output="jpeg"))
        if i >= lanes - 1:
            ctx.wait(tickets[i - (lanes - 1)])
    for t in tickets[-(lanes - 1):] if lanes > 1 else []:
        ctx.wait(t)
    dt = time.perf_counter() - t0
    print(f"{lanes} lane(s): {dt / n * 1e3:.1f} ms/job ({n} jobs)")
ctx.close()
