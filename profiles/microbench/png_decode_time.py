"""Timing of PNG-input jobs (png_decode.cu): 9 files -> stitched RGBA canvas on the device / PNG file out.
usage: python profiles/microbench/png_decode_time.py [photo|shot] [jobs]"""
import io
import sys
import time

import numpy as np
from PIL import Image

sys.path.insert(0, ".")
from imagestitching_b200 import StitchContext  # noqa: E402
from tests.test_png_decode import _picture, _screenshot  # noqa: E402


def files(kind):
    out = []
    for i in range(3):   # three distinct files, used three times each
        b = io.BytesIO()
        if kind == "photo":
            Image.fromarray(_picture(4032, 3024, 3, seed=i), "RGB").save(b, "PNG", compress_level=6)
        else:
            Image.fromarray(_screenshot(1290, 2796, seed=i), "RGBA").save(b, "PNG")
        out.append(b.getvalue())
    return [out[i % 3] for i in range(9)]


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "photo"
    jobs = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    fs = files(kind)
    print(kind, "file bytes", [len(f) for f in fs[:3]], flush=True)
    ctx = StitchContext(0)
    t = time.perf_counter(); ctx.decode_png(fs[0]); print("first decode_png (allocations): %.1f ms" % ((time.perf_counter() - t) * 1e3))
    for _ in range(3):
        t = time.perf_counter(); ctx.decode_png(fs[0]); print("decode_png one file: %.2f ms" % ((time.perf_counter() - t) * 1e3))
    if jobs == 0:
        ctx.close()
        return
    imgs = [{"file": f} for f in fs]
    for output in ("rgba", "png"):
        ctx.set_lanes(1)
        ctx.run(imgs, "vertical", output=output)
        t = time.perf_counter(); ctx.run(imgs, "vertical", output=output); one = (time.perf_counter() - t) * 1e3
        ctx.set_lanes(3)
        ts = [ctx.submit(imgs, "vertical", output=output) for _ in range(3)]
        [ctx.wait(x) for x in ts]
        t = time.perf_counter()
        ts = [ctx.submit(imgs, "vertical", output=output) for _ in range(jobs)]
        [ctx.wait(x) for x in ts]
        per = (time.perf_counter() - t) * 1e3 / jobs
        print("9 %s PNG files -> %s: single job %.2f ms, %.2f ms/job over 3 lanes" % (kind, output, one, per), flush=True)
    ctx.close()


main()
