"""File-to-file timing of the widened path: 9 JPEG files in -> GPU decode -> stitch -> RGBA canvas or JPEG file out.

Not the contract bench (bench.py measures BASELINE's metric on RGBA bitmaps); this shows what the (f) rows buy
end to end: the upload shrinks from 439 MB of bitmaps to the files themselves.  Run on the GPU box:
    python profiles/bench_files.py [--steps 10]
The CPU column is libjpeg-turbo through PIL on one host core (decode only), for scale.
"""
import argparse
import io
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from PIL import Image, ImageFile  # noqa: E402

ImageFile.MAXBLOCK = 1 << 26


def picture(w, h, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    a = np.stack([128 + 100 * np.sin(x / 170 + y / 290 + seed), 128 + 90 * np.cos(x / 230 - y / 130), (x + y * 2) % 256], -1)
    a = a + rng.integers(-12, 13, a.shape)
    return Image.fromarray(np.clip(a, 0, 255).astype(np.uint8))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--quality", type=int, default=90)
    args = ap.parse_args()
    from imagestitching_b200 import StitchContext, prepare_image
    files = []
    for i in range(9):
        b = io.BytesIO()
        picture(4032, 3024, i).save(b, "JPEG", quality=args.quality, subsampling=2)
        files.append(b.getvalue())
    recs = [dict(prepare_image(f), file=f) for f in files]
    t = time.perf_counter()
    for f in files[:3]:
        np.asarray(Image.open(io.BytesIO(f)).convert("RGB"))
    cpu_ms = (time.perf_counter() - t) / 3 * 1e3
    ctx = StitchContext(0)
    out = {"files_MB": round(sum(map(len, files)) / 1e6, 2), "libjpeg_turbo_decode_ms_per_image_1core": round(cpu_ms, 1)}
    # decode only
    ctx.decode_jpeg(files[0])
    t = time.perf_counter()
    for k in range(args.steps):
        ctx.decode_jpeg(files[k % 9])
    out["gpu_decode_ms_per_image_incl_d2h"] = round((time.perf_counter() - t) / args.steps * 1e3, 2)
    for output in ("rgba", "jpeg", "png"):
        kw = dict(jpeg_quality=90, jpeg_capacity=64 << 20) if output == "jpeg" else {}
        ctx.run(recs, "vertical", mode="min", platform="ios", output=output, **kw)
        l0 = ctx.kernel_launches
        t = time.perf_counter()
        for _ in range(args.steps):
            r = ctx.run(recs, "vertical", mode="min", platform="ios", output=output, **kw)
        if output == "png":
            out["png_MB"] = round(len(r.png) / 1e6, 2)
        ms = (time.perf_counter() - t) / args.steps * 1e3
        out[f"files_to_{output}_ms_per_job"] = round(ms, 2)
        out[f"files_to_{output}_MPps"] = round(r.width * r.height / 1e6 / (ms / 1e3), 1)
        out[f"files_to_{output}_launches_per_job"] = (ctx.kernel_launches - l0) / args.steps
    ctx.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
