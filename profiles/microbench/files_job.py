"""One JPEG-files-in / JPEG-file-out job of bench.py's e2e_files leg (9 x 12 MP JPEG files), for a launch list."""
import sys
import time

sys.path.insert(0, ".")
import numpy as np  # noqa: E402

from imagestitching_b200 import StitchContext  # noqa: E402
from tests import util as U  # noqa: E402

ctx = StitchContext(0)
files = []
for i in range(3):
    bmp = U.synth_image(4032, 3024, 0x5717C4 + i, 0)
    o = ctx.run([{"naturalWidth": 4032, "naturalHeight": 3024, "bitmap": bmp}], "vertical", platform="other", deviceMaxCanvasSize=4032,
                deviceMaxCanvasPixels=4032 * 3024, output="jpeg", jpeg_quality=90)
    files.append(o.jpeg)
imgs = [{"file": files[i % 3]} for i in range(9)]
for out in ("jpeg", "png"):
    ctx.run(imgs, "vertical", output=out)
    t = time.perf_counter()
    o = ctx.run(imgs, "vertical", output=out)
    print("%s job: %.2f ms, canvas %dx%d" % (out, (time.perf_counter() - t) * 1e3, o.width, o.height))
ctx.close()
