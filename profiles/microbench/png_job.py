"""One PNG-in / PNG-out job of bench.py's e2e_files.png_in leg (9 screenshot files), for a launch list."""
import sys
import time

sys.path.insert(0, ".")
import bench  # noqa: E402
from imagestitching_b200 import StitchContext  # noqa: E402

files = [bench.synth_screenshot_png(1290, 2796, seed=i % 3) for i in range(3)]
files = [files[i % 3] for i in range(9)]
ctx = StitchContext(0)
imgs = [{"file": f} for f in files]
ctx.run(imgs, "vertical", output="png")
t = time.perf_counter()
o = ctx.run(imgs, "vertical", output="png")
print("job: %.2f ms, canvas %dx%d, png %d bytes" % ((time.perf_counter() - t) * 1e3, o.width, o.height, len(o.png)))
ctx.close()
