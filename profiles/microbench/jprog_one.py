"""One progressive 12 MP file decoded a few times (for an ncu launch list of the progressive decoder's kernels)."""
import io, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from PIL import Image, ImageFile
from imagestitching_b200 import StitchContext
ImageFile.MAXBLOCK = 1 << 26
w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4032, 3024)
noise = int(sys.argv[3]) if len(sys.argv) > 3 else 12
rng = np.random.default_rng(1)
y, x = np.mgrid[0:h, 0:w]
a = np.stack([128 + 100 * np.sin(x / 17 + y / 29), 128 + 90 * np.cos(x / 23 - y / 13), (x * 3 + y * 5) % 256], -1)
a = np.clip(a + rng.integers(-noise, noise + 1, a.shape), 0, 255).astype(np.uint8)
b = io.BytesIO()
Image.fromarray(a).save(b, "JPEG", quality=90, subsampling=2, progressive=True)
f = b.getvalue()
ctx = StitchContext(0)
ctx.decode_jpeg(f)
t0 = time.perf_counter()
ctx.decode_jpeg(f)
print(f"{w}x{h} noise {noise}: {len(f) / 1e6:.2f} MB, decode_jpeg {1e3 * (time.perf_counter() - t0):.1f} ms (incl. 48 MB download)")
ctx.close()
