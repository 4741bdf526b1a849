"""Decoder statistics of one photo-like and one screenshot-like PNG (STITCH_PNG_DEBUG=1 prints them)."""
import io
import os
import sys

os.environ["STITCH_PNG_DEBUG"] = "1"
sys.path.insert(0, ".")
from PIL import Image  # noqa: E402

from imagestitching_b200 import StitchContext  # noqa: E402
from tests.test_png_decode import _picture, _screenshot  # noqa: E402

ctx = StitchContext(0)
for name, arr, mode, kw in (("photo", _picture(4032, 3024, 3, seed=0), "RGB", {"compress_level": 6}),
                            ("photo_l9", _picture(2016, 1512, 3, seed=1, noise=2), "RGB", {"compress_level": 9}),
                            ("shot", _screenshot(1290, 2796, seed=0), "RGBA", {})):
    b = io.BytesIO()
    Image.fromarray(arr, mode).save(b, "PNG", **kw)
    print(name, len(b.getvalue()), flush=True)
    ctx.decode_png(b.getvalue())
ctx.close()
