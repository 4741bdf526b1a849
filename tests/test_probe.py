"""Header probe + prepared record (index.js:678-741, wx.getImageInfo).  No GPU.

Files are written by PIL (libjpeg / libpng / libwebp) and the probe is held to what PIL reads back.
"""
import io
import struct

import pytest
from PIL import Image

from imagestitching_b200 import StitchError, StitchPage, get_image_info, prepare_image
from imagestitching_b200 import _lib as L

NAMES = {1: "up", 2: "up-mirrored", 3: "down", 4: "down-mirrored", 5: "left-mirrored", 6: "right", 7: "right-mirrored",
         8: "left"}


def _file(fmt, size=(123, 77), exif=None, mode="RGB", **kw):
    im = Image.new(mode, size, (10, 20, 30) if mode == "RGB" else 0)
    b = io.BytesIO()
    if exif is not None:
        ex = Image.Exif()
        ex[0x0112] = exif
        kw["exif"] = ex
    im.save(b, fmt, **kw)
    return b.getvalue()


@pytest.mark.parametrize("o", range(1, 9))
def test_jpeg_exif_orientation(o):
    data = _file("JPEG", exif=o)
    info = get_image_info(data)
    assert (info["width"], info["height"], info["type"]) == (123, 77, "jpeg")
    assert info["orientationExif"] == o and info["orientation"] == NAMES[o]
    assert Image.open(io.BytesIO(data)).getexif()[0x0112] == o


@pytest.mark.parametrize("fmt,typ,kw", [
    ("JPEG", "jpeg", {}), ("JPEG", "jpeg", {"progressive": True}), ("JPEG", "jpeg", {"optimize": True, "quality": 30}),
    ("PNG", "png", {}), ("GIF", "gif", {}), ("BMP", "bmp", {}),
    ("WEBP", "webp", {}), ("WEBP", "webp", {"lossless": True}),
])
@pytest.mark.parametrize("size", [(1, 1), (123, 77), (4032, 3024), (3, 9001)])
def test_sizes_and_types(fmt, typ, kw, size):
    if fmt == "WEBP" and max(size) > 16383:
        pytest.skip("WebP limit")
    mode = "P" if fmt == "GIF" else "RGB"
    data = _file(fmt, size, mode=mode, **kw)
    info = get_image_info(data)
    assert (info["width"], info["height"]) == Image.open(io.BytesIO(data)).size == size
    assert info["type"] == typ and info["orientationExif"] == 1 and info["orientation"] == "up"


def test_png_and_webp_exif():
    assert get_image_info(_file("PNG", exif=8))["orientation"] == "left"
    assert get_image_info(_file("WEBP", exif=3))["orientationExif"] == 3          # VP8X + EXIF chunk


def test_big_endian_tiff_and_long_type():
    # hand-built APP1: Motorola byte order, orientation stored as LONG
    tiff = b"MM\x00\x2a\x00\x00\x00\x08" + struct.pack(">H", 1) + struct.pack(">HHII", 0x0112, 4, 1, 5) + b"\0\0\0\0"
    app1 = b"\xff\xe1" + struct.pack(">H", 2 + 6 + len(tiff)) + b"Exif\0\0" + tiff
    base = _file("JPEG")
    data = base[:2] + app1 + base[2:]
    info = get_image_info(data)
    assert info["orientationExif"] == 5 and (info["width"], info["height"]) == (123, 77)


def test_truncated_and_garbage():
    data = _file("JPEG", exif=6)
    for cut in (0, 1, 3, 10, 30):
        with pytest.raises(StitchError) as e:
            get_image_info(data[:cut])
        assert e.value.status == L.E_BAD_IMAGE
    with pytest.raises(StitchError):
        get_image_info(b"hello world, not an image at all")
    with pytest.raises(StitchError):
        get_image_info(b"RIFF\x10\0\0\0WEBPVP8 ")
    # a corrupt EXIF block must not break the size
    bad = data.replace(b"Exif\0\0II", b"Exif\0\0IX").replace(b"Exif\0\0MM", b"Exif\0\0MX")
    assert bad != data
    assert get_image_info(bad)["orientationExif"] == 1


def test_prepared_record_feeds_the_page():
    """index.js:724-739: the record keeps WeChat's orientation STRING, so the draw is un-rotated (U:155,198)."""
    files = [_file("JPEG", (640, 480), exif=6), _file("PNG", (320, 200))]
    recs = [prepare_image(f, tempFilePath=f"mem://{i}") for i, f in enumerate(files)]
    assert recs[0]["orientation"] == "right" and recs[0]["naturalWidth"] == 640 and recs[0]["fileSize"] == len(files[0])
    assert recs[0]["preparedType"] == "jpeg" and recs[0]["prepared"] is True and recs[0]["tempFilePath"] == "mem://0"
    from imagestitching_b200 import compute_layout
    lay = compute_layout(recs, "vertical", "min")
    assert lay["natural"] == [[640, 480], [320, 200]] and lay["canvasW"] >= 320
