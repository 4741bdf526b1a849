"""The host-side file parsers (jpeg_parse.h, png_parse.h, bmp_parse.h, gif_parse.h) on damaged files, under
AddressSanitizer + UBSan: untrusted bytes must end in a status, never in a read outside the file (the kernels then rely
on what an ACCEPTED header promises: offsets and lengths inside the upload).  CPU only."""
import io
import os
import subprocess
import sys

import numpy as np
import pytest
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

DRIVER = r'''
import ctypes, sys, numpy as np
lib = ctypes.CDLL(sys.argv[1])
lib.parse_fuzz_one.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t]
kind, path, rounds = int(sys.argv[2]), sys.argv[3], int(sys.argv[4])
good = open(path, "rb").read()
assert lib.parse_fuzz_one(kind, good, len(good)) == 0, "the intact file must parse"
rng = np.random.default_rng(kind)
accepted = 0
for r in range(rounds):
    b = bytearray(good)
    mode = r % 4
    if mode == 0:                                   # cut anywhere
        b = b[: int(rng.integers(0, len(b) + 1))]
    elif mode == 1:                                 # flip bytes in the first kilobyte (headers)
        for pos in rng.integers(0, min(len(b), 1024), int(rng.integers(1, 6))):
            b[pos] = int(rng.integers(0, 256))
    elif mode == 2:                                 # flip bytes anywhere
        for pos in rng.integers(0, len(b), int(rng.integers(1, 12))):
            b[pos] = int(rng.integers(0, 256))
    else:                                           # both
        for pos in rng.integers(0, min(len(b), 1024), 3):
            b[pos] = int(rng.integers(0, 256))
        b = b[: int(rng.integers(len(b) // 2, len(b) + 1))]
    st = lib.parse_fuzz_one(kind, bytes(b), len(b))
    assert st < 100, f"round {r}: an accepted header points outside the file ({st})"
    accepted += st == 0
print("accepted", accepted, "of", rounds)
'''


def _build():
    out = os.path.join(ROOT, "tests", "_build", "libparse_fuzz.so")
    src = os.path.join(ROOT, "tests", "native", "parse_fuzz.cpp")
    inc = os.path.join(ROOT, "imagestitching_b200", "csrc")
    deps = [src] + [os.path.join(inc, f) for f in ("jpeg_parse.h", "png_parse.h", "bmp_parse.h", "gif_parse.h")]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O1", "-g", "-std=c++17", "-fPIC", "-shared", "-fsanitize=address,undefined",
                               "-fno-sanitize-recover=undefined", "-I" + inc, "-o", out, src])
    return out


def _picture(w, h):
    y, x = np.mgrid[0:h, 0:w]
    a = np.stack([128 + 100 * np.sin(x / 17 + y / 29), 128 + 90 * np.cos(x / 23 - y / 13), (x * 3 + y * 5) % 256], -1)
    return Image.fromarray(np.clip(a, 0, 255).astype(np.uint8))


def _files(tmp):
    im = _picture(97, 61)
    out = []
    for name, kind, kw in (("base.jpg", 0, {"format": "JPEG", "quality": 85}),
                           ("prog.jpg", 0, {"format": "JPEG", "quality": 85, "progressive": True}),
                           ("rst.jpg", 0, {"format": "JPEG", "quality": 85, "restart_marker_blocks": 3}),
                           ("a.png", 1, {"format": "PNG"}), ("a.bmp", 2, {"format": "BMP"}), ("a.gif", 3, {"format": "GIF"})):
        p = os.path.join(tmp, name)
        (im.quantize(64) if name.endswith(".gif") else im).save(p, **kw)
        out.append((kind, p))
    p = os.path.join(tmp, "pal.bmp")
    im.quantize(16).save(p, "BMP")
    out.append((2, p))
    p = os.path.join(tmp, "pal.png")
    im.quantize(16).save(p, "PNG")
    out.append((1, p))
    return out


def _asan_runtime():
    try:
        return subprocess.check_output(["g++", "-print-file-name=libasan.so"], text=True).strip()
    except Exception:
        return ""


def test_parsers_survive_damaged_files(tmp_path):
    lib = _build()
    asan = _asan_runtime()
    if not asan or not os.path.exists(asan):
        pytest.skip("no AddressSanitizer runtime in this image")
    env = dict(os.environ, LD_PRELOAD=asan, ASAN_OPTIONS="detect_leaks=0:abort_on_error=1")
    for kind, path in _files(str(tmp_path)):
        r = subprocess.run([sys.executable, "-c", DRIVER, lib, str(kind), path, "600"], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, f"{path}:\n{r.stdout[-500:]}\n{r.stderr[-3000:]}"


def _plain_lib():
    """The same source without the sanitizers (loads into this process)."""
    out = os.path.join(ROOT, "tests", "_build", "libparse_plain.so")
    src = os.path.join(ROOT, "tests", "native", "parse_fuzz.cpp")
    inc = os.path.join(ROOT, "imagestitching_b200", "csrc")
    deps = [src] + [os.path.join(inc, f) for f in ("jpeg_parse.h", "png_parse.h", "bmp_parse.h", "gif_parse.h")]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I" + inc, "-o", out, src])
    import ctypes
    lib = ctypes.CDLL(out)
    lib.parse_fields.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p]
    return lib


def _fields(lib, kind, data):
    out = np.zeros(8, np.int64)
    lib.parse_fields(kind, data, len(data), out.ctypes.data)
    return out.tolist()


def test_parsers_read_what_the_files_say():
    """Header fields of the library's host-side parsers against what PIL and the test writers put into the files."""
    from tests import test_bmp_decode as TB
    from tests import test_gif_decode as TG
    from tests.test_jpeg_decode import _jpeg
    lib = _plain_lib()
    for name, data in TB.files():
        im = Image.open(io.BytesIO(data))
        st, w, h, bpp, top_down, stride, off, _ = _fields(lib, 2, data)
        assert st == 0 and (w, h) == im.size, name
        assert stride == (w * bpp + 31) // 32 * 4 and off + stride * h <= len(data), name
        if name != "pal8_os2":                       # (that entry of the BMP test list repeats the top-down file)
            assert top_down == ("topdown" in name), name
    for name, data in TG.files():
        im = Image.open(io.BytesIO(data))
        st, w, h, laced, min_code, transparent, stream_len, _ = _fields(lib, 3, data)
        assert st == 0 and (w, h) == im.size, name
        assert transparent == im.info.get("transparency", -1), name
        if "'interlace': 0" in name or "rows_in_order" in name:
            assert laced == 0, name
        elif "'interlace': 1" in name:
            assert laced == (min(w, h) >= 16), name      # PIL does not interlace pictures under 16 px
        assert 2 <= min_code <= 8 and 0 < stream_len < len(data), name
    for kw, want in (({"subsampling": 0}, (3, 1, 1, 0, 0)), ({"subsampling": 1}, (3, 2, 1, 0, 0)), ({"subsampling": 2}, (3, 2, 2, 0, 0)),
                     ({"gray": True}, (1, 1, 1, 0, 0)), ({"subsampling": 2, "progressive": True}, (3, 2, 2, 1, 10)),
                     ({"gray": True, "progressive": True}, (1, 1, 1, 1, 6))):
        data = _jpeg(123, 77, quality=80, **kw)
        f = _fields(lib, 0, data)
        assert f[0] == 0 and f[1:3] == [123, 77] and tuple(f[3:8]) == want, (kw, f)
    b = io.BytesIO()
    _picture(50, 31).save(b, "PNG")
    f = _fields(lib, 1, b.getvalue())
    assert f[:7] == [0, 50, 31, 8, 2, 0, 150], f
