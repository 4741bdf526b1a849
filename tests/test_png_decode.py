"""PNG inputs (utils/canvas.js:27-121 loadImageFrom, index.js:1441-1463; P:4 SUPPORTED_IMAGE_TYPES).

CPU: the oracle decoder (oracle/png_dec_ref.c: its own bit-by-bit inflater, unfilter, colour conversion) is pinned to
     Python's zlib and to PIL / libpng on every colour type, bit depth, filter and deflate strategy; the pieces the GPU
     kernels are made of (imagestitching_b200/csrc/png_inflate_core.h) are driven sequentially by tests/native/pinf_host.cpp
     and held against zlib too.
GPU: stitch_decode_png and STITCH_MEM_HOST_PNG job inputs are byte-identical to the oracle decoder.
"""
import ctypes
import io
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest
from PIL import Image

from imagestitching_b200 import _lib as L
from oracle import png_dec_c as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _picture(w, h, c, seed=0, noise=6):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    base = (np.sin(x / 17.0) + np.cos(y / 23.0)) * 60 + 128
    a = np.stack([(base + 20 * k + rng.normal(0, noise, (h, w))).clip(0, 255) for k in range(c)], -1).astype(np.uint8)
    return a


def _screenshot(w, h, seed=0):
    """Flat areas, text-like speckle, a gradient: long matches and runs, like a phone screenshot."""
    rng = np.random.default_rng(seed)
    a = np.full((h, w, 4), 255, np.uint8)
    a[: h // 8] = (32, 120, 240, 255)
    for k in range(12):
        y0 = int(rng.integers(h // 8, max(h // 8 + 1, h - 20)))
        x0 = int(rng.integers(0, max(1, w - 60)))
        a[y0:y0 + 14, x0:x0 + 50, :3] = rng.integers(0, 2, (min(14, h - y0), min(50, w - x0), 1)) * 200
    a[h // 2: h // 2 + h // 6, :, 0] = np.linspace(0, 255, w).astype(np.uint8)
    return a


def _png(arr, mode, **kw):
    b = io.BytesIO()
    Image.fromarray(arr, mode).save(b, "PNG", **kw)
    return b.getvalue()


def _pil_rgba(data):
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGBA"))


def _chunk(typ, body):
    return struct.pack(">I", len(body)) + typ + body + struct.pack(">I", zlib.crc32(typ + body) & 0xFFFFFFFF)


def _raw_png(w, h, depth, ctype, rows, filters=None, extra=(), level=6, idat_split=None, strategy=zlib.Z_DEFAULT_STRATEGY):
    """A PNG file from already-packed rows (bytes per row), with the per-row filter types given (0-4) or chosen in turn."""
    bits = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype] * depth
    bpp = max(1, bits // 8)
    rb = (w * bits + 7) // 8
    out = bytearray()
    prev = np.zeros(rb, np.int32)
    for y in range(h):
        cur = np.frombuffer(bytes(rows[y]), np.uint8).astype(np.int32)
        assert cur.size == rb
        ft = filters[y] if filters is not None else y % 5
        left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]]) if rb > bpp else np.zeros(rb, np.int32)
        ul = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]]) if rb > bpp else np.zeros(rb, np.int32)
        if ft == 0:
            pred = np.zeros(rb, np.int32)
        elif ft == 1:
            pred = left
        elif ft == 2:
            pred = prev
        elif ft == 3:
            pred = (left + prev) >> 1
        else:
            p = left + prev - ul
            pa, pb, pc = abs(p - left), abs(p - prev), abs(p - ul)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, ul))
        out.append(ft)
        out += ((cur - pred) & 0xFF).astype(np.uint8).tobytes()
        prev = cur
    co = zlib.compressobj(level, zlib.DEFLATED, 15, 8, strategy)
    z = co.compress(bytes(out)) + co.flush()
    body = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 0))
    for typ, data in extra:
        body += _chunk(typ, data)
    if idat_split:
        for o in range(0, len(z), idat_split):
            body += _chunk(b"IDAT", z[o:o + idat_split])
    else:
        body += _chunk(b"IDAT", z)
    return body + _chunk(b"IEND", b"")


def _adam7_png(w, h, depth, ctype, rows, extra=(), level=6):
    """The interlaced (Adam7) form of the picture whose packed rows are `rows`: the pixels of every pass re-packed into
    that pass's own rows, each pass filtered on its own (filter types in turn), one zlib stream."""
    bits = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype] * depth
    bpp = max(1, bits // 8)
    px = np.stack([np.unpackbits(np.frombuffer(bytes(r), np.uint8))[: w * bits].reshape(w, bits) for r in rows])   # [h, w, bits]
    out = bytearray()
    for x0, y0, dx, dy, pw, ph, rb in P.adam7_passes(w, h, bits):
        sub = px[y0::dy, x0::dx].reshape(ph, pw * bits)
        prev = np.zeros(rb, np.int32)
        for y in range(ph):
            cur = np.packbits(sub[y]).astype(np.int32)
            assert cur.size == rb
            ft = (y + x0) % 5
            left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]]) if rb > bpp else np.zeros(rb, np.int32)
            ul = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]]) if rb > bpp else np.zeros(rb, np.int32)
            if ft == 0:
                pred = np.zeros(rb, np.int32)
            elif ft == 1:
                pred = left
            elif ft == 2:
                pred = prev
            elif ft == 3:
                pred = (left + prev) >> 1
            else:
                p = left + prev - ul
                pa, pb, pc = abs(p - left), abs(p - prev), abs(p - ul)
                pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, ul))
            out.append(ft)
            out += ((cur - pred) & 0xFF).astype(np.uint8).tobytes()
            prev = cur
    body = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 1))
    for typ, data in extra:
        body += _chunk(typ, data)
    return body + _chunk(b"IDAT", zlib.compress(bytes(out), level)) + _chunk(b"IEND", b"")


def _interlaced_matrix():
    """(name, interlaced file, the same picture not interlaced) over colour types, depths and awkward sizes."""
    rng = np.random.default_rng(17)
    files = []
    for (w, h) in ((67, 41), (1, 1), (3, 2), (8, 8), (9, 17), (5, 1), (1, 9)):
        for depth, ctype in ((1, 0), (2, 0), (4, 3), (8, 0), (8, 2), (8, 3), (8, 4), (8, 6), (16, 2), (16, 6)):
            if (w, h) != (67, 41) and (depth, ctype) not in ((1, 0), (4, 3), (8, 2), (8, 6)):
                continue
            ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
            rb = (w * ch * depth + 7) // 8
            rows = [rng.integers(0, 256, rb, dtype=np.uint8) for _ in range(h)]
            extra = []
            if ctype == 3:
                n_pal = min(1 << depth, 200)
                if depth == 8:
                    rows = [r % n_pal for r in rows]
                extra = [(b"PLTE", rng.integers(0, 256, n_pal * 3, dtype=np.uint8).tobytes()),
                         (b"tRNS", rng.integers(0, 256, max(1, n_pal // 2), dtype=np.uint8).tobytes())]
            files.append((f"{w}x{h}_ct{ctype}_d{depth}", _adam7_png(w, h, depth, ctype, rows, extra), _raw_png(w, h, depth, ctype, rows, extra=extra)))
    a = _picture(333, 211, 4, seed=21)
    rows = [a[y].tobytes() for y in range(211)]
    files.append(("photo_rgba", _adam7_png(333, 211, 8, 6, rows), _raw_png(333, 211, 8, 6, rows)))
    s = _screenshot(390, 300)
    rows = [s[y].tobytes() for y in range(300)]
    files.append(("screenshot", _adam7_png(390, 300, 8, 6, rows), _raw_png(390, 300, 8, 6, rows)))
    return files


def _file_matrix():
    """(name, file bytes) over colour types, depths, filters, tRNS, chunking and deflate strategies."""
    rng = np.random.default_rng(7)
    files = []
    for mode, c in (("RGB", 3), ("RGBA", 4), ("L", 1), ("LA", 2)):
        a = _picture(301, 217, c, seed=c)
        files.append((f"pil_{mode}", _png(a if c > 1 else a[..., 0], mode)))
    files.append(("pil_palette", _png_q(_picture(300, 200, 3, seed=9), 37)))
    files.append(("pil_level1", _png(_picture(257, 129, 3, seed=2), "RGB", compress_level=1)))
    files.append(("pil_level9_opt", _png(_picture(257, 129, 3, seed=3), "RGB", compress_level=9, optimize=True)))
    files.append(("pil_stored", _png(_picture(130, 67, 4, seed=4), "RGBA", compress_level=0)))
    files.append(("pil_screenshot", _png(_screenshot(390, 844), "RGBA")))
    files.append(("pil_1x1", _png(np.array([[[9, 8, 7]]], np.uint8), "RGB")))
    # hand-made files: every filter in turn, low bit depths, 16 bit, tRNS keys, split IDAT, fixed / RLE / Huffman-only
    w, h = 67, 41
    for depth in (1, 2, 4, 8, 16):
        rb = (w * depth + 7) // 8
        rows = [rng.integers(0, 256, rb, dtype=np.uint8) for _ in range(h)]
        extra = [(b"tRNS", struct.pack(">H", 1 if depth < 16 else 0x1234))] if depth in (2, 16) else []
        if depth == 16:
            rows[3][0:2] = (0x12, 0x34)
        files.append((f"gray{depth}", _raw_png(w, h, depth, 0, rows, extra=extra)))
    for depth in (1, 2, 4, 8):
        rb = (w * depth + 7) // 8
        n_pal = min(1 << depth, 200)
        rows = [rng.integers(0, 256, rb, dtype=np.uint8) for _ in range(h)]
        if depth == 8:
            rows = [r % n_pal for r in rows]
        pal = rng.integers(0, 256, n_pal * 3, dtype=np.uint8).tobytes()
        trns = rng.integers(0, 256, max(1, n_pal // 2), dtype=np.uint8).tobytes()
        files.append((f"palette{depth}", _raw_png(w, h, depth, 3, rows, extra=[(b"PLTE", pal), (b"tRNS", trns)])))
    for ctype, ch in ((2, 3), (4, 2), (6, 4)):
        for depth in (8, 16):
            rb = w * ch * depth // 8
            rows = [rng.integers(0, 256, rb, dtype=np.uint8) for _ in range(h)]
            extra = []
            if ctype == 2:
                keyb = bytes(rows[5][:rb // w]) if depth == 16 else bytes([0, rows[5][0], 0, rows[5][1], 0, rows[5][2]])
                extra = [(b"tRNS", keyb)]
            files.append((f"ct{ctype}_d{depth}", _raw_png(w, h, depth, ctype, rows, extra=extra)))
    smooth = _picture(200, 150, 3, seed=11)
    rows = [smooth[y].tobytes() for y in range(150)]
    for name, kw in (("fixed", {"strategy": zlib.Z_FIXED}), ("rle", {"strategy": zlib.Z_RLE}), ("huffman_only", {"strategy": zlib.Z_HUFFMAN_ONLY}),
                     ("split_idat_100", {"idat_split": 100}), ("split_idat_1", {"idat_split": 1}), ("paeth_only", {"filters": [4] * 150}),
                     ("avg_only", {"filters": [3] * 150}), ("level1", {"level": 1})):
        files.append((name, _raw_png(200, 150, 8, 2, rows, **kw)))
    return files


def _png_q(arr, colours):
    b = io.BytesIO()
    Image.fromarray(arr, "RGB").quantize(colours).save(b, "PNG")
    return b.getvalue()


_MATRIX = None


def matrix():
    global _MATRIX
    if _MATRIX is None:
        _MATRIX = _file_matrix()
    return _MATRIX


def _expected(data):
    """PIL's RGBA, except where PIL's conversion is not the stated one (16-bit gray goes through mode I)."""
    return _pil_rgba(data)


# PIL applies a tRNS colour key of these modes its own way (16-bit keys against 8-bit pixels, a 2-bit key against the
# widened value) and widens 16-bit gray through mode I: their alpha / values are checked from first principles below.
PIL_RGB_ONLY = {"gray2", "ct2_d16"}
PIL_SKIP = {"gray16"}


# ---------------- CPU ----------------
def test_oracle_equals_pil_and_zlib():
    for name, data in matrix():
        got = P.decode(data)
        assert np.array_equal(got, P.decode(data, use_zlib=True)), name      # the oracle's own inflater == zlib
        if name in PIL_SKIP:
            continue
        want = _expected(data)
        if name in PIL_RGB_ONLY:
            assert np.array_equal(got[..., :3], want[..., :3]), name
        else:
            assert np.array_equal(got, want), name


def test_oracle_interlaced_equals_plain_and_pil():
    """Adam7: the interlaced file decodes to the pixels of the plain file of the same picture -- in PIL / libpng (so the
    test's interlacer is sound) and in the oracle, whose own inflater agrees with zlib on it."""
    for name, laced, plain in _interlaced_matrix():
        assert np.array_equal(_pil_rgba(laced), _pil_rgba(plain)), name
        got = P.decode(laced)
        assert np.array_equal(got, P.decode(plain)), name
        assert np.array_equal(got, P.decode(laced, use_zlib=True)), name


def test_oracle_low_depth_gray_and_16_bit_values():
    """The modes PIL converts its own way: checked from first principles."""
    rng = np.random.default_rng(3)
    w, h = 19, 5
    for depth in (1, 2, 4):
        rb = (w * depth + 7) // 8
        rows = [rng.integers(0, 256, rb, dtype=np.uint8) for _ in range(h)]
        got = P.decode(_raw_png(w, h, depth, 0, rows))
        for y in range(h):
            bits = np.unpackbits(rows[y])[: w * depth].reshape(w, depth)
            v = (bits * (1 << np.arange(depth - 1, -1, -1))).sum(1)
            assert np.array_equal(got[y, :, 0], v * (255 // ((1 << depth) - 1)))
            assert (got[y, :, 3] == 255).all() and np.array_equal(got[y, :, 0], got[y, :, 2])
    rows = [rng.integers(0, 256, w * 2, dtype=np.uint8) for _ in range(h)]
    rows[1][4:6] = (0xAB, 0xCD)
    got = P.decode(_raw_png(w, h, 16, 0, rows, extra=[(b"tRNS", b"\xab\xcd")]))
    for y in range(h):
        assert np.array_equal(got[y, :, 1], rows[y][0::2])
    assert got[1, 2, 3] == 0 and (np.delete(got[..., 3].ravel(), 1 * w + 2) == 255).all()
    # colour keys: 2-bit gray (raw sample value), 16-bit RGB (all 48 bits)
    rows = [rng.integers(0, 256, (w * 2 + 7) // 8, dtype=np.uint8) for _ in range(h)]
    got = P.decode(_raw_png(w, h, 2, 0, rows, extra=[(b"tRNS", b"\x00\x01")]))
    for y in range(h):
        v = (np.unpackbits(rows[y])[: w * 2].reshape(w, 2) * (2, 1)).sum(1)
        assert np.array_equal(got[y, :, 3], np.where(v == 1, 0, 255))
    rows = [rng.integers(0, 256, w * 6, dtype=np.uint8) for _ in range(h)]
    rows[2][12:18] = rows[0][0:6]
    got = P.decode(_raw_png(w, h, 16, 2, rows, extra=[(b"tRNS", bytes(rows[0][0:6]))]))
    assert got[0, 0, 3] == 0 and got[2, 2, 3] == 0 and (got[..., 3] == 0).sum() == 2


def test_oracle_rejects_damage():
    data = bytearray(_png(_picture(64, 48, 3), "RGB"))
    bad = bytearray(data)
    bad[60] ^= 0xFF   # inside IDAT: CRC fails in the oracle's chunk walk
    with pytest.raises(P.PngError):
        P.decode(bytes(bad))
    with pytest.raises(P.PngError):
        P.decode(b"\x89PNG\r\n\x1a\n" + bytes(20))


def _host_lib():
    out = os.path.join(ROOT, "tests", "_build", "libpinf_host.so")
    src = os.path.join(ROOT, "tests", "native", "pinf_host.cpp")
    hdr = os.path.join(ROOT, "imagestitching_b200", "csrc", "png_inflate_core.h")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I" + os.path.dirname(hdr), src, "-o", out])
    lib = ctypes.CDLL(out)
    lib.pinf_host_inflate.argtypes = [ctypes.c_char_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
    return lib


def _host_inflate(lib, z, raw_len):
    out = np.zeros(max(raw_len, 1), np.uint8)
    st = np.zeros(6, np.uint32)
    rc = lib.pinf_host_inflate(z, len(z), raw_len, out.ctypes.data, st.ctypes.data)
    return rc, out[:raw_len].tobytes(), st


def test_kernel_pieces_on_the_cpu_equal_zlib():
    """find -> validate -> count -> walk -> write -> resolve -> bytes with the kernels' own inline functions."""
    lib = _host_lib()
    rng = np.random.default_rng(1)
    walk = (np.cumsum(rng.integers(-3, 4, 400000)) % 256).astype(np.uint8).tobytes()
    flat = (b"\x00" * 5000 + b"\x01\x02\x03" * 3000 + bytes(rng.integers(0, 256, 2000, dtype=np.uint8))) * 30
    noise = bytes(rng.integers(0, 256, 200000, dtype=np.uint8))
    cases = []
    for raw in (walk, flat, noise, b"", b"a", b"ab" * 40000):
        for kw in ({"level": 6}, {"level": 1}, {"level": 9, "memLevel": 9}, {"level": 0}, {"level": 6, "strategy": zlib.Z_FIXED},
                   {"level": 6, "strategy": zlib.Z_RLE}, {"level": 6, "strategy": zlib.Z_HUFFMAN_ONLY}, {"level": 6, "memLevel": 1}):
            cases.append((raw, kw))
    found_by_finder = 0
    for raw, kw in cases:
        co = zlib.compressobj(**kw)
        z = co.compress(raw) + co.flush()
        rc, got, st = _host_inflate(lib, z, len(raw))
        assert rc == 0 and got == raw, (kw, len(raw), rc, st.tolist())
        assert st[1] <= st[2], "a false candidate survived validation + count"   # not an error if it ever fires: the walk drops them
        found_by_finder += int(st[2] - st[3])
    assert found_by_finder > 200   # the dynamic blocks really are found by the header test, not by the walker
    # a stream cut short, a flipped bit, a wrong length: rejected
    co = zlib.compressobj(6)
    z = co.compress(walk) + co.flush()
    assert _host_inflate(lib, z[: len(z) // 2], len(walk))[0] != 0
    assert _host_inflate(lib, z, len(walk) - 1)[0] != 0
    assert _host_inflate(lib, z, len(walk) + 1)[0] != 0


def test_finder_table_test_equals_the_loop():
    """The block finder's table-driven header test == the plain one, on 2 M random windows and every offset of a stream."""
    lib = _host_lib()
    lib.pinf_host_prefilter_check.restype = ctypes.c_uint32
    lib.pinf_host_prefilter_check.argtypes = [ctypes.c_char_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p]
    rng = np.random.default_rng(2)
    z = zlib.compress((np.cumsum(rng.integers(-3, 4, 300000)) % 256).astype(np.uint8).tobytes(), 6)
    hits = ctypes.c_uint32(0)
    assert lib.pinf_host_prefilter_check(z, len(z), 2_000_000, ctypes.byref(hits)) == 0
    assert hits.value > 1000


# ---------------- GPU ----------------
@pytest.mark.gpu
def test_gpu_decoder_equals_oracle_on_the_matrix(ctx):
    for name, data in matrix():
        got = ctx.decode_png(data)
        assert np.array_equal(got, P.decode(data)), name


@pytest.mark.gpu
def test_gpu_decoder_interlaced(ctx):
    """Adam7 files: seven passes through transpose / unfilter / conversion, scattered into the bitmap."""
    for name, laced, plain in _interlaced_matrix():
        assert np.array_equal(ctx.decode_png(laced), P.decode(laced)), name
    # a job that mixes an interlaced and a plain file
    from tests import util as U
    _, laced, plain = _interlaced_matrix()[-2]
    a = ctx.run([{"file": laced}, {"file": plain}], "vertical", platform="ios")
    b = ctx.run([{"file": plain}, {"file": plain}], "vertical", platform="ios")
    assert np.array_equal(a.rgba, b.rgba)


@pytest.mark.gpu
def test_gpu_decoder_sizes_and_shapes(ctx):
    for i, (w, h) in enumerate([(1, 1), (2, 1), (1, 70), (33, 31), (32, 32), (640, 481), (1, 4000), (5000, 3), (1290, 2796)]):
        for c, mode in ((3, "RGB"), (4, "RGBA")):
            data = _png(_picture(w, h, c, seed=i), mode)
            assert np.array_equal(ctx.decode_png(data), P.decode(data)), (w, h, mode)


@pytest.mark.gpu
def test_gpu_decoder_full_size_photo_and_screenshot(ctx):
    photo = _png(_picture(4032, 3024, 3, seed=5), "RGB", compress_level=6)
    assert np.array_equal(ctx.decode_png(photo), P.decode(photo))
    shot = _png(_screenshot(1290, 2796, seed=6), "RGBA")
    assert np.array_equal(ctx.decode_png(shot), P.decode(shot))


@pytest.mark.gpu
def test_gpu_png_job_inputs_equal_bitmap_job(ctx):
    """A job given PNG files == the same job given the oracle's bitmaps of those files (sync, async and mixed with JPEG)."""
    dims = [(403, 302), (390, 844), (640, 480)]
    files = [_png(_picture(403, 302, 3, seed=1), "RGB"), _png(_screenshot(390, 844, seed=2), "RGBA"),
             _png(_picture(640, 480, 4, seed=3), "RGBA")]
    bitmaps = [P.decode(f) for f in files]
    imgs_f = [{"naturalWidth": w, "naturalHeight": h, "file": f} for (w, h), f in zip(dims, files)]
    imgs_b = [{"naturalWidth": w, "naturalHeight": h, "bitmap": b} for (w, h), b in zip(dims, bitmaps)]
    for direction in ("vertical", "horizontal"):
        want = ctx.run(imgs_b, direction)
        got = ctx.run(imgs_f, direction)
        assert np.array_equal(got.rgba, want.rgba), direction
        # dims read from the file when the record carries none
        got2 = ctx.run([{"file": f} for f in files], direction)
        assert np.array_equal(got2.rgba, want.rgba)
    ctx.set_lanes(3)
    try:
        tickets = [ctx.submit(imgs_f, "vertical", output="png") for _ in range(5)]
        outs = [ctx.wait(t) for t in tickets]
        want = ctx.run(imgs_b, "vertical", output="png")
        for o in outs:
            assert o.png == want.png
    finally:
        ctx.set_lanes(1)
    # JPEG and PNG files in one job
    b = io.BytesIO()
    Image.fromarray(_picture(403, 302, 3, seed=8), "RGB").save(b, "JPEG", quality=90)
    from oracle import jpeg_c as J
    mixed_f = [{"file": b.getvalue()}, {"file": files[1]}]
    mixed_b = [{"naturalWidth": 403, "naturalHeight": 302, "bitmap": J.decode(b.getvalue())},
               {"naturalWidth": 390, "naturalHeight": 844, "bitmap": bitmaps[1]}]
    assert np.array_equal(ctx.run(mixed_f, "vertical").rgba,
                          ctx.run(mixed_b, "vertical").rgba)


@pytest.mark.gpu
def test_gpu_decoder_errors(ctx):
    from imagestitching_b200 import StitchError
    good = _png(_picture(200, 120, 3, seed=4), "RGB")
    # the interlace flag set on a stream that is not interlaced: the inflated data has the wrong length
    lace = bytearray(good)
    lace[28] = 1
    lace[29:33] = struct.pack(">I", zlib.crc32(bytes(lace[12:29])) & 0xFFFFFFFF)
    cases = {"interlaced": bytes(lace), "not png": b"\xff\xd8\xff\xe0" + bytes(64), "cut": good[: len(good) // 2],
             "bad IHDR crc": good[:20] + bytes([good[20] ^ 1]) + good[21:]}
    # a damaged deflate stream with correct chunk framing: only the inflater can notice
    idat_at = good.index(b"IDAT")
    (n,) = struct.unpack(">I", good[idat_at - 4: idat_at])
    body = bytearray(good[idat_at + 4: idat_at + 4 + n])
    for k in range(40, min(len(body), 400), 7):
        body[k] ^= 0x5A
    cases["damaged stream"] = good[: idat_at - 4] + _chunk(b"IDAT", bytes(body)) + good[idat_at + 8 + n:]
    # stream that inflates to the wrong amount of data
    raw = zlib.decompress(bytes(good[idat_at + 4: idat_at + 4 + n]))
    cases["short data"] = good[: idat_at - 4] + _chunk(b"IDAT", zlib.compress(raw[:-10])) + good[idat_at + 8 + n:]
    for name, data in cases.items():
        with pytest.raises(StitchError) as ei:
            ctx.decode_png(data)
        assert ei.value.status == L.E_BAD_IMAGE, name
    # the context still works afterwards
    assert np.array_equal(ctx.decode_png(good), P.decode(good))


@pytest.mark.gpu
def test_gpu_decoder_survives_damaged_streams(ctx):
    """Bit flips inside the deflate stream (chunk framing and CRCs repaired, so only the inflater can notice): the GPU
    decoder and the oracle agree on accept / reject, on the bytes when both accept, and the context keeps working."""
    from imagestitching_b200 import StitchError
    rng = np.random.default_rng(11)
    seeds = [_png(_picture(160, 120, 3, seed=21), "RGB"), _png(_screenshot(200, 300, seed=22), "RGBA"),
             _png(_picture(97, 61, 4, seed=23), "RGBA", compress_level=1)]
    rejected = accepted = 0
    for base in seeds:
        at = base.index(b"IDAT")
        (n,) = struct.unpack(">I", base[at - 4: at])
        for trial in range(14):
            body = bytearray(base[at + 4: at + 4 + n])
            for _ in range(int(rng.integers(1, 4))):
                body[int(rng.integers(2, n))] ^= 1 << int(rng.integers(0, 8))
            if trial == 13:
                body = body[: n // 3]            # the stream simply stops
            data = base[: at - 4] + _chunk(b"IDAT", bytes(body)) + base[at + 8 + n:]
            try:
                want = P.decode(data)
            except P.PngError:
                want = None
            try:
                got = ctx.decode_png(data)
            except StitchError as e:
                assert e.status == L.E_BAD_IMAGE
                got = None
            assert (got is None) == (want is None), (trial, "GPU and oracle disagree on whether the stream is valid")
            if want is None:
                rejected += 1
            else:
                accepted += 1
                assert np.array_equal(got, want), trial
    assert rejected >= 10 and accepted + rejected == 42
    good = seeds[0]
    assert np.array_equal(ctx.decode_png(good), P.decode(good))


@pytest.mark.gpu
def test_gpu_png_async_error_and_batch_of_unequal_files(ctx):
    """A damaged PNG among good ones fails the job at stitch_wait with the page's message; the next job on that lane is
    clean.  Nine files of nine different shapes and colour types in one job."""
    from imagestitching_b200 import StitchError
    good = _png(_picture(200, 120, 3, seed=4), "RGB")
    at = good.index(b"IDAT")
    (n,) = struct.unpack(">I", good[at - 4: at])
    body = bytearray(good[at + 4: at + 4 + n])
    for k in range(40, min(len(body), 400), 7):
        body[k] ^= 0x5A
    bad = good[: at - 4] + _chunk(b"IDAT", bytes(body)) + good[at + 8 + n:]
    ctx.set_lanes(2)
    try:
        t_bad = ctx.submit([{"file": good}, {"file": bad}], "vertical")
        t_ok = ctx.submit([{"file": good}, {"file": good}], "vertical")
        with pytest.raises(StitchError) as ei:
            ctx.wait(t_bad)
        assert ei.value.status == L.E_BAD_IMAGE and "解码异常" in ei.value.message
        ok = ctx.wait(t_ok)
        again = ctx.wait(ctx.submit([{"file": good}, {"file": good}], "vertical"))
        assert np.array_equal(ok.rgba, again.rgba)
    finally:
        ctx.set_lanes(1)
    files = [data for name, data in matrix() if name in ("pil_RGB", "pil_RGBA", "pil_L", "pil_LA", "pil_palette", "gray1", "palette4",
                                                         "ct6_d16", "fixed")]
    assert len(files) == 9
    bitmaps = [P.decode(f) for f in files]
    want = ctx.run([{"naturalWidth": b.shape[1], "naturalHeight": b.shape[0], "bitmap": b} for b in bitmaps], "horizontal")
    got = ctx.run([{"file": f} for f in files], "horizontal")
    assert np.array_equal(got.rgba, want.rgba)
