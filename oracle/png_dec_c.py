"""TEST INFRASTRUCTURE (oracle): PNG file -> RGBA8 bitmap on the CPU, the restatement of what loadImageFrom
(utils/canvas.js:27-121; pages/index/index.js:1441-1463) gets from the platform decoder.  Chunk walk here (ISO 15948),
inflate / unfilter / colour conversion in oracle/png_dec_ref.c.  Only tests/, smoke() and bench.py's CPU legs may
import this."""
from __future__ import annotations

import ctypes as C
import struct
import zlib

import numpy as np

from oracle.raster_c import load as _load

SIG = b"\x89PNG\r\n\x1a\n"
CHANNELS = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}


def _lib():
    lib = _load()
    if not getattr(lib, "_pngdec_bound", False):
        lib.png_dec_ref_inflate.restype = C.c_int
        lib.png_dec_ref_inflate.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t]
        lib.png_dec_ref_unfilter.restype = C.c_int
        lib.png_dec_ref_unfilter.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_int]
        lib.png_dec_ref_to_rgba.restype = None
        lib.png_dec_ref_to_rgba.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                            C.c_void_p, C.c_void_p]
        lib._pngdec_bound = True
    return lib


class PngError(ValueError):
    pass


def chunks(data: bytes):
    """(type, payload) of every chunk up to IEND; CRC of each one checked."""
    if data[:8] != SIG:
        raise PngError("not a PNG file")
    o = 8
    while o + 12 <= len(data):
        (n,) = struct.unpack(">I", data[o:o + 4])
        typ = data[o + 4:o + 8]
        body = data[o + 8:o + 8 + n]
        if len(body) != n or o + 12 + n > len(data):
            raise PngError("truncated chunk")
        (crc,) = struct.unpack(">I", data[o + 8 + n:o + 12 + n])
        if zlib.crc32(typ + body) & 0xFFFFFFFF != crc:
            raise PngError(f"bad CRC in {typ!r}")
        yield typ, body
        if typ == b"IEND":
            return
        o += 12 + n


def header(data: bytes) -> dict:
    it = chunks(data)
    typ, body = next(it)
    if typ != b"IHDR" or len(body) != 13:
        raise PngError("IHDR missing")
    w, h, depth, ctype, comp, flt, lace = struct.unpack(">IIBBBBB", body)
    if ctype not in CHANNELS or comp or flt:
        raise PngError("bad IHDR")
    bits = CHANNELS[ctype] * depth
    return {"w": w, "h": h, "depth": depth, "ctype": ctype, "interlace": lace, "bpp": max(1, bits // 8),
            "row_bytes": (w * bits + 7) // 8}


def inflate(z: bytes, out_len: int) -> np.ndarray:
    """zlib stream -> exactly out_len bytes, through the oracle's own sequential inflater."""
    out = np.empty(out_len, np.uint8)
    rc = _lib().png_dec_ref_inflate(z, len(z), out.ctypes.data, out_len)
    if rc:
        raise PngError(f"inflate failed ({rc})")
    return out


# Adam7 (ISO 15948 clause 8.2): pass -> (x0, y0, dx, dy); a pass holds the pixels (x0 + i dx, y0 + j dy)
ADAM7 = ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2))


def adam7_passes(w: int, h: int, bits: int):
    """(x0, y0, dx, dy, pass width, pass height, row bytes) of the passes that hold pixels, in stream order."""
    out = []
    for x0, y0, dx, dy in ADAM7:
        pw, ph = (w - x0 + dx - 1) // dx, (h - y0 + dy - 1) // dy
        if pw > 0 and ph > 0:
            out.append((x0, y0, dx, dy, pw, ph, (pw * bits + 7) // 8))
    return out


def decode(data: bytes, use_zlib: bool = False) -> np.ndarray:
    """RGBA8 [h, w, 4], straight alpha.  use_zlib=True swaps the oracle's inflater for Python's zlib (the pin)."""
    H = header(data)
    if H["interlace"] > 1:
        raise PngError("unknown interlace method")
    palette = np.zeros((256, 4), np.uint8)
    palette[:, 3] = 255
    key = np.zeros(3, np.uint16)
    has_key = 0
    idat = []
    for typ, body in chunks(data):
        if typ == b"PLTE":
            n = len(body) // 3
            palette[:n, :3] = np.frombuffer(body, np.uint8).reshape(n, 3)
        elif typ == b"tRNS" and not idat:
            if H["ctype"] == 3:
                palette[:len(body), 3] = np.frombuffer(body, np.uint8)
            elif H["ctype"] == 0 and len(body) == 2:
                has_key, key[0] = 1, struct.unpack(">H", body)[0]
            elif H["ctype"] == 2 and len(body) == 6:
                has_key, key[:] = 1, struct.unpack(">HHH", body)
        elif typ == b"IDAT":
            idat.append(body)
    z = b"".join(idat)
    bits = CHANNELS[H["ctype"]] * H["depth"]
    passes = adam7_passes(H["w"], H["h"], bits) if H["interlace"] else [(0, 0, 1, 1, H["w"], H["h"], H["row_bytes"])]
    raw_len = sum((rb + 1) * ph for *_, ph, rb in passes)
    if use_zlib:
        raw = np.frombuffer(zlib.decompress(z), np.uint8).copy()
        if raw.size != raw_len:
            raise PngError("wrong amount of image data")
    else:
        raw = inflate(z, raw_len)
    lib = _lib()
    out = np.empty((H["h"], H["w"], 4), np.uint8)
    off = 0
    for x0, y0, dx, dy, pw, ph, rb in passes:        # every pass is a small image of its own: filtered and unfiltered alone
        part = raw[off:off + (rb + 1) * ph]
        off += (rb + 1) * ph
        if lib.png_dec_ref_unfilter(part.ctypes.data, ph, rb, H["bpp"]):
            raise PngError("bad filter type")
        px = np.empty((ph, pw, 4), np.uint8)
        lib.png_dec_ref_to_rgba(part.ctypes.data, pw, ph, rb, H["depth"], H["ctype"], palette.ctypes.data, has_key,
                                key.ctypes.data, px.ctypes.data)
        out[y0::dy, x0::dx] = px
    return out
