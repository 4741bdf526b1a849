"""TEST INFRASTRUCTURE (oracle): ctypes wrapper of oracle/png_ref.c (PNG restatement of the 'png' export,
utils/canvas.js:205-242, pages/index/index.js:1577-1579).  Only tests/ and smoke() may import this."""
from __future__ import annotations

import ctypes as C

import numpy as np

from oracle.raster_c import load as _load


def _lib():
    lib = _load()
    if not getattr(lib, "_png_bound", False):
        lib.png_ref_encode_rgba.restype = C.c_int
        lib.png_ref_encode_rgba.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_size_t,
                                            C.POINTER(C.c_size_t)]
        lib.png_ref_crc32.restype = C.c_uint32
        lib.png_ref_crc32.argtypes = [C.c_void_p, C.c_size_t]
        lib.png_ref_code_lengths.restype = None
        lib.png_ref_code_lengths.argtypes = [C.c_void_p, C.c_void_p]
        lib._png_bound = True
    return lib


def encode_rgba(rgba: np.ndarray) -> bytes:
    a = np.ascontiguousarray(rgba, dtype=np.uint8)
    h, w = a.shape[:2]
    cap = h * (w * 4 + 1) * 2 + 4096 + 64 * ((h + 7) // 8)
    out = np.empty(cap, np.uint8)
    need = C.c_size_t(0)
    rc = _lib().png_ref_encode_rgba(a.ctypes.data, w, h, a.strides[0], out.ctypes.data, cap, C.byref(need))
    if rc:
        raise ValueError(f"png_ref_encode_rgba failed ({rc}, need {need.value})")
    return out[: need.value].tobytes()


def crc32(data: bytes) -> int:
    return int(_lib().png_ref_crc32(data, len(data)))


def code_lengths(freq) -> np.ndarray:
    f = np.ascontiguousarray(freq, dtype=np.uint32)
    assert f.shape == (286,)
    out = np.zeros(286, np.uint8)
    _lib().png_ref_code_lengths(f.ctypes.data, out.ctypes.data)
    return out
