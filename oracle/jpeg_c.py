"""TEST INFRASTRUCTURE (oracle): ctypes wrapper of oracle/jpeg_ref.c (baseline JPEG restatement of the
'jpg' export, utils/canvas.js:205-242).  Only tests/, smoke() and bench.py's CPU legs may import this."""
from __future__ import annotations

import ctypes as C

import numpy as np

from oracle.raster_c import load as _load


def _lib():
    lib = _load()
    if not getattr(lib, "_jpeg_bound", False):
        lib.jpeg_ref_encode_i420.restype = C.c_int
        lib.jpeg_ref_encode_i420.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        lib.jpeg_ref_header.restype = C.c_size_t
        lib.jpeg_ref_header.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        lib.jpeg_ref_quant_table.restype = None
        lib.jpeg_ref_quant_table.argtypes = [C.c_int, C.c_int, C.c_void_p]
        lib.jpeg_ref_huff_tables.restype = None
        lib.jpeg_ref_huff_tables.argtypes = [C.c_void_p, C.c_void_p]
        lib.jpeg_ref_block.restype = None
        lib.jpeg_ref_block.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        lib.jpeg_ref_info.restype = C.c_int
        lib.jpeg_ref_info.argtypes = [C.c_void_p, C.c_size_t] + [C.POINTER(C.c_int)] * 5
        lib.jpeg_ref_decode.restype = C.c_int
        lib.jpeg_ref_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64]
        lib.jpeg_ref_coefficients.restype = C.c_int
        lib.jpeg_ref_coefficients.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        lib._jpeg_bound = True
    return lib


def encode_i420(y: np.ndarray, cb: np.ndarray, cr: np.ndarray, quality: int = 100, restart_interval: int = 8) -> bytes:
    """I420 planes -> JFIF bytes (sequential CPU encoder)."""
    h, w = y.shape
    assert cb.shape == ((h + 1) // 2, (w + 1) // 2) and cr.shape == cb.shape
    y, cb, cr = (np.ascontiguousarray(a, dtype=np.uint8) for a in (y, cb, cr))
    cap = 1024 + ((w + 15) // 16) * ((h + 15) // 16) * 6 * 440
    out = np.empty(cap, np.uint8)
    n = C.c_size_t()
    rc = _lib().jpeg_ref_encode_i420(y.ctypes.data, cb.ctypes.data, cr.ctypes.data, w, h, quality, restart_interval,
                                     out.ctypes.data, cap, C.byref(n))
    assert rc == 0, "oracle JPEG buffer too small"
    return out[: n.value].tobytes()


def quant_table(chroma: int, quality: int) -> np.ndarray:
    q = np.empty(64, np.uint8)
    _lib().jpeg_ref_quant_table(chroma, quality, q.ctypes.data)
    return q


def huff_tables():
    code = np.empty((4, 256), np.uint16)
    length = np.empty((4, 256), np.uint8)
    _lib().jpeg_ref_huff_tables(code.ctypes.data, length.ctypes.data)
    return code, length


def info(data: bytes):
    """(width, height, components, luma h-sampling, luma v-sampling) of a baseline JPEG, or raises ValueError."""
    lib = _lib()
    v = [C.c_int() for _ in range(5)]
    rc = lib.jpeg_ref_info(data, len(data), *[C.byref(x) for x in v])
    if rc:
        raise ValueError(f"oracle JPEG decoder: unsupported or corrupt file ({rc})")
    return tuple(x.value for x in v)


def decode(data: bytes) -> np.ndarray:
    """oracle/jpeg_dec_ref.c: RGBA8 [h, w, 4] bitmap (alpha 255) of a baseline JPEG file."""
    lib = _lib()
    w, h, *_ = info(data)
    out = np.empty((h, w, 4), np.uint8)
    rc = lib.jpeg_ref_decode(data, len(data), out.ctypes.data, 0)
    if rc:
        raise ValueError(f"oracle JPEG decoder failed ({rc})")
    return out


def coefficients(data: bytes):
    """Quantised coefficients per component: int16 [blocks_y, blocks_x, 64] over the MCU-padded planes."""
    lib = _lib()
    w, h, nc, hs, vs = info(data)
    mcux, mcuy = -(-w // (8 * hs)), -(-h // (8 * vs))
    shapes = [(mcuy * vs, mcux * hs, 64)] + [(mcuy, mcux, 64)] * (nc - 1)
    arrs = [np.zeros(s, np.int16) for s in shapes]
    ptrs = [a.ctypes.data for a in arrs] + [None] * (3 - nc)
    rc = lib.jpeg_ref_coefficients(data, len(data), *ptrs)
    if rc:
        raise ValueError(f"oracle JPEG decoder failed ({rc})")
    return arrs
