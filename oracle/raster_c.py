"""ctypes wrapper of the C oracle (oracle/raster_ref.c).  TEST INFRASTRUCTURE ONLY.

fp64 entry = fast checker for full-size parity tests; fp32 entry = the timed CPU
baseline (`bench.py` cpu_baseline / --impl reference).  Built by oracle/Makefile.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Sequence

import numpy as np

from .layout_ref import Layout, placement_for

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")


class OrImage(C.Structure):
    _fields_ = [("data", C.c_void_p), ("pitch", C.c_int64), ("bw", C.c_int32), ("bh", C.c_int32),
                ("X0", C.c_double), ("Y0", C.c_double), ("W", C.c_double), ("H", C.c_double),
                ("transpose", C.c_int32), ("flip_u", C.c_int32), ("flip_v", C.c_int32), ("pad", C.c_int32)]


_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB):
        subprocess.check_call(["make", "-C", HERE], stdout=subprocess.DEVNULL)
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        for name in ("oracle_render_f64", "oracle_render_f32"):
            fn = getattr(_lib, name)
            fn.restype = C.c_int
            fn.argtypes = [C.POINTER(OrImage), C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                           C.c_void_p, C.c_void_p]
        for name in ("oracle_preview_f64", "oracle_preview_f32"):
            fn = getattr(_lib, name)
            fn.restype = C.c_int
            fn.argtypes = [C.POINTER(OrImage), C.c_int, C.c_int, C.c_void_p, C.c_int64]
    return _lib


def set_threads(n: int) -> None:
    """OpenMP threads the C port uses from now on (bench.py: 1-thread and all-cores CPU baselines)."""
    load().oracle_set_threads(int(n))


def render_preview(bmp: np.ndarray, pl, precision: str = "f64") -> np.ndarray:
    """C twin of raster_ref.render_preview: straight-alpha RGBA8 [canvasH, canvasW, 4]."""
    from .layout_ref import preview_placement
    lib = load()
    a = np.ascontiguousarray(np.asarray(bmp))
    p = preview_placement(pl)
    im = OrImage()
    im.data, im.pitch, im.bh, im.bw = a.ctypes.data, a.strides[0], a.shape[0], a.shape[1]
    im.X0, im.Y0, im.W, im.H = p.X0, p.Y0, p.W, p.H
    out = np.empty((int(pl.canvasH), int(pl.canvasW), 4), np.uint8)
    fn = lib.oracle_preview_f64 if precision == "f64" else lib.oracle_preview_f32
    if out.size and fn(C.byref(im), int(pl.canvasW), int(pl.canvasH), out.ctypes.data, 0) != 0:
        raise MemoryError("oracle_preview failed")
    return out


def render(layout: Layout, bitmaps: Sequence[np.ndarray], orientations: Sequence[object], want_rgba: bool = True,
           want_yuv: bool = False, precision: str = "f64", return_float: bool = False):
    """Returns dict with 'rgba' [h,w,4], 'y','cb','cr' planes and optionally 'float' [h,w,3]."""
    lib = load()
    cw, ch = int(layout.canvasW), int(layout.canvasH)
    ss = layout.superSample if layout.superSample != 1 else 1.0
    n = len(bitmaps)
    arr = (OrImage * n)()
    keep = []
    for i, (rect, bmp, o) in enumerate(zip(layout.rects, bitmaps, orientations)):
        a = np.asarray(bmp)
        assert a.dtype == np.uint8 and a.ndim == 3 and a.shape[2] == 4
        if a.strides[2] != 1 or a.strides[1] != 4:
            a = np.ascontiguousarray(a)
        keep.append(a)
        p = placement_for(rect, o, ss)
        arr[i].data = a.ctypes.data
        arr[i].pitch = a.strides[0]
        arr[i].bh, arr[i].bw = a.shape[0], a.shape[1]
        arr[i].X0, arr[i].Y0, arr[i].W, arr[i].H = p.X0, p.Y0, p.W, p.H
        arr[i].transpose, arr[i].flip_u, arr[i].flip_v = int(p.transpose), int(p.flip_u), int(p.flip_v)
    out = {}
    rgba = np.empty((ch, cw, 4), np.uint8) if want_rgba else None
    cwh, chh = (cw + 1) // 2, (ch + 1) // 2
    y = np.empty((ch, cw), np.uint8) if want_yuv else None
    cb = np.empty((chh, cwh), np.uint8) if want_yuv else None
    cr = np.empty((chh, cwh), np.uint8) if want_yuv else None
    fl = None
    if return_float:
        fl = np.empty((ch, cw, 3), np.float64 if precision == "f64" else np.float32)
    fn = lib.oracle_render_f64 if precision == "f64" else lib.oracle_render_f32
    rc = fn(arr, n, cw, ch, rgba.ctypes.data if want_rgba else None, 0, y.ctypes.data if want_yuv else None,
            cb.ctypes.data if want_yuv else None, cr.ctypes.data if want_yuv else None,
            fl.ctypes.data if fl is not None else None)
    if rc != 0:
        raise MemoryError("oracle_render failed")
    out["rgba"], out["y"], out["cb"], out["cr"], out["float"] = rgba, y, cb, cr, fl
    return out
