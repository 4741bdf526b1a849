"""ctypes binding of libstitch.so (include/stitch.h).  No torch types cross this boundary.

The library is loaded from the package directory (built in-tree by
``__graft_entry__.build()`` / ``make -C imagestitching_b200/csrc``).  If it is
missing the import fails loudly: there is no Python / CPU fallback for pixels.
"""
from __future__ import annotations

import ctypes as C
import os

STITCH_MAX_IMAGES = 9

(OK, E_EMPTY, E_BAD_DIMS, E_BAD_OUT, E_NO_SURFACE, E_BAD_IMAGE, E_TOO_MANY, E_BAD_ARG, E_CUDA, E_BUSY) = range(10)
VERTICAL, HORIZONTAL = 0, 1
MODE_MIN, MODE_MAX, MODE_ORIGINAL, MODE_OTHER = 0, 1, 2, 3
IOS, ANDROID, OTHER = 0, 1, 2
MEM_HOST, MEM_DEVICE, MEM_HOST_JPEG, MEM_HOST_PNG, MEM_HOST_BMP, MEM_HOST_FILE, MEM_HOST_GIF = 0, 1, 2, 3, 4, 5, 6
OUT_RGBA, OUT_YUV420, OUT_JPEG, OUT_PNG = 1, 2, 4, 8
NOALLOC_DIGEST_MAX_BYTES = 1 << 30


class StitchProfile(C.Structure):
    _fields_ = [("platform", C.c_int32), ("max_canvas_size", C.c_int32), ("max_canvas_pixels", C.c_int64)]


class StitchImageDesc(C.Structure):
    _fields_ = [("natural_width", C.c_int32), ("natural_height", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("orientation", C.c_int32), ("reserved", C.c_int32), ("file_size", C.c_int64)]


class StitchParams(C.Structure):
    _fields_ = [("direction", C.c_int32), ("mode", C.c_int32), ("gap", C.c_double), ("profile", StitchProfile)]


class StitchRect(C.Structure):
    _fields_ = [("dx", C.c_double), ("dy", C.c_double), ("dw", C.c_double), ("dh", C.c_double)]


class StitchLayout(C.Structure):
    _fields_ = [("n", C.c_int32), ("big_task", C.c_int32), ("max_side", C.c_double), ("max_pixels", C.c_double),
                ("pre_w", C.c_double), ("pre_h", C.c_double), ("scale_down", C.c_double), ("target_w", C.c_double),
                ("target_h", C.c_double), ("super_sample", C.c_double), ("canvas_w", C.c_double), ("canvas_h", C.c_double),
                ("scaled_gap", C.c_double), ("rects", StitchRect * STITCH_MAX_IMAGES),
                ("natural_w", C.c_int32 * STITCH_MAX_IMAGES), ("natural_h", C.c_int32 * STITCH_MAX_IMAGES)]


class StitchImage(C.Structure):
    _fields_ = [("desc", StitchImageDesc), ("data", C.c_void_p), ("bmp_width", C.c_int32), ("bmp_height", C.c_int32),
                ("pitch_bytes", C.c_int64), ("mem_kind", C.c_int32), ("reserved", C.c_int32)]


class StitchJob(C.Structure):
    _fields_ = [("images", C.POINTER(StitchImage)), ("n", C.c_int32), ("params", StitchParams),
                ("output_flags", C.c_int32), ("out_mem_kind", C.c_int32), ("out_rgba", C.c_void_p),
                ("out_rgba_pitch", C.c_int64), ("out_yuv", C.c_void_p), ("out_capacity_rgba", C.c_int64),
                ("out_capacity_yuv", C.c_int64), ("out_jpeg", C.c_void_p), ("out_capacity_jpeg", C.c_int64),
                ("jpeg_quality", C.c_int32), ("jpeg_restart_interval", C.c_int32),
                ("out_png", C.c_void_p), ("out_capacity_png", C.c_int64)]


class StitchResult(C.Structure):
    _fields_ = [("layout", StitchLayout), ("canvas_w", C.c_int32), ("canvas_h", C.c_int32), ("rgba_bytes", C.c_int64),
                ("yuv_bytes", C.c_int64), ("kernel_launches", C.c_int32), ("path_info", C.c_int32),
                ("jpeg_bytes", C.c_int64), ("png_bytes", C.c_int64)]


class StitchImageInfo(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("type", C.c_int32), ("orientation_exif", C.c_int32)]


class StitchPreviewLayout(C.Structure):
    _fields_ = [("canvas_w", C.c_int32), ("canvas_h", C.c_int32), ("pv_w", C.c_int32), ("pv_h", C.c_int32),
                ("scale_fit", C.c_double), ("dx", C.c_double), ("dy", C.c_double), ("dpr", C.c_double)]


PROGRESS_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int32)

# name -> (restype, argtypes); every symbol include/stitch.h declares
SIGNATURES = {
    "stitch_profile_preset": (C.c_int, [C.c_int32, C.POINTER(StitchProfile)]),
    "stitch_layout": (C.c_int, [C.POINTER(StitchImageDesc), C.c_int32, C.POINTER(StitchParams), C.POINTER(StitchLayout)]),
    "stitch_progress_sequence": (C.c_int, [C.c_int32, C.POINTER(C.c_int32), C.c_int32]),
    "stitch_ctx_create": (C.c_int, [C.c_int32, C.c_void_p, C.POINTER(C.c_void_p)]),
    "stitch_ctx_destroy": (C.c_int, [C.c_void_p]),
    "stitch_ctx_stream": (C.c_void_p, [C.c_void_p]),
    "stitch_ctx_set_lanes": (C.c_int, [C.c_void_p, C.c_int32]),
    "stitch_ctx_join": (C.c_int, [C.c_void_p]),
    "stitch_set_progress_cb": (C.c_int, [C.c_void_p, PROGRESS_FN, C.c_void_p]),
    "stitch_output_size": (C.c_int, [C.POINTER(StitchLayout), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "stitch_jpeg_capacity": (C.c_int64, [C.POINTER(StitchLayout), C.c_int32]),
    "stitch_png_capacity": (C.c_int64, [C.POINTER(StitchLayout)]),
    "stitch_run": (C.c_int, [C.c_void_p, C.POINTER(StitchJob), C.POINTER(StitchResult)]),
    "stitch_submit": (C.c_int, [C.c_void_p, C.POINTER(StitchJob), C.POINTER(StitchResult), C.POINTER(C.c_uint64)]),
    "stitch_wait": (C.c_int, [C.c_void_p, C.c_uint64]),
    "stitch_probe_image": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(StitchImageInfo)]),
    "stitch_orientation_name": (C.c_char_p, [C.c_int32]),
    "stitch_type_name": (C.c_char_p, [C.c_int32]),
    "stitch_prepare_record": (C.c_int, [C.POINTER(StitchImageInfo), C.c_int64, C.POINTER(StitchImageDesc)]),
    "stitch_decode_jpeg": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                     C.POINTER(StitchImageInfo)]),
    "stitch_decode_png": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                    C.POINTER(StitchImageInfo)]),
    "stitch_decode_bmp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                    C.POINTER(StitchImageInfo)]),
    "stitch_decode_gif": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                    C.POINTER(StitchImageInfo)]),
    "stitch_decode_image": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                    C.POINTER(StitchImageInfo)]),
    "stitch_preview_layout": (C.c_int, [C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double,
                                        C.POINTER(StitchPreviewLayout)]),
    "stitch_preview": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_double, C.c_double,
                                 C.c_double, C.c_void_p, C.c_int64, C.c_int32, C.POINTER(StitchPreviewLayout)]),
    "stitch_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "stitch_host_free": (C.c_int, [C.c_void_p]),
    "stitch_device_alloc": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "stitch_device_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "stitch_memcpy_h2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "stitch_memcpy_d2h": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "stitch_generate_image": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int64, C.c_uint64, C.c_int32]),
    "stitch_device_digest": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)]),
    "stitch_device_digest_noalloc": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_char_p, C.c_size_t]),
    "stitch_pool_create": (C.c_int, [C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "stitch_pool_destroy": (C.c_int, [C.c_void_p]),
    "stitch_pool_device_count": (C.c_int32, [C.c_void_p]),
    "stitch_pool_ctx": (C.c_void_p, [C.c_void_p, C.c_int32]),
    "stitch_pool_submit": (C.c_int, [C.c_void_p, C.POINTER(StitchJob), C.POINTER(StitchResult), C.c_int32, C.POINTER(C.c_uint64)]),
    "stitch_pool_digest": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "stitch_pool_wait": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(C.c_int32)]),
    "stitch_pool_wait_all": (C.c_int, [C.c_void_p]),
    "stitch_pool_jobs_run": (C.c_int64, [C.c_void_p, C.c_int32]),
    "stitch_pool_last_error": (C.c_char_p, [C.c_void_p]),
    "stitch_strerror": (C.c_char_p, [C.c_int]),
    "stitch_last_error": (C.c_char_p, [C.c_void_p]),
    "stitch_abi_version": (C.c_int, []),
    "stitch_kernel_launch_count": (C.c_int64, [C.c_void_p]),
}

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # see stitch_ctx_create (must precede CUDA's initialisation)
LIB_PATH = os.environ.get("STITCH_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libstitch.so")
_lib = None


def load() -> C.CDLL:
    """Load libstitch.so and bind every declared symbol; raises if the build is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C imagestitching_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
