"""imagestitching_b200: B200-native stitch path of Iamctb/ImageStitching behind a C ABI.

Only what the path needs lives here: ``csrc/`` (CUDA kernels + the C-ABI library),
``_lib`` (ctypes binding), ``stitch`` (host-side mirror of the reference's page
interface) and ``ts/`` (TypeScript facade + N-API shim sources).
"""
from .stitch import (DeviceBitmap, StitchContext, StitchError, StitchOutput, StitchPage, StitchPool, compute_layout, get_image_info, prepare_image,
                     preview_layout,
                     stitch)

__all__ = ["DeviceBitmap", "StitchContext", "StitchError", "StitchOutput", "StitchPage", "StitchPool", "compute_layout", "get_image_info", "prepare_image", "preview_layout", "stitch"]
