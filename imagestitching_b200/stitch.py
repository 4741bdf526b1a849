"""Host-side mirror of the reference's stitch interface, on top of the C ABI.

The reference (a WeChat Page, pages/index/index.js under
/root/reference/miniprogram-stitch/miniprogram/) exposes the stitch path as page
state + one method:
    in : data.images[] ({naturalWidth,naturalHeight,width,height,orientation,fileSize,...},
         index.js:724-739), data.direction ('vertical'|'horizontal', :16),
         data.gap (:17), data.verticalStitchMode / horizontalStitchMode (:19-20),
         platform + deviceMaxCanvasSize / deviceMaxCanvasPixels (:126-156)
    out: data.stitchedTempPath, data.stitchProgress, data.isStitching (:1611)
    fn : onStitchVertical / onStitchHorizontal -> onStitch (:771-788, :1186-1633)
``StitchPage`` keeps those names, argument meanings and error behaviour;
``stitch()`` / ``compute_layout()`` are the functional form the TypeScript facade
(imagestitching_b200/ts/stitch.ts) also exposes.  Pixels come only from libstitch.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

from . import _lib as L

_DIRECTIONS = {"vertical": L.VERTICAL, "horizontal": L.HORIZONTAL}
_MODES = {"min": L.MODE_MIN, "max": L.MODE_MAX, "original": L.MODE_ORIGINAL}
_PLATFORMS = {"ios": L.IOS, "android": L.ANDROID}

# reference Error messages by status (index.js:1254,1320,1412,1513)
_MESSAGES = {
    L.E_BAD_DIMS: "图片尺寸不可用",
    L.E_BAD_OUT: "输出尺寸计算失败",
    L.E_NO_SURFACE: "无法获取绘图上下文",
}


class StitchError(Exception):
    def __init__(self, status: int, message: str):
        super().__init__(message)
        self.status = status
        self.message = message


@dataclass
class DeviceBitmap:
    """An RGBA8 bitmap already resident in GPU memory (e.g. a torch tensor's data_ptr())."""
    ptr: int
    width: int
    height: int
    pitch: int = 0
    keepalive: Any = None


@dataclass
class StitchOutput:
    width: int
    height: int
    layout: Dict[str, Any]
    rgba: Any = None          # numpy [h,w,4] (host) or device pointer int
    y: Any = None
    cb: Any = None
    cr: Any = None
    jpeg: Optional[bytes] = None   # the exported JFIF file (output "...+jpeg" / "jpeg")
    png: Optional[bytes] = None    # the exported PNG file (output "...+png" / "png"), what the page asks for
    kernel_launches: int = 0
    path_info: int = 0
    progress: List[int] = field(default_factory=list)
    device: int = -1               # StitchPool: index of the worker / GPU that ran the job


def _output_flags(output: str) -> int:
    flags = 0
    for part in output.split("+"):
        flags |= {"rgba": L.OUT_RGBA, "yuv420": L.OUT_YUV420, "both": L.OUT_RGBA | L.OUT_YUV420, "jpeg": L.OUT_JPEG,
                  "png": L.OUT_PNG}[part]
    return flags


def _profile_struct(platform: str = "ios", deviceMaxCanvasSize: Optional[int] = None,
                    deviceMaxCanvasPixels: Optional[int] = None, fallback: bool = True) -> L.StitchProfile:
    """Profile as the page holds it after onLoad (index.js:342): stored values or the fallback."""
    lib = L.load()
    p = L.StitchProfile()
    plat = _PLATFORMS.get(platform, L.OTHER)
    if fallback:
        lib.stitch_profile_preset(plat, C.byref(p))
    p.platform = plat
    if deviceMaxCanvasSize:
        p.max_canvas_size = int(deviceMaxCanvasSize)
        if not deviceMaxCanvasPixels and fallback:
            # applyFallback() derives pixels from the stored size (index.js:131-137)
            side = p.max_canvas_size
            cap = side * min(side, 4096) if plat == L.ANDROID else side * min(side, 12288)
            p.max_canvas_pixels = max(cap, 4096 * 2048)
    if deviceMaxCanvasPixels:
        p.max_canvas_pixels = int(deviceMaxCanvasPixels)
    return p


def _orientation_value(o: Any) -> int:
    # utils/canvas.js:155-199: strict === against numbers 1..8; strings like 'up' draw plain
    if isinstance(o, bool) or not isinstance(o, (int, float)):
        return 0
    return int(o) if float(o).is_integer() and 1 <= o <= 8 else 0


def _desc(img: Dict[str, Any]) -> L.StitchImageDesc:
    d = L.StitchImageDesc()
    d.natural_width = int(img.get("naturalWidth") or 0)
    d.natural_height = int(img.get("naturalHeight") or 0)
    d.width = int(img.get("width") or 0)
    d.height = int(img.get("height") or 0)
    d.orientation = _orientation_value(img.get("orientation"))
    d.file_size = int(img.get("fileSize") or 0)
    return d


def _params(direction: str, mode: Optional[str], gap, profile: L.StitchProfile) -> L.StitchParams:
    p = L.StitchParams()
    if direction not in _DIRECTIONS:
        # index.js:1261: anything that is not 'vertical' takes the horizontal branch
        p.direction = L.HORIZONTAL
    else:
        p.direction = _DIRECTIONS[direction]
    mode = mode or "min"                       # index.js:1257
    p.mode = _MODES.get(mode, L.MODE_OTHER)
    try:
        p.gap = float(gap or 0)                # index.js:1256
    except (TypeError, ValueError):
        p.gap = 0.0
    p.profile = profile
    return p


def _layout_dict(lay: L.StitchLayout) -> Dict[str, Any]:
    n = lay.n
    return {
        "n": n, "bigTask": bool(lay.big_task), "maxSide": lay.max_side, "maxPixels": lay.max_pixels,
        "preW": lay.pre_w, "preH": lay.pre_h, "scaleDown": lay.scale_down,
        "targetW": lay.target_w, "targetH": lay.target_h, "superSample": lay.super_sample,
        "canvasW": lay.canvas_w, "canvasH": lay.canvas_h, "scaledGap": lay.scaled_gap,
        "rects": [[lay.rects[i].dx, lay.rects[i].dy, lay.rects[i].dw, lay.rects[i].dh] for i in range(n)],
        "natural": [[lay.natural_w[i], lay.natural_h[i]] for i in range(n)],
    }


def _raise(status: int, ctx=None):
    lib = L.load()
    msg = ""
    if ctx:
        msg = (lib.stitch_last_error(ctx) or b"").decode("utf-8", "replace")
    if not msg:
        msg = _MESSAGES.get(status) or (lib.stitch_strerror(status) or b"").decode("utf-8", "replace")
    raise StitchError(status, msg)


def compute_layout(images: Sequence[Dict[str, Any]], direction: str, mode: Optional[str] = "min", gap=0,
                   platform: str = "ios", deviceMaxCanvasSize: Optional[int] = None,
                   deviceMaxCanvasPixels: Optional[int] = None) -> Dict[str, Any]:
    """Exact twin of index.js:1251-1386 + 1522-1554 (CPU only; no GPU needed)."""
    lib = L.load()
    n = len(images)
    descs = (L.StitchImageDesc * max(n, 1))(*[_desc(i) for i in images])
    prm = _params(direction, mode, gap, _profile_struct(platform, deviceMaxCanvasSize, deviceMaxCanvasPixels))
    lay = L.StitchLayout()
    st = lib.stitch_layout(descs, n, C.byref(prm), C.byref(lay))
    if st != L.OK:
        _raise(st)
    d = _layout_dict(lay)
    seq = (C.c_int32 * 32)()
    k = lib.stitch_progress_sequence(n, seq, 32)
    d["progress"] = [seq[i] for i in range(k)]
    return d


def get_image_info(file_bytes: bytes) -> Dict[str, Any]:
    """wx.getImageInfo for an in-memory file: {width, height, type, orientation} with WeChat's orientation string,
    plus the numeric EXIF value as orientationExif."""
    lib = L.load()
    info = L.StitchImageInfo()
    buf = (C.c_char * max(1, len(file_bytes))).from_buffer_copy(file_bytes or b"\0")
    st = lib.stitch_probe_image(buf, len(file_bytes), C.byref(info))
    if st != L.OK:
        _raise(st)
    return {"width": info.width, "height": info.height, "type": lib.stitch_type_name(info.type).decode(),
            "orientation": lib.stitch_orientation_name(info.orientation_exif).decode(),
            "orientationExif": info.orientation_exif}


def prepare_image(file_bytes: bytes, **extra) -> Dict[str, Any]:
    """_prepareSingleImage (index.js:678-741) for an in-memory file: the prepared record the stitch path reads."""
    lib = L.load()
    info = L.StitchImageInfo()
    buf = (C.c_char * max(1, len(file_bytes))).from_buffer_copy(file_bytes or b"\0")
    st = lib.stitch_probe_image(buf, len(file_bytes), C.byref(info))
    if st != L.OK:
        _raise(st)
    d = L.StitchImageDesc()
    lib.stitch_prepare_record(C.byref(info), len(file_bytes), C.byref(d))
    rec = dict(extra)
    rec.update({"naturalWidth": d.natural_width, "naturalHeight": d.natural_height, "width": d.width, "height": d.height,
                "orientation": lib.stitch_orientation_name(info.orientation_exif).decode(), "fileSize": int(d.file_size),
                "preparedType": lib.stitch_type_name(info.type).decode(), "prepared": True})
    return rec


def preview_layout(export_w: int, export_h: int, view_w: float, view_h: float, dpr=1) -> Dict[str, Any]:
    """createHighResCanvas + the fit-scale of index.js:1599-1603 (pure CPU, no GPU needed)."""
    pl = L.StitchPreviewLayout()
    st = L.load().stitch_preview_layout(int(export_w), int(export_h), float(view_w), float(view_h), float(dpr or 0), C.byref(pl))
    if st != L.OK:
        _raise(st)
    return {"canvasW": pl.canvas_w, "canvasH": pl.canvas_h, "scaleFit": pl.scale_fit, "pvW": pl.pv_w, "pvH": pl.pv_h,
            "dx": pl.dx, "dy": pl.dy, "dpr": pl.dpr}


class StitchContext:
    """One per GPU: owns a StitchCtx (stream, staging).  Not thread-safe."""

    def __init__(self, device: int = 0, stream: int = 0):
        self.lib = L.load()
        h = C.c_void_p()
        st = self.lib.stitch_ctx_create(device, C.c_void_p(stream or None), C.byref(h))
        if st != L.OK:
            raise StitchError(st, "stitch_ctx_create failed (no usable CUDA device; there is no CPU fallback)")
        self.handle = h
        self.device = device
        self._cb = None

    def close(self):
        if self.handle:
            self.lib.stitch_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self) -> int:
        return int(self.lib.stitch_ctx_stream(self.handle) or 0)

    def set_lanes(self, n: int) -> None:
        """Number of job lanes stitch_submit rotates over (include/stitch.h: stitch_ctx_set_lanes)."""
        st = self.lib.stitch_ctx_set_lanes(self.handle, int(n))
        if st != L.OK:
            raise StitchError(st, (self.lib.stitch_last_error(self.handle) or b"").decode("utf-8", "replace"))

    def join(self) -> None:
        """Make the ctx stream wait (device-side) for everything submitted on the other lanes."""
        st = self.lib.stitch_ctx_join(self.handle)
        if st != L.OK:
            raise StitchError(st, (self.lib.stitch_last_error(self.handle) or b"").decode("utf-8", "replace"))

    @property
    def kernel_launches(self) -> int:
        return int(self.lib.stitch_kernel_launch_count(self.handle))

    # -- job construction ----------------------------------------------------
    def preview(self, exported, view_w: float, view_h: float, dpr=1) -> np.ndarray:
        """index.js:1593-1603: the exported canvas (uint8 [h, w, 4] array or DeviceBitmap) drawn fit-scaled and
        centred on the cleared preview canvas.  Returns straight-alpha RGBA8 [canvasH, canvasW, 4]."""
        lib = L.load()
        if isinstance(exported, DeviceBitmap):
            ptr, kind, w, h, pitch = exported.ptr, L.MEM_DEVICE, exported.width, exported.height, exported.pitch
        else:
            exported = np.ascontiguousarray(exported, dtype=np.uint8)
            ptr, kind, (h, w), pitch = exported.ctypes.data, L.MEM_HOST, exported.shape[:2], exported.strides[0]
        pl = L.StitchPreviewLayout()
        st = lib.stitch_preview_layout(int(w), int(h), float(view_w), float(view_h), float(dpr or 0), C.byref(pl))
        if st != L.OK:
            _raise(st, self.handle)
        out = np.empty((pl.canvas_h, pl.canvas_w, 4), dtype=np.uint8)
        st = lib.stitch_preview(self.handle, ptr, kind, int(w), int(h), int(pitch), float(view_w), float(view_h),
                                float(dpr or 0), out.ctypes.data, out.nbytes, L.MEM_HOST, C.byref(pl))
        if st != L.OK:
            _raise(st, self.handle)
        return out

    def decode_jpeg(self, file_bytes: bytes) -> np.ndarray:
        """loadImageFrom (utils/canvas.js:27-121) for a JPEG file (baseline or progressive): RGBA8 [h, w, 4], decoded on the GPU."""
        info = get_image_info(file_bytes)
        out = np.empty((info["height"], info["width"], 4), dtype=np.uint8)
        buf = (C.c_char * max(1, len(file_bytes))).from_buffer_copy(file_bytes or b"\0")
        st = self.lib.stitch_decode_jpeg(self.handle, buf, len(file_bytes), out.ctypes.data, out.nbytes, 0, L.MEM_HOST, None)
        if st != L.OK:
            _raise(st, self.handle)
        return out

    def decode_image(self, file_bytes: bytes) -> np.ndarray:
        """loadImageFrom (utils/canvas.js:27-121) for a file of any decodable type (JPEG, PNG, BMP, GIF: by its signature)."""
        info = get_image_info(file_bytes)
        out = np.empty((max(1, info["height"]), max(1, info["width"]), 4), dtype=np.uint8)
        buf = (C.c_char * max(1, len(file_bytes))).from_buffer_copy(file_bytes or b"\0")
        st = self.lib.stitch_decode_image(self.handle, buf, len(file_bytes), out.ctypes.data, out.nbytes, 0, L.MEM_HOST, None)
        if st != L.OK:
            _raise(st, self.handle)
        return out

    def decode_gif(self, file_bytes: bytes) -> np.ndarray:
        """loadImageFrom for a GIF file: RGBA8 [h, w, 4] of its first frame (gif_decode.cuh)."""
        info = get_image_info(file_bytes)
        out = np.empty((info["height"], info["width"], 4), dtype=np.uint8)
        buf = (C.c_char * max(1, len(file_bytes))).from_buffer_copy(file_bytes or b"\0")
        st = self.lib.stitch_decode_gif(self.handle, buf, len(file_bytes), out.ctypes.data, out.nbytes, 0, L.MEM_HOST, None)
        if st != L.OK:
            _raise(st, self.handle)
        return out

    def decode_bmp(self, file_bytes: bytes) -> np.ndarray:
        """loadImageFrom for a BMP file: RGBA8 [h, w, 4] (bmp_decode.cuh)."""
        info = get_image_info(file_bytes)
        out = np.empty((info["height"], info["width"], 4), dtype=np.uint8)
        buf = (C.c_char * max(1, len(file_bytes))).from_buffer_copy(file_bytes or b"\0")
        st = self.lib.stitch_decode_bmp(self.handle, buf, len(file_bytes), out.ctypes.data, out.nbytes, 0, L.MEM_HOST, None)
        if st != L.OK:
            _raise(st, self.handle)
        return out

    def decode_png(self, file_bytes: bytes) -> np.ndarray:
        """loadImageFrom (utils/canvas.js:27-121) for a PNG file: RGBA8 [h, w, 4] with straight alpha, inflated,
        unfiltered and converted on the GPU (png_decode.cu)."""
        info = get_image_info(file_bytes)
        out = np.empty((info["height"], info["width"], 4), dtype=np.uint8)
        buf = (C.c_char * max(1, len(file_bytes))).from_buffer_copy(file_bytes or b"\0")
        st = self.lib.stitch_decode_png(self.handle, buf, len(file_bytes), out.ctypes.data, out.nbytes, 0, L.MEM_HOST, None)
        if st != L.OK:
            _raise(st, self.handle)
        return out

    def build_job(self, images: Sequence[Dict[str, Any]], direction: str, mode: Optional[str], gap,
                  profile: L.StitchProfile, output: str, out_device: bool = False, out_rgba_ptr: int = 0,
                  out_yuv_ptr: int = 0, out_rgba_capacity: int = 0, out_yuv_capacity: int = 0, out_jpeg_ptr: int = 0,
                  out_jpeg_capacity: int = 0, jpeg_quality: int = 0, jpeg_restart_interval: int = 0):
        n = len(images)
        arr = (L.StitchImage * max(n, 1))()
        keep = []
        for i, img in enumerate(images):
            arr[i].desc = _desc(img)
            bmp = img.get("bitmap")
            if bmp is None and img.get("file") is not None:      # a JPEG / PNG FILE: decoded on the GPU (loadImageFrom)
                data = img["file"] if isinstance(img["file"], bytes) else bytes(img["file"])
                arr[i].data = C.cast(C.c_char_p(data), C.c_void_p)      # no copy: `data` is kept alive below
                arr[i].pitch_bytes = len(data)
                arr[i].mem_kind = L.MEM_HOST_FILE      # the library looks at the signature: JPEG, PNG, BMP or GIF
                keep.append(data)
            elif isinstance(bmp, DeviceBitmap):
                arr[i].data = bmp.ptr
                arr[i].bmp_width, arr[i].bmp_height = bmp.width, bmp.height
                arr[i].pitch_bytes = bmp.pitch
                arr[i].mem_kind = L.MEM_DEVICE
                keep.append(bmp)
            elif bmp is not None:
                a = np.asarray(bmp)
                if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 4:
                    raise StitchError(L.E_BAD_ARG, f"bitmap {i} must be uint8 [h,w,4] RGBA")
                if a.strides[2] != 1 or a.strides[1] != 4:
                    a = np.ascontiguousarray(a)
                arr[i].data = a.ctypes.data if a.size else None
                arr[i].bmp_height, arr[i].bmp_width = a.shape[0], a.shape[1]
                arr[i].pitch_bytes = a.strides[0] if a.shape[0] > 1 else a.shape[1] * 4
                arr[i].mem_kind = L.MEM_HOST
                keep.append(a)
        job = L.StitchJob()
        job.images = arr
        job.n = n
        job.params = _params(direction, mode, gap, profile)
        job.output_flags = _output_flags(output)
        job.out_mem_kind = L.MEM_DEVICE if out_device else L.MEM_HOST
        job.out_rgba = out_rgba_ptr or None
        job.out_yuv = out_yuv_ptr or None
        job.out_capacity_rgba = out_rgba_capacity
        job.out_capacity_yuv = out_yuv_capacity
        job.out_jpeg = out_jpeg_ptr or None
        job.out_capacity_jpeg = out_jpeg_capacity
        job.jpeg_quality = jpeg_quality
        job.jpeg_restart_interval = jpeg_restart_interval
        return job, (arr, keep)

    def _prepare(self, images, direction, mode, gap, platform, deviceMaxCanvasSize, deviceMaxCanvasPixels, output,
                 jpeg_quality, jpeg_restart_interval, jpeg_capacity, png_capacity):
        """Layout twin -> host output buffers -> StitchJob.  Everything a job needs to stay alive is in the returned dict."""
        lib = self.lib
        profile = _profile_struct(platform, deviceMaxCanvasSize, deviceMaxCanvasPixels)
        descs = (L.StitchImageDesc * len(images))(*[_desc(i) for i in images])
        for d, img in zip(descs, images):   # a FILE without a prepared record: the library reads its header (P:1227-1231)
            if img.get("file") is not None and not (d.natural_width and d.natural_height):
                info = get_image_info(img["file"] if isinstance(img["file"], bytes) else bytes(img["file"]))
                d.natural_width = d.width = info["width"]
                d.natural_height = d.height = info["height"]
                d.file_size = d.file_size or len(img["file"])
        prm = _params(direction, mode, gap, profile)
        lay = L.StitchLayout()
        st = lib.stitch_layout(descs, len(images), C.byref(prm), C.byref(lay))
        if st != L.OK:
            _raise(st)
        rb, yb = C.c_int64(), C.c_int64()
        lib.stitch_output_size(C.byref(lay), C.byref(rb), C.byref(yb))
        cw, ch = int(lay.canvas_w), int(lay.canvas_h)
        flags = _output_flags(output)
        rgba = np.empty((ch, cw, 4), dtype=np.uint8) if flags & L.OUT_RGBA else None
        yuv = np.empty(yb.value, dtype=np.uint8) if flags & L.OUT_YUV420 else None
        jpg = None
        if flags & L.OUT_JPEG:
            jpg = np.empty(jpeg_capacity or int(lib.stitch_jpeg_capacity(C.byref(lay), jpeg_restart_interval)), dtype=np.uint8)
        png = np.empty(png_capacity or int(lib.stitch_png_capacity(C.byref(lay))), dtype=np.uint8) if flags & L.OUT_PNG else None
        job, keep = self.build_job(images, direction, mode, gap, profile, output, False,
                                   rgba.ctypes.data if rgba is not None else 0,
                                   yuv.ctypes.data if yuv is not None else 0,
                                   rgba.nbytes if rgba is not None else 0, yuv.nbytes if yuv is not None else 0,
                                   jpg.ctypes.data if jpg is not None else 0, jpg.nbytes if jpg is not None else 0,
                                   jpeg_quality, jpeg_restart_interval)
        if png is not None:
            job.out_png, job.out_capacity_png = png.ctypes.data, png.nbytes
        return {"job": job, "keep": keep, "res": L.StitchResult(), "cw": cw, "ch": ch, "rgba": rgba, "yuv": yuv, "jpg": jpg,
                "png": png, "progress": []}

    def _collect(self, pj, st) -> StitchOutput:
        res, cw, ch = pj["res"], pj["cw"], pj["ch"]
        if st != L.OK:
            if st == L.E_NO_SURFACE and res.jpeg_bytes:
                raise StitchError(st, f"JPEG output buffer too small: {res.jpeg_bytes} bytes needed")
            if st == L.E_NO_SURFACE and res.png_bytes:
                raise StitchError(st, f"PNG output buffer too small: {res.png_bytes} bytes needed")
            _raise(st, self.handle)
        out = StitchOutput(cw, ch, _layout_dict(res.layout), kernel_launches=res.kernel_launches, path_info=res.path_info,
                           progress=pj["progress"])
        out.rgba = pj["rgba"]
        if pj["jpg"] is not None:
            out.jpeg = pj["jpg"][: res.jpeg_bytes].tobytes()
        if pj["png"] is not None:
            out.png = pj["png"][: res.png_bytes].tobytes()
        yuv = pj["yuv"]
        if yuv is not None:
            cwh, chh = (cw + 1) // 2, (ch + 1) // 2
            out.y = yuv[: cw * ch].reshape(ch, cw)
            out.cb = yuv[cw * ch: cw * ch + cwh * chh].reshape(chh, cwh)
            out.cr = yuv[cw * ch + cwh * chh:].reshape(chh, cwh)
        return out

    def run(self, images: Sequence[Dict[str, Any]], direction: str, *, mode: Optional[str] = "min", gap=0,
            platform: str = "ios", deviceMaxCanvasSize: Optional[int] = None,
            deviceMaxCanvasPixels: Optional[int] = None, output: str = "rgba",
            on_progress: Optional[Callable[[int], None]] = None, jpeg_quality: int = 0,
            jpeg_restart_interval: int = 0, jpeg_capacity: int = 0, png_capacity: int = 0) -> Optional[StitchOutput]:
        """Synchronous stitch with host outputs (numpy).  Returns None for an empty list (index.js:1189).
        `output`: "rgba" | "yuv420" | "both", optionally "+jpeg" / "+png", or "jpeg" / "png" alone (the exported file)."""
        lib = self.lib
        if not images:
            return None
        pj = self._prepare(images, direction, mode, gap, platform, deviceMaxCanvasSize, deviceMaxCanvasPixels, output,
                           jpeg_quality, jpeg_restart_interval, jpeg_capacity, png_capacity)
        progress = pj["progress"]

        def _cb(_user, pct):
            progress.append(int(pct))
            if on_progress:
                on_progress(int(pct))

        cb = L.PROGRESS_FN(_cb)
        lib.stitch_set_progress_cb(self.handle, cb, None)
        try:
            st = lib.stitch_run(self.handle, C.byref(pj["job"]), C.byref(pj["res"]))
        finally:
            lib.stitch_set_progress_cb(self.handle, C.cast(None, L.PROGRESS_FN), None)
        return self._collect(pj, st)

    def submit(self, images: Sequence[Dict[str, Any]], direction: str, *, mode: Optional[str] = "min", gap=0,
               platform: str = "ios", deviceMaxCanvasSize: Optional[int] = None, deviceMaxCanvasPixels: Optional[int] = None,
               output: str = "rgba", jpeg_quality: int = 0, jpeg_restart_interval: int = 0, jpeg_capacity: int = 0,
               png_capacity: int = 0):
        """Asynchronous stitch (stitch_submit): the job is enqueued on the next lane and the call returns without waiting
        for the GPU -- also for file inputs and file outputs.  Pass the returned handle to wait()."""
        pj = self._prepare(images, direction, mode, gap, platform, deviceMaxCanvasSize, deviceMaxCanvasPixels, output,
                           jpeg_quality, jpeg_restart_interval, jpeg_capacity, png_capacity)
        ticket = C.c_uint64(0)
        st = self.lib.stitch_submit(self.handle, C.byref(pj["job"]), C.byref(pj["res"]), C.byref(ticket))
        if st != L.OK:
            self._collect(pj, st)
        pj["ticket"] = ticket.value
        return pj

    def wait(self, pj) -> StitchOutput:
        """Completes a submitted job (and every job submitted before it) and returns its outputs."""
        return self._collect(pj, self.lib.stitch_wait(self.handle, pj["ticket"]))


class StitchPool:
    """Every GPU of the box from one process (include/stitch.h: stitch_pool_*): one worker thread + ctx per GPU, a
    shared job queue, no collective.  Jobs with host buffers go to whichever GPU has a free lane."""

    def __init__(self, devices: Optional[Sequence[int]] = None, lanes_per_device: int = 2):
        self.lib = L.load()
        h = C.c_void_p()
        arr = (C.c_int32 * len(devices))(*devices) if devices else None
        st = self.lib.stitch_pool_create(arr, len(devices) if devices else 0, lanes_per_device, C.byref(h))
        if st != L.OK:
            _raise(st)
        self.handle = h
        self._jobs: Dict[int, Any] = {}
        self._helper = StitchContext.__new__(StitchContext)   # build_job / _prepare / _collect need only .lib
        self._helper.lib = self.lib
        self._helper.handle = None

    @property
    def device_count(self) -> int:
        return int(self.lib.stitch_pool_device_count(self.handle))

    def ctx_handle(self, index: int):
        return self.lib.stitch_pool_ctx(self.handle, index)

    def jobs_run(self, index: int) -> int:
        return int(self.lib.stitch_pool_jobs_run(self.handle, index))

    def submit(self, images: Sequence[Dict[str, Any]], direction: str, *, device: int = -1, mode: Optional[str] = "min",
               gap=0, platform: str = "ios", deviceMaxCanvasSize: Optional[int] = None,
               deviceMaxCanvasPixels: Optional[int] = None, output: str = "rgba", jpeg_quality: int = 0,
               jpeg_restart_interval: int = 0, jpeg_capacity: int = 0, png_capacity: int = 0):
        pj = self._helper._prepare(images, direction, mode, gap, platform, deviceMaxCanvasSize, deviceMaxCanvasPixels,
                                   output, jpeg_quality, jpeg_restart_interval, jpeg_capacity, png_capacity)
        t = C.c_uint64(0)
        st = self.lib.stitch_pool_submit(self.handle, C.byref(pj["job"]), C.byref(pj["res"]), device, C.byref(t))
        if st != L.OK:
            _raise(st)
        pj["ticket"] = t.value
        return pj

    def wait(self, pj) -> StitchOutput:
        dev = C.c_int32(-1)
        st = self.lib.stitch_pool_wait(self.handle, pj["ticket"], C.byref(dev))
        if st != L.OK:
            msg = self.lib.stitch_pool_last_error(self.handle)
            raise StitchError(st, (msg or b"").decode("utf-8", "replace") or self.lib.stitch_strerror(st).decode())
        out = self._helper._collect(pj, st)
        out.device = int(dev.value)
        return out

    def close(self):
        if getattr(self, "handle", None):
            self.lib.stitch_pool_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default_ctx: Dict[int, StitchContext] = {}


def stitch(images: Sequence[Dict[str, Any]], direction: str, *, device: int = 0, **kw) -> Optional[StitchOutput]:
    """Functional entry: ordered image list + 'vertical'/'horizontal' in, exported buffer out."""
    ctx = _default_ctx.get(device)
    if ctx is None:
        ctx = _default_ctx[device] = StitchContext(device)
    return ctx.run(images, direction, **kw)


class StitchPage:
    """The stitch-relevant slice of the reference's Page object, same field and method names."""

    MAX_IMAGES = 9  # index.js:311

    def __init__(self, device: int = 0, platform: str = "ios", output: str = "rgba"):
        self.data: Dict[str, Any] = {
            "images": [], "direction": "vertical", "gap": 0,
            "verticalStitchMode": "min", "horizontalStitchMode": "min",
            "stitchedTempPath": "", "isStitching": False, "stitchProgress": 0,
        }
        self.platform = platform                     # wx.getSystemInfoSync().platform
        self.deviceMaxCanvasSize: Optional[int] = None
        self.deviceMaxCanvasPixels: Optional[int] = None
        self.output = output
        self.progressLog: List[int] = []
        self.lastToast = ""
        self.result: Optional[StitchOutput] = None
        self._device = device
        self._ctx: Optional[StitchContext] = None

    def setData(self, d: Dict[str, Any]):
        self.data.update(d)
        if "stitchProgress" in d:
            self.progressLog.append(d["stitchProgress"])

    def onStitchVertical(self):                      # index.js:771-779
        if self.data["isStitching"]:
            return
        if not self.data["images"]:
            self.lastToast = "请先添加图片"
            return
        self.setData({"direction": "vertical"})
        self.onStitch()

    def onStitchHorizontal(self):                    # index.js:780-788
        if self.data["isStitching"]:
            return
        if not self.data["images"]:
            self.lastToast = "请先添加图片"
            return
        self.setData({"direction": "horizontal"})
        self.onStitch()

    def onStitch(self):                              # index.js:1186-1633
        images = [dict(it) for it in (self.data["images"] or [])]
        if not images:
            return                                   # :1189
        direction, gap = self.data["direction"], self.data["gap"]
        self.progressLog = []
        self.data.update({"isStitching": True, "stitchedTempPath": ""})
        mode = (self.data["verticalStitchMode"] if direction == "vertical" else self.data["horizontalStitchMode"]) or "min"
        try:
            if self._ctx is None:
                self._ctx = StitchContext(self._device)
            out = self._ctx.run(images[: self.MAX_IMAGES] if len(images) <= self.MAX_IMAGES else images, direction,
                                mode=mode, gap=gap, platform=self.platform,
                                deviceMaxCanvasSize=self.deviceMaxCanvasSize,
                                deviceMaxCanvasPixels=self.deviceMaxCanvasPixels, output=self.output,
                                on_progress=lambda p: self.setData({"stitchProgress": p}) if p not in (100,) else None)
            self.result = out
            self.setData({"stitchedTempPath": f"mem://stitch/{id(out):x}", "stitchProgress": 100, "isStitching": False})  # :1611
        except StitchError as err:                   # :1618-1624
            self.lastToast = "拼图失败：" + err.message if err.message else "拼图失败"
            self.progressLog = [p for p in self.progressLog if p != 0]
            self.setData({"isStitching": False, "stitchProgress": 0})
