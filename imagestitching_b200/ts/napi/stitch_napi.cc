// N-API shim between imagestitching_b200/ts/stitch.ts and libstitch.so (include/stitch.h).
// Stable C N-API only.  stitch_run() executes on a libuv worker (napi_async_work); progress values are
// marshalled to JS through a thread-safe function.
//
// NOT BUILT IN THIS REPO'S CI: the image has no node_api.h.  Build with node-gyp:
//   { "targets": [{ "target_name": "stitch_napi", "sources": ["stitch_napi.cc"],
//       "include_dirs": ["../../../include"], "libraries": ["-L../../..", "-l:libstitch.so"] }] }
#include <node_api.h>

#include <cstring>
#include <string>
#include <vector>

#include "stitch.h"

namespace {

struct RunData {
  StitchCtx* ctx = nullptr;
  std::vector<StitchImage> images;
  std::vector<napi_ref> keep;          // the Uint8Arrays stay alive until completion
  StitchJob job{};
  StitchResult result{};
  std::vector<uint8_t> rgba, yuv, jpeg, png;
  int status = 0;
  std::string message;
  napi_deferred deferred = nullptr;
  napi_async_work work = nullptr;
  napi_threadsafe_function tsfn = nullptr;
};

int32_t get_i32(napi_env env, napi_value obj, const char* key, int32_t dflt = 0) {
  napi_value v;
  bool has = false;
  napi_has_named_property(env, obj, key, &has);
  if (!has) return dflt;
  napi_get_named_property(env, obj, key, &v);
  double d = dflt;
  if (napi_get_value_double(env, v, &d) != napi_ok) return dflt;
  return (int32_t)d;
}
double get_f64(napi_env env, napi_value obj, const char* key, double dflt = 0) {
  napi_value v;
  bool has = false;
  napi_has_named_property(env, obj, key, &has);
  if (!has) return dflt;
  napi_get_named_property(env, obj, key, &v);
  double d = dflt;
  napi_get_value_double(env, v, &d);
  return d;
}

void progress_to_js(napi_env env, napi_value cb, void*, void* data) {
  if (!env || !cb) return;
  napi_value undef, pct;
  napi_get_undefined(env, &undef);
  napi_create_int32(env, (int32_t)(intptr_t)data, &pct);
  napi_call_function(env, undef, cb, 1, &pct, nullptr);
}
void on_progress(void* user, int32_t pct) {   // called on the worker thread inside stitch_run
  auto* d = static_cast<RunData*>(user);
  if (d->tsfn) napi_call_threadsafe_function(d->tsfn, (void*)(intptr_t)pct, napi_tsfn_nonblocking);
}

void execute(napi_env, void* p) {
  auto* d = static_cast<RunData*>(p);
  stitch_set_progress_cb(d->ctx, on_progress, d);
  d->status = stitch_run(d->ctx, &d->job, &d->result);
  if (d->status != STITCH_OK) d->message = stitch_last_error(d->ctx);
  stitch_set_progress_cb(d->ctx, nullptr, nullptr);
}

void complete(napi_env env, napi_status, void* p) {
  auto* d = static_cast<RunData*>(p);
  napi_value out, v;
  napi_create_object(env, &out);
  napi_create_int32(env, d->status, &v); napi_set_named_property(env, out, "status", v);
  napi_create_string_utf8(env, d->message.c_str(), NAPI_AUTO_LENGTH, &v); napi_set_named_property(env, out, "message", v);
  napi_create_int32(env, d->result.canvas_w, &v); napi_set_named_property(env, out, "width", v);
  napi_create_int32(env, d->result.canvas_h, &v); napi_set_named_property(env, out, "height", v);
  auto attach = [&](const char* key, std::vector<uint8_t>& buf) {
    if (buf.empty()) return;
    void* dst; napi_value ab, ta;
    napi_create_arraybuffer(env, buf.size(), &dst, &ab);
    std::memcpy(dst, buf.data(), buf.size());
    napi_create_typedarray(env, napi_uint8_array, buf.size(), ab, 0, &ta);
    napi_set_named_property(env, out, key, ta);
  };
  if (d->status == STITCH_OK) { attach("rgba", d->rgba); attach("yuv", d->yuv);
    d->jpeg.resize((size_t)d->result.jpeg_bytes);   // the file is the first jpeg_bytes of the capacity buffer
    attach("jpeg", d->jpeg);
    d->png.resize((size_t)d->result.png_bytes);
    attach("png", d->png); }
  for (napi_ref r : d->keep) napi_delete_reference(env, r);
  if (d->tsfn) napi_release_threadsafe_function(d->tsfn, napi_tsfn_release);
  napi_resolve_deferred(env, d->deferred, out);
  napi_delete_async_work(env, d->work);
  delete d;
}

napi_value CreateContext(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  int32_t device = 0;
  if (argc > 0) napi_get_value_int32(env, argv[0], &device);
  StitchCtx* ctx = nullptr;
  if (stitch_ctx_create(device, nullptr, &ctx) != STITCH_OK) {
    napi_throw_error(env, nullptr, "stitch_ctx_create failed (no CUDA device; there is no CPU fallback)");
    return nullptr;
  }
  napi_value ext;
  napi_create_external(env, ctx, nullptr, nullptr, &ext);
  return ext;
}

napi_value DestroyContext(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  void* p = nullptr;
  napi_get_value_external(env, argv[0], &p);
  if (p) stitch_ctx_destroy(static_cast<StitchCtx*>(p));
  return nullptr;
}

napi_value Run(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value argv[3];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  auto* d = new RunData();
  void* p = nullptr;
  napi_get_value_external(env, argv[0], &p);
  d->ctx = static_cast<StitchCtx*>(p);
  napi_value job = argv[1], imgs;
  napi_get_named_property(env, job, "images", &imgs);
  uint32_t n = 0;
  napi_get_array_length(env, imgs, &n);
  d->images.resize(n);
  for (uint32_t i = 0; i < n; ++i) {
    napi_value im, data;
    napi_get_element(env, imgs, i, &im);
    StitchImage& s = d->images[i];
    std::memset(&s, 0, sizeof s);
    s.desc.natural_width = get_i32(env, im, "naturalWidth");
    s.desc.natural_height = get_i32(env, im, "naturalHeight");
    s.desc.width = get_i32(env, im, "width");
    s.desc.height = get_i32(env, im, "height");
    s.desc.orientation = get_i32(env, im, "orientation");   // non-numeric ('up', 'right') -> 0 -> plain draw
    s.desc.file_size = (int64_t)get_f64(env, im, "fileSize");
    s.bmp_width = get_i32(env, im, "bmpWidth");
    s.bmp_height = get_i32(env, im, "bmpHeight");
    s.pitch_bytes = get_i32(env, im, "pitchBytes");
    s.mem_kind = STITCH_MEM_HOST;
    bool has_file = false;
    napi_has_named_property(env, im, "file", &has_file);
    napi_get_named_property(env, im, has_file ? "file" : "data", &data);
    napi_typedarray_type tt; size_t len = 0; void* ptr = nullptr; napi_value ab; size_t off = 0;
    if (napi_get_typedarray_info(env, data, &tt, &len, &ptr, &ab, &off) == napi_ok) {
      s.data = ptr;
      if (has_file) {   // the FILE (JPEG, PNG, BMP or GIF: the library looks at its signature): decoded on the GPU
        s.mem_kind = STITCH_MEM_HOST_FILE;
        s.pitch_bytes = (int64_t)len;
      }
      napi_ref ref; napi_create_reference(env, data, 1, &ref); d->keep.push_back(ref);
    }
  }
  d->job.images = d->images.data();
  d->job.n = (int32_t)n;
  d->job.params.direction = get_i32(env, job, "direction");
  d->job.params.mode = get_i32(env, job, "mode");
  d->job.params.gap = get_f64(env, job, "gap");
  d->job.params.profile.platform = get_i32(env, job, "platform");
  d->job.params.profile.max_canvas_size = get_i32(env, job, "maxCanvasSize");
  d->job.params.profile.max_canvas_pixels = (int64_t)get_f64(env, job, "maxCanvasPixels");
  d->job.output_flags = get_i32(env, job, "output", STITCH_OUT_RGBA);
  d->job.out_mem_kind = STITCH_MEM_HOST;
  // size the outputs with the layout twin
  std::vector<StitchImageDesc> descs(n);
  for (uint32_t i = 0; i < n; ++i) descs[i] = d->images[i].desc;
  StitchLayout lay{};
  int64_t rb = 0, yb = 0;
  if (stitch_layout(descs.data(), (int32_t)n, &d->job.params, &lay) == STITCH_OK) stitch_output_size(&lay, &rb, &yb);
  if (d->job.output_flags & STITCH_OUT_RGBA) d->rgba.resize((size_t)rb);
  if (d->job.output_flags & STITCH_OUT_YUV420) d->yuv.resize((size_t)yb);
  if (d->job.output_flags & STITCH_OUT_JPEG) {
    d->job.jpeg_quality = get_i32(env, job, "jpegQuality", 0);
    d->job.jpeg_restart_interval = get_i32(env, job, "jpegRestartInterval", 0);
    d->jpeg.resize((size_t)stitch_jpeg_capacity(&lay, d->job.jpeg_restart_interval));
    d->job.out_jpeg = d->jpeg.data();
    d->job.out_capacity_jpeg = (int64_t)d->jpeg.size();
  }
  if (d->job.output_flags & STITCH_OUT_PNG) {
    d->png.resize((size_t)stitch_png_capacity(&lay));
    d->job.out_png = d->png.data();
    d->job.out_capacity_png = (int64_t)d->png.size();
  }
  d->job.out_rgba = d->rgba.empty() ? nullptr : d->rgba.data();
  d->job.out_yuv = d->yuv.empty() ? nullptr : d->yuv.data();
  d->job.out_capacity_rgba = (int64_t)d->rgba.size();
  d->job.out_capacity_yuv = (int64_t)d->yuv.size();

  napi_value promise, name;
  napi_create_promise(env, &d->deferred, &promise);
  napi_create_string_utf8(env, "stitch_run", NAPI_AUTO_LENGTH, &name);
  napi_valuetype t = napi_undefined;
  if (argc > 2) napi_typeof(env, argv[2], &t);
  if (t == napi_function) napi_create_threadsafe_function(env, argv[2], nullptr, name, 0, 1, nullptr, nullptr, nullptr, progress_to_js, &d->tsfn);
  napi_create_async_work(env, nullptr, name, execute, complete, d, &d->work);
  napi_queue_async_work(env, d->work);
  return promise;
}

// probe(file) -> {width, height, type, orientation, orientationExif}   (stitch_probe_image; synchronous, CPU only)
napi_value Probe(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  napi_typedarray_type tt; size_t len = 0; void* ptr = nullptr; napi_value ab; size_t off = 0;
  if (argc < 1 || napi_get_typedarray_info(env, argv[0], &tt, &len, &ptr, &ab, &off) != napi_ok) {
    napi_throw_type_error(env, nullptr, "probe(file: Uint8Array)");
    return nullptr;
  }
  StitchImageInfo ii;
  const int st = stitch_probe_image(ptr, len, &ii);
  if (st != STITCH_OK) { napi_throw_error(env, nullptr, stitch_strerror(st)); return nullptr; }
  napi_value out, v;
  napi_create_object(env, &out);
  napi_create_int32(env, ii.width, &v); napi_set_named_property(env, out, "width", v);
  napi_create_int32(env, ii.height, &v); napi_set_named_property(env, out, "height", v);
  napi_create_int32(env, ii.orientation_exif, &v); napi_set_named_property(env, out, "orientationExif", v);
  napi_create_string_utf8(env, stitch_type_name(ii.type), NAPI_AUTO_LENGTH, &v); napi_set_named_property(env, out, "type", v);
  napi_create_string_utf8(env, stitch_orientation_name(ii.orientation_exif), NAPI_AUTO_LENGTH, &v);
  napi_set_named_property(env, out, "orientation", v);
  return out;
}

// preview(ctx, rgba, exportW, exportH, viewW, viewH, dpr) -> Uint8Array   (stitch_preview; synchronous)
napi_value Preview(napi_env env, napi_callback_info info) {
  size_t argc = 7; napi_value argv[7];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  void* p = nullptr;
  napi_get_value_external(env, argv[0], &p);
  napi_typedarray_type tt; size_t len = 0; void* src = nullptr; napi_value ab; size_t off = 0;
  napi_get_typedarray_info(env, argv[1], &tt, &len, &src, &ab, &off);
  int32_t ew = 0, eh = 0; double vw = 0, vh = 0, dpr = 1;
  napi_get_value_int32(env, argv[2], &ew); napi_get_value_int32(env, argv[3], &eh);
  napi_get_value_double(env, argv[4], &vw); napi_get_value_double(env, argv[5], &vh); napi_get_value_double(env, argv[6], &dpr);
  StitchPreviewLayout pl;
  int st = stitch_preview_layout(ew, eh, vw, vh, dpr, &pl);
  if (st != STITCH_OK || (size_t)ew * eh * 4 > len) { napi_throw_error(env, nullptr, stitch_strerror(st ? st : STITCH_E_BAD_ARG)); return nullptr; }
  const size_t bytes = (size_t)pl.canvas_w * pl.canvas_h * 4;
  void* dst; napi_value out_ab, out;
  napi_create_arraybuffer(env, bytes, &dst, &out_ab);
  st = stitch_preview(static_cast<StitchCtx*>(p), src, STITCH_MEM_HOST, ew, eh, 0, vw, vh, dpr, dst, (int64_t)bytes, STITCH_MEM_HOST, &pl);
  if (st != STITCH_OK) { napi_throw_error(env, nullptr, stitch_last_error(static_cast<StitchCtx*>(p))); return nullptr; }
  napi_create_typedarray(env, napi_uint8_array, bytes, out_ab, 0, &out);
  return out;
}

// decodeImage(ctx, file) -> { width, height, data: Uint8Array }   (stitch_decode_image; synchronous): loadImageFrom
// (canvas.js:27-121) for a JPEG / PNG / BMP / GIF file, decoded on the GPU into straight-alpha RGBA8
napi_value DecodeImage(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  void* p = nullptr;
  napi_get_value_external(env, argv[0], &p);
  napi_typedarray_type tt; size_t len = 0; void* src = nullptr; napi_value ab; size_t off = 0;
  if (napi_get_typedarray_info(env, argv[1], &tt, &len, &src, &ab, &off) != napi_ok) {
    napi_throw_type_error(env, nullptr, "decodeImage: file must be a Uint8Array");
    return nullptr;
  }
  StitchImageInfo pi;
  int st = stitch_probe_image(src, len, &pi);
  if (st != STITCH_OK || pi.width <= 0 || pi.height <= 0) { napi_throw_error(env, nullptr, stitch_strerror(st ? st : STITCH_E_BAD_IMAGE)); return nullptr; }
  const size_t bytes = (size_t)pi.width * (size_t)pi.height * 4;
  void* dst; napi_value out_ab, data, out, v;
  napi_create_arraybuffer(env, bytes, &dst, &out_ab);
  st = stitch_decode_image(static_cast<StitchCtx*>(p), src, len, dst, (int64_t)bytes, 0, STITCH_MEM_HOST, nullptr);
  if (st != STITCH_OK) { napi_throw_error(env, nullptr, stitch_last_error(static_cast<StitchCtx*>(p))); return nullptr; }
  napi_create_typedarray(env, napi_uint8_array, bytes, out_ab, 0, &data);
  napi_create_object(env, &out);
  napi_create_int32(env, pi.width, &v); napi_set_named_property(env, out, "width", v);
  napi_create_int32(env, pi.height, &v); napi_set_named_property(env, out, "height", v);
  napi_set_named_property(env, out, "data", data);
  return out;
}

napi_value Init(napi_env env, napi_value exports) {
  napi_property_descriptor props[] = {
      {"createContext", nullptr, CreateContext, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"destroyContext", nullptr, DestroyContext, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"run", nullptr, Run, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"probe", nullptr, Probe, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"preview", nullptr, Preview, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"decodeImage", nullptr, DecodeImage, nullptr, nullptr, nullptr, napi_default, nullptr},
  };
  napi_define_properties(env, exports, 6, props);
  return exports;
}

}  // namespace

NAPI_MODULE(NODE_GYP_MODULE_NAME, Init)
