// Layout twin: bit-exact C++ restatement of the reference's double arithmetic.
// Build with -ffp-contract=off (no FMA contraction) and never -ffast-math.
//
// Reference (under /root/reference/miniprogram-stitch/miniprogram/):
//   P = pages/index/index.js
//   P:1211-1213 bigTask | P:1243-1244 normalise | P:1251-1320 base size |
//   P:1324-1336 limits  | P:1337-1357 clamp     | P:1360-1386 supersample |
//   P:1432-1433,1522-1554 placement | P:126-156 fallback profile |
//   P:1193,1247,1358,1556-1557,1581,1611 progress values.
#include <cmath>
#include <cstring>

#include "stitch.h"

namespace {

// ECMA-262 Math.round: nearest integer, ties toward +infinity (not C round()).
inline double js_round(double x) {
  double f = std::floor(x);
  return (x - f >= 0.5) ? f + 1.0 : f;
}
inline double dmin(double a, double b) { return a < b ? a : b; }
inline double dmax(double a, double b) { return a > b ? a : b; }
// JS `a || b || c` on numbers.
inline int32_t first_truthy(int32_t a, int32_t b, int32_t c) { return a ? a : (b ? b : c); }

}  // namespace

extern "C" int stitch_profile_preset(int32_t platform, StitchProfile* out) {
  if (!out || platform < STITCH_IOS || platform > STITCH_OTHER) return STITCH_E_BAD_ARG;
  // P:128-138 applyFallback()
  int64_t size = platform == STITCH_IOS ? 12288 : 4096;
  int64_t cap = platform == STITCH_ANDROID ? size * (size < 4096 ? size : 4096) : size * (size < 12288 ? size : 12288);
  int64_t floor_px = 4096LL * 2048LL;
  out->platform = platform;
  out->max_canvas_size = (int32_t)size;
  out->max_canvas_pixels = cap > floor_px ? cap : floor_px;
  return STITCH_OK;
}

extern "C" int stitch_progress_sequence(int32_t n, int32_t* out, int32_t cap) {
  if (n <= 0) return 0;
  int k = 0;
  auto push = [&](int v) {
    if (out && k < cap) out[k] = v;
    ++k;
  };
  push(1);  // P:1193
  for (int done = 1; done <= n; ++done) {  // P:1247
    double p = js_round(1 + ((double)done / (double)n) * 24);
    push((int)dmin(25, p));
  }
  push(30);  // P:1358
  for (int idx = 0; idx < n; ++idx) {  // P:1556-1557
    double p2 = 30 + js_round(((double)(idx + 1) / (double)n) * 60);
    push((int)dmin(90, p2));
  }
  push(96);   // P:1581
  push(100);  // P:1611
  return k;
}

extern "C" int stitch_layout(const StitchImageDesc* images, int32_t n, const StitchParams* params, StitchLayout* L) {
  if (!params || !L) return STITCH_E_BAD_ARG;
  if (n <= 0) return STITCH_E_EMPTY;  // P:1189
  if (!images) return STITCH_E_BAD_ARG;
  if (n > STITCH_MAX_IMAGES) return STITCH_E_TOO_MANY;  // P:311
  if (params->direction != STITCH_VERTICAL && params->direction != STITCH_HORIZONTAL) return STITCH_E_BAD_ARG;
  if (params->mode < STITCH_MODE_MIN || params->mode > STITCH_MODE_OTHER) return STITCH_E_BAD_ARG;
  const int32_t platform = params->profile.platform;
  if (platform < STITCH_IOS || platform > STITCH_OTHER) return STITCH_E_BAD_ARG;

  std::memset(L, 0, sizeof(*L));
  L->n = n;

  double totalBytes = 0;  // P:1211
  for (int i = 0; i < n; ++i) totalBytes = totalBytes + (images[i].file_size ? (double)images[i].file_size : 0.0);
  const bool big = (n >= 7) || (totalBytes >= 25.0 * 1024 * 1024);  // P:1212
  L->big_task = big ? 1 : 0;

  int32_t nw[STITCH_MAX_IMAGES], nh[STITCH_MAX_IMAGES];
  bool anyW = false, anyH = false;
  for (int i = 0; i < n; ++i) {  // P:1243-1244
    int32_t w = first_truthy(images[i].natural_width, images[i].width, 1);
    int32_t h = first_truthy(images[i].natural_height, images[i].height, 1);
    nw[i] = w > 1 ? w : 1;
    nh[i] = h > 1 ? h : 1;
    L->natural_w[i] = nw[i];
    L->natural_h[i] = nh[i];
    anyW = anyW || nw[i] > 0;  // P:1252 filter(n => n > 0)
    anyH = anyH || nh[i] > 0;
  }
  if (!anyW || !anyH) return STITCH_E_BAD_DIMS;  // P:1254

  const double gapIn = params->gap;
  const double gapPx = (gapIn != gapIn || gapIn == 0) ? 0.0 : gapIn;  // P:1256 `gap || 0`
  const int mode = params->mode;
  const bool vertical = params->direction == STITCH_VERTICAL;
  const bool minmax = (mode == STITCH_MODE_MIN || mode == STITCH_MODE_MAX);

  double outW = 0, outH = 0;
  {
    // Base extent along the common edge: min / max of that edge (P:1264,1273,1282,1291,1300,1309).
    double lo = vertical ? nw[0] : nh[0], hi = lo;
    for (int i = 1; i < n; ++i) {
      double v = vertical ? nw[i] : nh[i];
      lo = dmin(lo, v);
      hi = dmax(hi, v);
    }
    const double common = (mode == STITCH_MODE_MIN) ? lo : hi;
    double sum = 0;
    for (int i = 0; i < n; ++i) {
      const double iw = nw[i], ih = nh[i];
      double extent;
      if (minmax) {
        extent = vertical ? ih * (common / iw) : iw * (common / ih);  // P:1268,1277,1295,1304
      } else {
        extent = vertical ? ih : iw;  // P:1284-1285,1311-1312
      }
      sum = sum + extent + (i ? gapPx : 0.0);  // left-assoc, P:1269
    }
    if (vertical) {
      outW = common;
      outH = sum;
    } else {
      outH = common;
      outW = sum;
    }
  }

  outW = dmax(1, js_round(outW));  // P:1318
  outH = dmax(1, js_round(outH));  // P:1319
  if (!(outW != 0) || !(outH != 0) || outW != outW || outH != outH) return STITCH_E_BAD_OUT;  // P:1320
  L->pre_w = outW;
  L->pre_h = outH;

  // P:1325-1336
  const double SIDE = params->profile.max_canvas_size ? (double)params->profile.max_canvas_size
                                                      : (platform == STITCH_ANDROID ? 4096.0 : 12288.0);
  const double devPix = (double)params->profile.max_canvas_pixels;
  double PIX;
  if (platform == STITCH_IOS) {
    const double iosLimit = 16384.0 * 1400.0;
    PIX = dmin(devPix != 0 ? devPix : iosLimit, iosLimit);
  } else if (platform == STITCH_ANDROID) {
    const double lim = SIDE * dmin(SIDE, 8192);
    PIX = dmin(devPix != 0 ? devPix : lim, lim);
  } else {
    const double lim = SIDE * SIDE;
    PIX = dmin(devPix != 0 ? devPix : lim, lim);
  }
  L->max_side = SIDE;
  L->max_pixels = PIX;

  double scaleDown = 1;  // P:1337
  if (outW > SIDE || outH > SIDE) scaleDown = dmin(SIDE / outW, SIDE / outH);  // P:1340-1342
  const double totalPixels = outW * outH;  // P:1345
  if (totalPixels > PIX) {
    const double pixelScale = std::sqrt(PIX / totalPixels);  // P:1347
    scaleDown = dmin(scaleDown, pixelScale);
  }
  if (scaleDown < 1) {  // P:1351-1353
    outW = dmax(1, std::floor(outW * scaleDown));
    outH = dmax(1, std::floor(outH * scaleDown));
  }
  L->scale_down = scaleDown;

  const double targetW = outW, targetH = outH;  // P:1360-1361
  L->target_w = targetW;
  L->target_h = targetH;
  const double basePixels = targetW * targetH;
  const double MAX_SS = big ? 1.0 : (platform == STITCH_IOS ? 2.2 : 2.6);  // P:1363
  double ss = 1;
  if (basePixels > 0 && basePixels < PIX) {  // P:1365
    const double ratio = std::sqrt(PIX / basePixels);
    if (ratio > 1.01) {
      const double sideCap = dmin(SIDE / targetW, SIDE / targetH);
      ss = dmin(dmin(MAX_SS, ratio), sideCap);  // P:1369 Math.min(a,b,c)
    }
  }
  if (!std::isfinite(ss) || ss < 1) ss = 1;  // P:1372
  double cw = dmax(1, js_round(targetW * ss));  // P:1373
  double ch = dmax(1, js_round(targetH * ss));
  int guard = 0;
  while (cw * ch > PIX && guard < 20) {  // P:1377-1383
    ss *= 0.96;
    if (ss < 1) {
      ss = 1;
      break;  // cw/ch intentionally keep the previous iteration's values
    }
    cw = dmax(1, std::floor(targetW * ss));
    ch = dmax(1, std::floor(targetH * ss));
    ++guard;
  }
  L->super_sample = ss;
  L->canvas_w = cw;
  L->canvas_h = ch;

  const double scaledGap = gapPx * scaleDown;  // P:1432
  L->scaled_gap = scaledGap;
  double cursorX = 0, cursorY = 0;
  for (int i = 0; i < n; ++i) {  // P:1522-1554
    const double naturalW = nw[i], naturalH = nh[i];
    StitchRect& r = L->rects[i];
    if (vertical) {
      if (mode == STITCH_MODE_ORIGINAL) {
        const double dw = js_round(naturalW * scaleDown);
        const double dh = js_round(naturalH * scaleDown);
        const double dx = std::floor((outW - dw) / 2);
        r = StitchRect{dx, cursorY, dw, dh};
        cursorY += dh + scaledGap;
      } else {
        const double drawH = js_round(naturalH * (outW / naturalW));
        r = StitchRect{0, cursorY, outW, drawH};
        cursorY += drawH + scaledGap;
      }
    } else {
      if (mode == STITCH_MODE_ORIGINAL) {
        const double dw = js_round(naturalW * scaleDown);
        const double dh = js_round(naturalH * scaleDown);
        const double dy = std::floor((outH - dh) / 2);
        r = StitchRect{cursorX, dy, dw, dh};
        cursorX += dw + scaledGap;
      } else {
        const double drawW = js_round(naturalW * (outH / naturalH));
        r = StitchRect{cursorX, 0, drawW, outH};
        cursorX += drawW + scaledGap;
      }
    }
  }
  return STITCH_OK;
}

extern "C" int stitch_output_size(const StitchLayout* layout, int64_t* rgba_bytes, int64_t* yuv_bytes) {
  if (!layout) return STITCH_E_BAD_ARG;
  // U:206-209: w,h = max(1, floor(canvasOutW/H)); P:1577-1579 exports the whole canvas unscaled.
  const int64_t w = (int64_t)dmax(1, std::floor(layout->canvas_w));
  const int64_t h = (int64_t)dmax(1, std::floor(layout->canvas_h));
  if (rgba_bytes) *rgba_bytes = w * h * 4;
  if (yuv_bytes) *yuv_bytes = w * h + 2 * ((w + 1) / 2) * ((h + 1) / 2);
  return STITCH_OK;
}

extern "C" int64_t stitch_jpeg_capacity(const StitchLayout* layout, int32_t restart_interval) {
  if (!layout) return 0;
  const int64_t w = (int64_t)dmax(1, std::floor(layout->canvas_w));
  const int64_t h = (int64_t)dmax(1, std::floor(layout->canvas_h));
  const int64_t ri = restart_interval > 0 ? restart_interval : 8;
  const int64_t mcus = ((w + 15) / 16) * ((h + 15) / 16);
  const int64_t segs = (mcus + ri - 1) / ri;
  // header < 1 KB; 6 blocks of at most 432 bytes per MCU; per segment one padded byte (stuffed) + RSTn
  return 1024 + mcus * 6 * 432 + segs * 4 + 2;
}

extern "C" int64_t stitch_png_capacity(const StitchLayout* layout) {
  if (!layout) return 0;
  const int64_t w = (int64_t)layout->canvas_w, h = (int64_t)layout->canvas_h;
  const int64_t bands = (h + 7) / 8;
  // per band: chunk framing 12, block header (1239 bits) + end of block + stored block or Adler + zlib header < 192
  return 33 + 12 + bands * (12 + 192) + (h * (w * 4 + 1) * 15 + 7) / 8 + 64;
}

// Preview geometry: U:17-24 (createHighResCanvas) + P:1599-1603.
extern "C" int stitch_preview_layout(int32_t export_w, int32_t export_h, double view_w, double view_h, double dpr,
                                     StitchPreviewLayout* out) {
  if (!out) return STITCH_E_BAD_ARG;
  std::memset(out, 0, sizeof(*out));
  if (export_w <= 0 || export_h <= 0) return STITCH_E_BAD_DIMS;
  if (!(dpr == dpr) || dpr == 0.0) dpr = 1.0;  // `info.pixelRatio || 1` (0 and NaN are falsy)
  out->dpr = dpr;
  const double cw = std::floor(view_w * dpr), ch = std::floor(view_h * dpr);  // U:19-20
  if (!(cw >= 1.0) || !(ch >= 1.0) || cw > 2147483647.0 || ch > 2147483647.0) return STITCH_E_BAD_OUT;
  out->canvas_w = (int32_t)cw;
  out->canvas_h = (int32_t)ch;
  const double fx = view_w / (double)export_w, fy = view_h / (double)export_h;
  out->scale_fit = fx < fy ? fx : fy;                                   // P:1600
  out->pv_w = (int32_t)js_round((double)export_w * out->scale_fit);     // P:1601
  out->pv_h = (int32_t)js_round((double)export_h * out->scale_fit);     // P:1602
  out->dx = (view_w - (double)out->pv_w) / 2;                           // P:1603
  out->dy = (view_h - (double)out->pv_h) / 2;
  return STITCH_OK;
}

extern "C" const char* stitch_strerror(int status) {
  switch (status) {
    case STITCH_OK: return "ok";
    case STITCH_E_EMPTY: return "empty image list";
    case STITCH_E_BAD_DIMS: return "图片尺寸不可用";
    case STITCH_E_BAD_OUT: return "输出尺寸计算失败";
    case STITCH_E_NO_SURFACE: return "无法获取绘图上下文";
    case STITCH_E_BAD_IMAGE: return "图片解码异常";
    case STITCH_E_TOO_MANY: return "too many images (max 9)";
    case STITCH_E_BAD_ARG: return "bad argument";
    case STITCH_E_CUDA: return "CUDA error";
    case STITCH_E_BUSY: return "stitch already in progress";
    default: return "unknown status";
  }
}

extern "C" int stitch_abi_version(void) { return STITCH_ABI_VERSION; }
