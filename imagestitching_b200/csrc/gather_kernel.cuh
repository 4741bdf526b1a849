// General compose kernel ("gather"): one thread per canvas pixel, loops over the
// images in painter's order and evaluates the stated AB1 filter
// directly from global memory.  Handles everything the reference's draw loop can
// produce: fractional device rects (superSample CTM, P:1426-1428; scaledGap, P:1432),
// overlap / clip (dx=-1, last image past the canvas), EXIF 1-8 (U:153-202), straight
// alpha, upscale (bilinear) and downscale (area).  White background (P:1423-1424)
// is the initial value of the accumulator, so no separate fill pass exists.
//
// Coordinates are evaluated in fp64 exactly as the oracle does; colour is
// accumulated in fp32.  The streaming unit (stream_unit.cuh) takes over the
// bandwidth-critical integer-aligned downscale case.
#pragma once
#include <cstdint>

#include "stitch_plan.h"

namespace stitch {

struct AxisTaps {
  float cov;     // coverage of this device pixel by the device rect, 0 -> skip
  int i0;        // first tap (bitmap index along this axis before flip); may be -1 for bilinear
  int ntaps;
  float s0r;     // area: start of the source interval relative to i0;  bilinear: fraction t
  float s1r;     // area: end of the source interval relative to i0
  float inv_tot; // area: 1/(s1-s0)
  int bilinear;
};

__device__ __forceinline__ AxisTaps axis_taps(int j, double d0, double dlen, int slen) {
  AxisTaps t;
  t.cov = 0.f;
  t.i0 = 0;
  t.ntaps = 0;
  t.s0r = t.s1r = t.inv_tot = 0.f;
  t.bilinear = 0;
  const double d1 = d0 + dlen;
  const double a = fmax((double)j, d0);
  const double b = fmin((double)j + 1.0, d1);
  const double c = b - a;
  if (!(c > 0.0)) return t;
  t.cov = (float)c;
  const double s = dlen / (double)slen;
  if (s < 1.0) {
    double s0 = fmin(fmax((a - d0) / s, 0.0), (double)slen);
    double s1 = fmin(fmax((b - d0) / s, 0.0), (double)slen);
    int i0 = min((int)floor(s0), slen - 1);
    int i1 = max(i0 + 1, min((int)ceil(s1), slen));
    double tot = s1 - s0;
    t.i0 = i0;
    if (!(tot > 0.0)) {
      t.ntaps = 1;
      t.s0r = 0.f;
      t.s1r = 1.f;
      t.inv_tot = 1.f;
    } else {
      t.ntaps = i1 - i0;
      t.s0r = (float)(s0 - (double)i0);
      t.s1r = (float)(s1 - (double)i0);
      t.inv_tot = (float)(1.0 / tot);
    }
  } else {
    const double u = ((a + b) * 0.5 - d0) / s - 0.5;
    const double f = floor(u);
    t.bilinear = 1;
    t.i0 = (int)f;
    t.ntaps = 2;
    t.s0r = (float)(u - f);
  }
  return t;
}

__device__ __forceinline__ float tap_weight(const AxisTaps& t, int k) {
  if (t.bilinear) return k == 0 ? 1.f - t.s0r : t.s0r;
  float lo = fmaxf(t.s0r, (float)k);
  float hi = fminf(t.s1r, (float)(k + 1));
  return fmaxf(hi - lo, 0.f) * t.inv_tot;
}

// Un-rounded composited colour of canvas pixel (x, y): premultiplied RGB and alpha, all on the 0..255
// scale.  Over the opaque white canvas A stays 255 and R, G, B are the final colour.
__device__ __forceinline__ void shade(const JobParam& P, int x, int y, float& R, float& G, float& B, float& A) {
  R = G = B = A = P.transparent ? 0.f : 255.f;
  for (int k = 0; k < P.n; ++k) {
    const ImgParam& im = P.img[k];
    if (x < im.x_lo || x >= im.x_hi || y < im.y_lo || y >= im.y_hi) continue;
    AxisTaps tx = axis_taps(x, im.X0, im.W, im.sx_len);
    AxisTaps ty = axis_taps(y, im.Y0, im.H, im.sy_len);
    if (tx.cov <= 0.f || ty.cov <= 0.f) continue;
    float fr = 0.f, fg = 0.f, fb = 0.f, fa = 0.f;
    for (int ky = 0; ky < ty.ntaps; ++ky) {
      const float wy = tap_weight(ty, ky);
      int iy = min(max(ty.i0 + ky, 0), im.sy_len - 1);
      if (im.flip_y) iy = im.sy_len - 1 - iy;
      for (int kx = 0; kx < tx.ntaps; ++kx) {
        const float w = wy * tap_weight(tx, kx);
        int ix = min(max(tx.i0 + kx, 0), im.sx_len - 1);
        if (im.flip_x) ix = im.sx_len - 1 - ix;
        // device (ix, iy) -> bitmap (col, row)
        const int col = im.transpose ? iy : ix;
        const int row = im.transpose ? ix : iy;
        const uint32_t px = __ldg(reinterpret_cast<const uint32_t*>(im.data + (int64_t)row * im.pitch) + col);
        const float a = (float)(px >> 24);
        const float wa = w * a * (1.f / 255.f);  // premultiply (App. C.6)
        fr = fmaf(wa, (float)(px & 0xffu), fr);
        fg = fmaf(wa, (float)((px >> 8) & 0xffu), fg);
        fb = fmaf(wa, (float)((px >> 16) & 0xffu), fb);
        fa = fmaf(w, a, fa);
      }
    }
    const float C = tx.cov * ty.cov;
    const float keep = 1.f - C * fa * (1.f / 255.f);
    R = fmaf(C, fr, keep * R);
    G = fmaf(C, fg, keep * G);
    B = fmaf(C, fb, keep * B);
    A = fmaf(C, fa, keep * A);
  }
}
__device__ __forceinline__ void shade(const JobParam& P, int x, int y, float& R, float& G, float& B) {
  float A;
  shade(P, x, y, R, G, B, A);
}

__device__ __forceinline__ uint32_t q8(float v) {  // round half up, clamp (App. C.7)
  return (uint32_t)min(max(__float2int_rd(v + 0.5f), 0), 255);
}

// RGBA8 of a shaded pixel: opaque over white, or straight alpha over the cleared canvas (colour =
// premultiplied / alpha before rounding; nothing where alpha rounds to 0).
__device__ __forceinline__ uint32_t finish_px(const JobParam& P, float R, float G, float B, float A) {
  if (!P.transparent) return q8(R) | (q8(G) << 8) | (q8(B) << 16) | 0xff000000u;
  const uint32_t a8 = q8(A);
  if (a8 == 0 || !(A > 0.f)) return 0u;
  const float inv = 255.f / A;
  return q8(R * inv) | (q8(G * inv) << 8) | (q8(B * inv) << 16) | (a8 << 24);
}

__global__ void __launch_bounds__(256) gather_rgba_kernel(const __grid_constant__ JobParam P) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31);
  const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= P.cw || y >= P.ch) return;
  float R, G, B, A;
  shade(P, x, y, R, G, B, A);
  *reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch + (int64_t)x * 4) = finish_px(P, R, G, B, A);
}

// One thread per 2x2 block: RGBA (optional) + JFIF BT.601 full-range I420 (App. C.9).
__global__ void __launch_bounds__(256) gather_yuv_kernel(const __grid_constant__ JobParam P) {
  const int bx = blockIdx.x * 32 + (threadIdx.x & 31);
  const int by = blockIdx.y * 8 + (threadIdx.x >> 5);
  const int cwh = (P.cw + 1) >> 1, chh = (P.ch + 1) >> 1;
  if (bx >= cwh || by >= chh) return;
  float cb = 0.f, cr = 0.f;
#pragma unroll
  for (int dy = 0; dy < 2; ++dy) {
#pragma unroll
    for (int dx = 0; dx < 2; ++dx) {
      const int x = min(bx * 2 + dx, P.cw - 1);  // edge replication for odd sizes
      const int y = min(by * 2 + dy, P.ch - 1);
      float R, G, B;
      shade(P, x, y, R, G, B);
      const bool inside = (bx * 2 + dx < P.cw) && (by * 2 + dy < P.ch);
      if (inside) {
        const float Y = 0.299f * R + 0.587f * G + 0.114f * B;
        P.out_y[(int64_t)y * P.cw + x] = (uint8_t)q8(Y);
        if (P.out_rgba) {
          const uint32_t v = q8(R) | (q8(G) << 8) | (q8(B) << 16) | 0xff000000u;
          *reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch + (int64_t)x * 4) = v;
        }
      }
      cb += 128.f - 0.168736f * R - 0.331264f * G + 0.5f * B;
      cr += 128.f + 0.5f * R - 0.418688f * G - 0.081312f * B;
    }
  }
  P.out_cb[(int64_t)by * cwh + bx] = (uint8_t)q8(cb * 0.25f);
  P.out_cr[(int64_t)by * cwh + bx] = (uint8_t)q8(cr * 0.25f);
}

}  // namespace stitch
