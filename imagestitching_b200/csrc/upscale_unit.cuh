// The bilinear upscale unit of the fused compose kernel (see compose_kernel.cuh).
//
// Replaces ctx.drawImage (U:156-199) where the device rect is at least as large as the bitmap on both
// axes -- BASELINE configs[0] (2.2x supersampled canvas, P:1360-1386 + ctx.scale(ss,ss) P:1426-1428) and
// the enlarged images of configs[2] -- for EXIF 1-4.  The device rect may be fractional; the region
// holds only canvas pixels the rect covers completely, so coverage is 1 and the filter is the plain
// bilinear one of SURVEY App. C.4.  The fractional edge rows / columns are GATHER regions.
//
// One warp, one unit = 32 aligned quads of 4 output pixels (128 px) x tile_h output rows.  The work is
// output-bound (one 16-byte store per lane and row); source pixels come through L1 with 32-bit loads:
// a lane keeps, for the two source rows its output row blends, the horizontally interpolated values of
// its 4 pixels in registers and reloads a row only when the output row moves on to the next one.
//
// Horizontal pass: the 8 column taps of a lane's 4 pixels lie inside a WINDOW of WIN consecutive source
// pixels (WIN = 4 when the scale is >= 1.5, else 5: the upper taps of pixels x and x+3 are at most
// ceil(3 / scale) apart).  The lane loads and expands the window once per source row (WIN loads, 3 WIN PRMT
// instead of 8 loads and 24 PRMT for the taps one by one) and applies a per-lane 4 x WIN weight matrix that
// holds (1 - t) and t at each pixel's two taps and zeros elsewhere.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "stitch_plan.h"
#include "stream_unit.cuh"

namespace stitch {

#define UPSCALE_TILE_W 128
#define UPSCALE_TILE_H 32
#define UPSCALE_TAIL_H 16  // unit height at the end of the hand-out order (launch.cuh)

// fp64 tap position of device pixel j on one axis, exactly as gather_kernel.cuh::axis_taps evaluates it
// for a fully covered pixel: u = ((j + 0.5) - d0) / s - 0.5 with s = dlen / slen.
__device__ __forceinline__ void bilinear_tap(int j, double d0, double s, int slen, bool flip, int* ia, int* ib, float* t) {
  const double a = (double)j, b = (double)j + 1.0;
  const double u = ((a + b) * 0.5 - d0) / s - 0.5;
  const double f = floor(u);
  const int i0 = (int)f;
  *t = (float)(u - f);
  int i_a = min(max(i0, 0), slen - 1), i_b = min(max(i0 + 1, 0), slen - 1);
  if (flip) {
    i_a = slen - 1 - i_a;
    i_b = slen - 1 - i_b;
  }
  *ia = i_a;
  *ib = i_b;
}

template <int MODE, int WIN>
__device__ __forceinline__ void upscale_unit(const JobParam& P, const Region& R, int tx, int ty, int lane) {
  constexpr bool RGBA = MODE != STREAM_MODE_YUV, YUV = MODE != STREAM_MODE_RGBA;
  const ImgParam& im = P.img[R.img];
  const double sx = im.W / (double)im.sx_len, sy = im.H / (double)im.sy_len;
  const int y0 = R.y0 + ty * R.tile_h, y1 = min(y0 + R.tile_h, R.y1);
  const int q = (R.x0 >> 2) + tx * 32 + lane;  // this lane's aligned quad of canvas pixels
  const int xq = 4 * q;

  // ---- column taps of the lane's 4 pixels -> window start and weight matrix ----
  // wt[c][j]: weight of window pixel j in output pixel c.  Stored so that (wt[0][j], wt[1][j]) and
  // (wt[2][j], wt[3][j]) are register pairs: the blue channel is interpolated two pixels at a time.
  float wt[4][WIN];
  bool ok[4];
  uint32_t offW;
  {
    int ia[4], ib[4];
    float tcol[4];
    int wmin = 0x7fffffff;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int x = xq + k;
      ok[k] = x >= R.x0 && x < R.x1;
      bilinear_tap(min(max(x, R.x0), R.x1 - 1), im.X0, sx, im.sx_len, im.flip_x != 0, &ia[k], &ib[k], &tcol[k]);
      wmin = min(wmin, min(ia[k], ib[k]));
    }
    wmin = max(min(wmin, im.sx_len - WIN), 0);   // the window stays inside the row (launch.cuh: sx_len >= 5)
    offW = (uint32_t)wmin * 4u;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int ja = ia[k] - wmin, jb = ib[k] - wmin;
#pragma unroll
      for (int j = 0; j < WIN; ++j) wt[k][j] = (ja == j ? 1.f - tcol[k] : 0.f) + (jb == j ? tcol[k] : 0.f);
    }
  }
  // ---- row taps: lane r evaluates output row y0 + r once, the row loop fetches them by shuffle.
  // k_r is the UNCLAMPED index of the upper tap (floor(u), may be -1); the device rect is at least as tall
  // as the bitmap, so k advances by at most one per output row and the unit walks the source rows k, k+1, ...
  // once: it holds rows k and k+1 (horizontally interpolated) in registers, emits every output row that
  // blends them, then replaces the older one by row k+2 -- which has been in flight since the previous step.
  int k_r;
  float trow;
  {
    const int j = min(y0 + lane, y1 - 1);
    const double u = (((double)j + ((double)j + 1.0)) * 0.5 - im.Y0) / sy - 0.5;   // == bilinear_tap's u
    const double f = floor(u);
    k_r = (int)f;
    trow = (float)(u - f);
  }
  auto src_row = [&](int k) {  // clamp to edge, then the EXIF mirror
    const int c = min(max(k, 0), im.sy_len - 1);
    return im.flip_y ? im.sy_len - 1 - c : c;
  };

  uint32_t and_acc = 0xffffffffu;
  auto issue = [&](int k, uint32_t* raw) {
    const uint32_t* base = reinterpret_cast<const uint32_t*>(im.data + (int64_t)src_row(k) * im.pitch + (size_t)offW);
    asm("" : "+l"(base));  // keep ONE row pointer; the loads take immediate offsets
#pragma unroll
    for (int j = 0; j < WIN; ++j) raw[j] = __ldg(base + j);
  };
  // h layout: h[2c], h[2c+1] = red, green of pixel c; h[8 + c] = blue of pixel c (1 + v * 2^-15 form)
  auto convert = [&](const uint32_t* raw, float* h) {
    float fr[WIN], fg[WIN], fb[WIN];
#pragma unroll
    for (int j = 0; j < WIN; ++j) {
      and_acc &= raw[j];
      fr[j] = expand<0>(raw[j]);
      fg[j] = expand<1>(raw[j]);
      fb[j] = expand<2>(raw[j]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      mul2_s(h[2 * c], h[2 * c + 1], fr[0], fg[0], wt[c][0]);
#pragma unroll
      for (int j = 1; j < WIN; ++j) fma2_s(h[2 * c], h[2 * c + 1], fr[j], fg[j], wt[c][j]);
    }
#pragma unroll
    for (int c = 0; c < 4; c += 2) {
      mul2_s(h[8 + c], h[9 + c], wt[c][0], wt[c + 1][0], fb[0]);
#pragma unroll
      for (int j = 1; j < WIN; ++j) fma2_s(h[8 + c], h[9 + c], wt[c][j], wt[c + 1][j], fb[j]);
    }
  };

  // output pointers
  uint8_t* orow = RGBA ? P.out_rgba + (int64_t)y0 * P.out_pitch + (int64_t)xq * 4 : nullptr;
  const bool vec_ok = RGBA && ok[0] && ok[3] && ((((uintptr_t)P.out_rgba) | (uintptr_t)P.out_pitch) & 15) == 0;
  uint8_t* yrow = nullptr;
  uint8_t *cbp = nullptr, *crp = nullptr;
  float ccb[2] = {0.f, 0.f}, ccr[2] = {0.f, 0.f};
  int cwh = 0;
  if (YUV) {
    cwh = (P.cw + 1) >> 1;
    yrow = P.out_y + (int64_t)y0 * P.cw + xq;
    cbp = P.out_cb + (int64_t)(y0 >> 1) * cwh + (xq >> 1);
    crp = P.out_cr + (int64_t)(y0 >> 1) * cwh + (xq >> 1);
  }

  int y = y0;
  auto emit = [&](const float* A, const float* B, float tv) {
    float v[12];
    const float wa = 1.f - tv;
#pragma unroll
    for (int k = 0; k < 12; k += 2) {
      mul2_s(v[k], v[k + 1], A[k], A[k + 1], wa);
      fma2_s(v[k], v[k + 1], B[k], B[k + 1], tv);
    }
    if (RGBA) {
      // red / green pairs in the plain form, the blue pairs in the negated one (pack_rgb takes alpha from its sign)
      uint32_t q[12];
#pragma unroll
      for (int k = 0; k < 8; k += 2) quant2_one(q[k], q[k + 1], v[k], v[k + 1]);
      quant2_neg_one(q[8], q[9], v[8], v[9]);
      quant2_neg_one(q[10], q[11], v[10], v[11]);
      uint32_t px[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) px[k] = pack_rgb(q[2 * k], q[2 * k + 1], q[8 + k]);
      if (vec_ok) {
        *reinterpret_cast<uint4*>(orow) = make_uint4(px[0], px[1], px[2], px[3]);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (ok[k]) reinterpret_cast<uint32_t*>(orow)[k] = px[k];
      }
      orow += P.out_pitch;
    }
    if (YUV) {
      uint32_t yq[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float R_ = v[2 * k], G_ = v[2 * k + 1], B_ = v[8 + k];
        yq[k] = quant_bits_one(fmaf(0.299f, R_, fmaf(0.587f, G_, 0.114f * B_))) & 0xffu;
        ccb[k >> 1] += fmaf(-0.168736f, R_, fmaf(-0.331264f, G_, 0.5f * B_));
        ccr[k >> 1] += fmaf(0.5f, R_, fmaf(-0.418688f, G_, -0.081312f * B_));
      }
      if (ok[0] && ok[3] && (((uintptr_t)yrow) & 3) == 0) {
        *reinterpret_cast<uint32_t*>(yrow) = yq[0] | (yq[1] << 8) | (yq[2] << 16) | (yq[3] << 24);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (ok[k]) yrow[k] = (uint8_t)yq[k];
      }
      yrow += P.cw;
      if ((y & 1) != 0) {  // second row of the chroma block row (y0 is even)
        const float bias = 1.f + 128.f / 32768.f;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (ok[2 * c]) {   // the region is even-aligned: a pair is inside or outside as a whole
            cbp[c] = (uint8_t)quant_bits_one(fmaf(0.25f, ccb[c], bias));
            crp[c] = (uint8_t)quant_bits_one(fmaf(0.25f, ccr[c], bias));
          }
          ccb[c] = ccr[c] = 0.f;
        }
        cbp += cwh;
        crp += cwh;
      }
    }
    ++y;
  };

  float hA[12], hB[12];
  uint32_t raw[WIN];
  int k = __shfl_sync(0xffffffffu, k_r, 0);
  // bit r: output row r starts a new source-row pair (k advances by exactly one there)
  const int k_prev = __shfl_up_sync(0xffffffffu, k_r, 1);
  const uint32_t step = __ballot_sync(0xffffffffu, lane > 0 && k_r != k_prev);
  const int nrows = y1 - y0;
  {
    uint32_t raw2[WIN];
    issue(k, raw);
    issue(k + 1, raw2);
    convert(raw, hA);
    convert(raw2, hB);
  }
  issue(k + 2, raw);
  int r = 0;
  float tv = __shfl_sync(0xffffffffu, trow, 0);
  for (;;) {
    // rows k (hA) and k + 1 (hB); the next row's weight is fetched one row ahead
    do {
      ++r;
      const float tn = __shfl_sync(0xffffffffu, trow, r & 31);
      emit(hA, hB, tv);
      tv = tn;
    } while (r < nrows && !((step >> r) & 1u));
    if (r >= nrows) break;
    convert(raw, hA);
    ++k;
    issue(k + 2, raw);
    // rows k (hB) and k + 1 (hA)
    do {
      ++r;
      const float tn = __shfl_sync(0xffffffffu, trow, r & 31);
      emit(hB, hA, tv);
      tv = tn;
    } while (r < nrows && !((step >> r) & 1u));
    if (r >= nrows) break;
    convert(raw, hB);
    ++k;
    issue(k + 2, raw);
  }
  // translucent source pixels: redo the unit with the exact premultiplied source-over evaluation
  const bool need_fix = (and_acc >> 24) != 0xffu;
  if (__any_sync(0xffffffffu, need_fix)) {
    const int xa = max(R.x0, 4 * ((R.x0 >> 2) + tx * 32)), xb = min(R.x1, 4 * ((R.x0 >> 2) + tx * 32 + 32));
    redo_rect<MODE>(P, xa, y0, xb, y1, lane, 0);
  }
}

}  // namespace stitch
