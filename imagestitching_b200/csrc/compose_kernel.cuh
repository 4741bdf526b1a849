// Fused resample + compose kernel (K1): ONE launch draws every image of a job into the canvas.
//
// Replaces the per-image ctx.drawImage loop (P:1439-1572 -> U:153-202) and the canvas
// pre-fill (P:1423-1424).  The canvas is partitioned on the host (plan.h) into regions,
// every region into warp-sized units; each warp of the grid executes one unit:
//
//   REGION_STREAM  integer-aligned area downscale (1 < ratio_x <= 4, ratio_y > 1), EXIF 1-4.
//       The bandwidth path.  A warp owns a band of <=31 source groups (4 px = 16 B each) and
//       streams DOWN the source rows with one coalesced 128-bit load per lane per row
//       (register prefetch ring, no shared-memory staging of pixels).  Vertical area weights
//       are warp-uniform (row table in smem); the horizontal pass runs in registers when an
//       output row completes: each lane reduces its 4 columns into the <=MAXO outputs that
//       start in its group plus the tail of its left neighbour's last output, which travels
//       by one shuffle.  Every source byte is loaded exactly once per warp.
//       Arithmetic: a byte b is expanded with ONE PRMT to the float 1 + b*2^-15
//       (0x3F80bb00); because the filter weights sum to 1 the accumulators stay in that
//       form, and one FFMA.RM against 2^23 - 2^15 + 0.5 leaves floor(value + 0.5) in the
//       low mantissa byte.  Resolution of the form is 2^-8 LSB per rounding.
//       Pixels whose filtered alpha is not 255 fall back to the exact gather path.
//   REGION_FILL    white.
//   REGION_GATHER  per-pixel evaluation (gather_kernel.cuh shade()) for upscale / EXIF 5-8 /
//       ratios outside the streaming range.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "gather_kernel.cuh"
#include "stitch_plan.h"

namespace stitch {

#ifndef STREAM_NPF
#define STREAM_NPF 4  // rows of prefetch per lane
#endif
#ifndef STREAM_MINB
#define STREAM_MINB 2  // resident CTAs per SM the register allocation targets
#endif

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// byte ch of px -> float 1 + b * 2^-15
template <int CH>
__device__ __forceinline__ float expand(uint32_t px) {
  return __uint_as_float(__byte_perm(px, 0x3F800000u, 0x7604u | (CH << 4)));
}

// Exact (gather) evaluation of one canvas pixel, kept out of line so the rare slow path does not
// inflate the streaming loop's register allocation.
__device__ __noinline__ uint32_t shade_px(const JobParam& P, int x, int y) {
  float Rr, Gg, Bb;
  shade(P, x, y, Rr, Gg, Bb);
  return q8(Rr) | (q8(Gg) << 8) | (q8(Bb) << 16) | 0xff000000u;
}

__device__ __forceinline__ int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

template <int MAXO>
__device__ __forceinline__ void stream_unit(const JobParam& P, const Region& R, int tx, int ty, float2* tab, int lane) {
  const ImgParam& im = P.img[R.img];
  const int sw = im.sx_len, sh = im.sy_len;
  const int dw = (int)im.W, dh = (int)im.H;
  const int X0 = (int)im.X0, Y0 = (int)im.Y0;
  const int ox0 = R.x0 + tx * R.tile_w, ox1 = min(ox0 + R.tile_w, R.x1);
  const int oy0 = R.y0 + ty * R.tile_h, oy1 = min(oy0 + R.tile_h, R.y1);
  const float inv_sw = 1.0f / (float)sw, inv_sh = 1.0f / (float)sh;

  // ---- vertical: source rows [rs, re) and the warp's row table ----
  const int rs = (int)(((int64_t)oy0 * sh) / dh);
  const int re = (int)min((int64_t)sh, ceil_div64((int64_t)oy1 * sh, dh));
  const int nrows = re - rs;
  for (int k = lane; k < nrows; k += 32) {
    const int64_t i = rs + k;
    const int64_t lo = i * dh, hi = lo + dh;          // row i in units of 1/dh ... scaled by dh*sh
    const int64_t Ya = lo / sh;                        // output row containing the start of row i
    const int64_t E = (Ya + 1) * sh;                   // where output row Ya ends
    float wc, wn;
    if (E >= hi) {                                     // row entirely inside output Ya
      wc = (float)dh * inv_sh;
      wn = (E == hi) ? 0.f : -1.f;                     // exact boundary at the row end: emit, next starts clean
    } else {
      wc = (float)(E - lo) * inv_sh;
      wn = (float)(hi - E) * inv_sh;
    }
    if (Ya < oy0) {                                    // contribution to the previous unit's row is dropped
      wc = (wn > 0.f) ? wn : 0.f;
      wn = -1.f;
    }
    tab[k] = make_float2(wc, wn);
  }
  __syncwarp();

  // ---- horizontal: this lane's 4-pixel source group and the outputs that start in it ----
  const int g0 = (int)((((int64_t)ox0 * sw) / dw) >> 2);
  const int g = g0 + lane;
  const int64_t c0 = (int64_t)4 * g * dw;              // group start in scaled units
  const int Xfirst = (int)ceil_div64(c0, sw);
  const int Xnext = (int)ceil_div64(c0 + (int64_t)4 * dw, sw);
  const int cnt = min(Xnext - Xfirst, MAXO);
  float w[MAXO][4], wprev[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int64_t a = c0 + (int64_t)c * dw, b = a + dw;
    {
      const int64_t s = (int64_t)(Xfirst - 1) * sw, e = s + sw;
      const int64_t ov = min(e, b) - max(s, a);
      wprev[c] = ov > 0 ? (float)ov * inv_sw : 0.f;
    }
#pragma unroll
    for (int o = 0; o < MAXO; ++o) {
      const int64_t s = (int64_t)(Xfirst + o) * sw, e = s + sw;
      const int64_t ov = min(e, b) - max(s, a);
      w[o][c] = (ov > 0 && Xfirst + o < dw) ? (float)ov * inv_sw : 0.f;
    }
  }
  const bool active = ((int64_t)4 * g * dw < (int64_t)ox1 * sw) && (4 * g < sw);
  bool st_ok[MAXO];
  int xdev[MAXO];
#pragma unroll
  for (int o = 0; o < MAXO; ++o) {
    const int X = Xfirst + o;
    st_ok[o] = active && (o < cnt) && (X >= ox0) && (X < ox1);
    xdev[o] = X0 + (im.flip_x ? dw - 1 - X : X);
  }

  const uint8_t* base = im.data + (int64_t)rs * im.pitch + (int64_t)g * 16;
  uint4 ring[STREAM_NPF];
#pragma unroll
  for (int u = 0; u < STREAM_NPF; ++u) {
    ring[u] = make_uint4(0, 0, 0, 0);
    if (active && u < nrows) ring[u] = ldg_stream(reinterpret_cast<const uint4*>(base + (int64_t)u * im.pitch));
  }
  float acc[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) acc[k] = 0.f;
  int Y = oy0;
  bool need_fix = false;

  for (int i = 0; i < nrows; i += STREAM_NPF) {
#pragma unroll
    for (int u = 0; u < STREAM_NPF; ++u) {
      const int ii = i + u;
      if (ii < nrows) {
        const uint4 px = ring[u];
        if (active && ii + STREAM_NPF < nrows)
          ring[u] = ldg_stream(reinterpret_cast<const uint4*>(base + (int64_t)(ii + STREAM_NPF) * im.pitch));
        const float2 t = tab[ii];
        float f[16];
        f[0] = expand<0>(px.x); f[1] = expand<1>(px.x); f[2] = expand<2>(px.x); f[3] = expand<3>(px.x);
        f[4] = expand<0>(px.y); f[5] = expand<1>(px.y); f[6] = expand<2>(px.y); f[7] = expand<3>(px.y);
        f[8] = expand<0>(px.z); f[9] = expand<1>(px.z); f[10] = expand<2>(px.z); f[11] = expand<3>(px.z);
        f[12] = expand<0>(px.w); f[13] = expand<1>(px.w); f[14] = expand<2>(px.w); f[15] = expand<3>(px.w);
#pragma unroll
        for (int k = 0; k < 16; ++k) acc[k] = fmaf(t.x, f[k], acc[k]);
        if (t.y >= 0.f) {
          // ---- output row Y complete: horizontal pass in registers ----
          float prev[4], out[MAXO][4];
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            prev[ch] = wprev[0] * acc[ch];
#pragma unroll
            for (int c = 1; c < 4; ++c) prev[ch] = fmaf(wprev[c], acc[c * 4 + ch], prev[ch]);
#pragma unroll
            for (int o = 0; o < MAXO; ++o) {
              out[o][ch] = w[o][0] * acc[ch];
#pragma unroll
              for (int c = 1; c < 4; ++c) out[o][ch] = fmaf(w[o][c], acc[c * 4 + ch], out[o][ch]);
            }
          }
          const int ydev = Y0 + (im.flip_y ? dh - 1 - Y : Y);
          uint8_t* orow = P.out_rgba + (int64_t)ydev * P.out_pitch;
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            float rcv = __shfl_down_sync(0xffffffffu, prev[ch], 1);
            if (lane == 31) rcv = 0.f;
#pragma unroll
            for (int o = 0; o < MAXO; ++o)
              if (o == cnt - 1) out[o][ch] += rcv;
          }
#pragma unroll
          for (int o = 0; o < MAXO; ++o) {
            if (st_ok[o]) {
              // filtered alpha must be 255 for the plain (non-premultiplied) average to be exact
              need_fix |= !(fmaf(out[o][3], 32768.f, -32768.f) > 254.95f);
              const uint32_t r = __float_as_uint(__fmaf_rd(out[o][0], 32768.f, 8355840.5f));
              const uint32_t gg = __float_as_uint(__fmaf_rd(out[o][1], 32768.f, 8355840.5f));
              const uint32_t b = __float_as_uint(__fmaf_rd(out[o][2], 32768.f, 8355840.5f));
              const uint32_t v = __byte_perm(__byte_perm(r, gg, 0x0040u), b, 0x0410u) | 0xff000000u;
              *reinterpret_cast<uint32_t*>(orow + (int64_t)xdev[o] * 4) = v;
            }
          }
          ++Y;
#pragma unroll
          for (int k = 0; k < 16; ++k) acc[k] = t.y * f[k];
        }
      }
    }
  }
  // Translucent bitmaps (rare: photos are opaque): the unit is redone with the exact premultiplied
  // source-over evaluation.  Outside the hot loop so it costs the streaming path no registers.
  if (__any_sync(0xffffffffu, need_fix)) {
    const int ax = im.flip_x ? dw - ox1 : ox0, bx = im.flip_x ? dw - ox0 : ox1;
    const int ay = im.flip_y ? dh - oy1 : oy0, by = im.flip_y ? dh - oy0 : oy1;
    for (int y = Y0 + ay; y < Y0 + by; ++y) {
      uint32_t* row = reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch);
      for (int x = X0 + ax + lane; x < X0 + bx; x += 32) row[x] = shade_px(P, x, y);
    }
  }
}

__device__ __forceinline__ void fill_unit(const JobParam& P, const Region& R, int tx, int ty, int lane) {
  const int x0 = R.x0 + tx * R.tile_w, x1 = min(x0 + R.tile_w, R.x1);
  const int y0 = R.y0 + ty * R.tile_h, y1 = min(y0 + R.tile_h, R.y1);
  for (int y = y0; y < y1; ++y) {
    uint32_t* row = reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch);
    for (int x = x0 + lane; x < x1; x += 32) row[x] = 0xffffffffu;
  }
}

__device__ __forceinline__ void gather_unit(const JobParam& P, const Region& R, int tx, int ty, int lane) {
  const int x0 = R.x0 + tx * R.tile_w, x1 = min(x0 + R.tile_w, R.x1);
  const int y0 = R.y0 + ty * R.tile_h, y1 = min(y0 + R.tile_h, R.y1);
  for (int y = y0; y < y1; ++y) {
    uint32_t* row = reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch);
    for (int x = x0 + lane; x < x1; x += 32) {
      row[x] = shade_px(P, x, y);
    }
  }
}

// MAXO = max over the job's streaming regions of ceil(4 / ratio_x): 2 for ratio > 2 (the 9x12MP
// config), 3 for ratio > 4/3, 4 for ratio > 1.
template <int MAXO>
__global__ void __launch_bounds__(STREAM_WARPS * 32, STREAM_MINB)
compose_kernel(const __grid_constant__ JobParam P, const __grid_constant__ Plan plan) {
  __shared__ float2 tabs[STREAM_WARPS][STREAM_TAB_MAX];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int unit = blockIdx.x * STREAM_WARPS + warp;
  if (unit >= plan.n_units) return;
  int ri = 0;
  while (ri + 1 < plan.n_regions && unit >= plan.r[ri + 1].first_unit) ++ri;
  const Region& R = plan.r[ri];
  const int local = unit - R.first_unit;
  const int ty = local / R.tiles_x, tx = local - ty * R.tiles_x;
  if (R.kind == REGION_STREAM) {
    stream_unit<MAXO>(P, R, tx, ty, tabs[warp], lane);
  } else if (R.kind == REGION_FILL) {
    fill_unit(P, R, tx, ty, lane);
  } else {
    gather_unit(P, R, tx, ty, lane);
  }
}

}  // namespace stitch
