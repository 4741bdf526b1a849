// The streaming area-downscale unit of the fused compose kernel (see compose_kernel.cuh).
//
// One warp, one unit = a band of <=32 source groups (4 px = 16 B per lane) x tile_h output rows
// of one image.  Replaces ctx.drawImage (U:156-199) for integer-aligned downscales.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "stitch_plan.h"

namespace stitch {

#ifndef STREAM_NST
#define STREAM_NST 4  // ring depth == row-loop unroll: source rows in flight per lane
#endif
#define STREAM_STAGE_BYTES 512

struct WarpSmem {
  alignas(128) uint8_t ring[STREAM_NST][STREAM_STAGE_BYTES];
  alignas(8) float2 tab[STREAM_TAB_MAX];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Per-lane 16-byte async copy global -> shared (LDGSTS), L1 bypassed.  A lane only ever refills the
// slot it has itself just read with LDS; both go through the lane's LSU queue in program order and
// the copy's shared-memory write lands a global-memory latency later, so no barrier is needed and
// no other lane ever touches the slot.
__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global.L2::128B [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}

// byte CH of px -> float 1 + b * 2^-15  (bit pattern 0x3F80bb00): one PRMT
template <int CH>
__device__ __forceinline__ float expand(uint32_t px) {
  return __uint_as_float(__byte_perm(px, 0x3F800000u, 0x7604u | (CH << 4)));
}
// float 1 + v * 2^-15 (v in [0,255]) -> 0x4B0000kk with kk = floor(v + 0.5): one FFMA.RM
__device__ __forceinline__ uint32_t quant_bits(float x) { return __float_as_uint(__fmaf_rd(x, 32768.f, 8355840.5f)); }

// Integer geometry in 32 bits when every product fits (the common case), else 64 bits.
template <typename T>
struct Geo {
  static __device__ __forceinline__ T fdiv(T n, T d) { return n / d; }
  static __device__ __forceinline__ T cdiv(T n, T d) { return (n + d - 1) / d; }
};

// Builds the warp's row table for source rows [rs, rs + nrows): entry = (weight into the current
// output row, weight into the next one or -1 when the row does not complete an output row).
// Each lane fills a contiguous chunk, so it needs one division and then steps incrementally.
template <typename T>
__device__ __forceinline__ void build_row_table(float2* tab, int lane, int rs, int nrows, int nrows_pad, int sh, int dh,
                                                int oy0, float inv_sh) {
  const int chunk = (nrows_pad + 31) >> 5;
  int k = lane * chunk;
  const int k_end = min(k + chunk, nrows_pad);
  if (k >= k_end) return;
  T lo = (T)(rs + k) * (T)dh;      // start of row (rs+k) in source units scaled by dh
  T Ya = lo / (T)sh;               // output row that contains it
  T E = (Ya + 1) * (T)sh;          // where that output row ends
  for (; k < k_end; ++k) {
    float wc, wn;
    if (k >= nrows) {              // padding up to a multiple of the unroll: contributes nothing
      wc = 0.f;
      wn = -1.f;
    } else {
      const T hi = lo + (T)dh;
      if (E >= hi) {               // row entirely inside output row Ya
        wc = (float)dh * inv_sh;
        wn = (E == hi) ? 0.f : -1.f;   // boundary exactly at the row end: emit, the next row starts clean
      } else {
        wc = (float)(E - lo) * inv_sh;
        wn = (float)(hi - E) * inv_sh;
      }
      if ((int)Ya < oy0) {         // the part that belongs to the previous unit's output row is dropped
        wc = (wn > 0.f) ? wn : 0.f;
        wn = -1.f;
      }
      lo = hi;
      if (E <= lo) {               // next row starts in the next output row
        ++Ya;
        E += (T)sh;
      }
    }
    tab[k] = make_float2(wc, wn);
  }
}

// MAXO = max number of outputs that can START inside one 4-pixel source group = ceil(4 / ratio_x):
// 2 for ratio_x > 2 (the 9x12MP config), 3 for ratio_x > 4/3, 4 for ratio_x > 1.
template <int MAXO, typename T>
__device__ __forceinline__ void stream_unit(const JobParam& P, const Region& R, int tx, int ty, WarpSmem& ws, int lane) {
  const ImgParam& im = P.img[R.img];
  const int sw = im.sx_len, sh = im.sy_len;
  const int dw = (int)im.W, dh = (int)im.H;
  const int X0 = (int)im.X0, Y0 = (int)im.Y0;
  const int ox0 = R.x0 + tx * R.tile_w, ox1 = min(ox0 + R.tile_w, R.x1);
  const int oy0 = R.y0 + ty * R.tile_h, oy1 = min(oy0 + R.tile_h, R.y1);
  const float inv_sw = 1.0f / (float)sw, inv_sh = 1.0f / (float)sh;

  // ---- vertical: source rows [rs, re) ----
  const int rs = (int)(((T)oy0 * (T)sh) / (T)dh);
  const int re = min(sh, (int)Geo<T>::cdiv((T)oy1 * (T)sh, (T)dh));
  const int nrows = re - rs;
  const int nrows_pad = (nrows + STREAM_NST - 1) / STREAM_NST * STREAM_NST;
  __syncwarp();  // every lane is done reading the previous unit's table
  build_row_table<T>(ws.tab, lane, rs, nrows, nrows_pad, sh, dh, oy0, inv_sh);

  // ---- horizontal: this lane's 4-pixel source group and the outputs that start in it ----
  const int g0 = (int)((((T)ox0 * (T)sw) / (T)dw) >> 2);
  const int g = g0 + lane;
  const T c0 = (T)4 * (T)g * (T)dw;                  // group start, in source units scaled by dw
  const int Xfirst = (int)Geo<T>::cdiv(c0, (T)sw);
  const int Xnext = (int)Geo<T>::cdiv(c0 + (T)4 * (T)dw, (T)sw);
  const int cnt = min(Xnext - Xfirst, MAXO);
  float w[MAXO][4], wprev[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const T a = c0 + (T)c * (T)dw, b = a + (T)dw;    // column c of the group
    {
      const T e = (T)Xfirst * (T)sw;                 // end of output Xfirst-1 (it started in an earlier group)
      wprev[c] = (e > a) ? (float)(min(e, b) - a) * inv_sw : 0.f;
    }
#pragma unroll
    for (int o = 0; o < MAXO; ++o) {
      const T s = (T)(Xfirst + o) * (T)sw, e = s + (T)sw;
      const T lo = max(s, a), hi = min(e, b);
      w[o][c] = (hi > lo && Xfirst + o < dw) ? (float)(hi - lo) * inv_sw : 0.f;
    }
  }
  // groups the warp needs: up to the end of the tile's last output, inside the bitmap
  const int g_end = min((int)Geo<T>::cdiv((T)ox1 * (T)sw, (T)4 * (T)dw), (sw + 3) >> 2);
  const int nact = min(32, g_end - g0);
  const bool active = lane < nact;
  bool st_ok[MAXO];
  float last[MAXO];  // 1 for the last output that starts in this group: it takes the right neighbour's tail
#pragma unroll
  for (int o = 0; o < MAXO; ++o) {
    const int X = Xfirst + o;
    st_ok[o] = active && (o < cnt) && (X >= ox0) && (X < ox1);
    last[o] = (o == cnt - 1 && lane != 31) ? 1.f : 0.f;
  }
  // columns past the bitmap edge (row padding) must not trip the alpha check
  const uint32_t m0 = (4 * g + 0 >= sw) ? 0xff000000u : 0u, m1 = (4 * g + 1 >= sw) ? 0xff000000u : 0u,
                 m2 = (4 * g + 2 >= sw) ? 0xff000000u : 0u, m3 = (4 * g + 3 >= sw) ? 0xff000000u : 0u;
  // output pointer walks with the rows: device position of this lane's output 0, +-4 bytes per output
  const int xstep = im.flip_x ? -4 : 4;
  const int64_t ystep = im.flip_y ? -P.out_pitch : P.out_pitch;
  uint8_t* optr = P.out_rgba + (int64_t)(Y0 + (im.flip_y ? dh - 1 - oy0 : oy0)) * P.out_pitch +
                  (int64_t)(X0 + (im.flip_x ? dw - 1 - Xfirst : Xfirst)) * 4;

  // ---- ring prologue: each lane stages its own 16 B per row, one commit group per row ----
  const int64_t pitch = im.pitch;
  const uint8_t* lsrc = im.data + (int64_t)rs * pitch + (int64_t)g * 16;
  const uint32_t lring = smem_u32(&ws.ring[0][lane * 16]);
#pragma unroll
  for (int u = 0; u < STREAM_NST; ++u) {
    if (active && u < nrows) cp_async16(lring + u * STREAM_STAGE_BYTES, lsrc);
    lsrc += pitch;
    cp_async_commit();
  }
  __syncwarp();  // row table complete

  float acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = 0.f;
  uint32_t and_acc = 0xffffffffu;
  const float2* tab = ws.tab;

  for (int i = 0; i < nrows_pad; i += STREAM_NST) {
#pragma unroll
    for (int u = 0; u < STREAM_NST; ++u) {
      cp_async_wait<STREAM_NST - 1>();   // row i+u has landed (this lane's own copy)
      const uint4 px = lds128(lring + u * STREAM_STAGE_BYTES);
      const float2 t = tab[i + u];
      if (active && i + u + STREAM_NST < nrows) cp_async16(lring + u * STREAM_STAGE_BYTES, lsrc);
      lsrc += pitch;
      cp_async_commit();
      and_acc &= (px.x | m0);
      and_acc &= (px.y | m1);
      and_acc &= (px.z | m2);
      and_acc &= (px.w | m3);
      float f[12];
      f[0] = expand<0>(px.x); f[1] = expand<1>(px.x); f[2] = expand<2>(px.x);
      f[3] = expand<0>(px.y); f[4] = expand<1>(px.y); f[5] = expand<2>(px.y);
      f[6] = expand<0>(px.z); f[7] = expand<1>(px.z); f[8] = expand<2>(px.z);
      f[9] = expand<0>(px.w); f[10] = expand<1>(px.w); f[11] = expand<2>(px.w);
#pragma unroll
      for (int k = 0; k < 12; ++k) acc[k] = fmaf(t.x, f[k], acc[k]);
      if (t.y >= 0.f) {
        // ---- an output row is complete: horizontal pass in registers ----
        float out[MAXO][3];
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
          float prev = wprev[0] * acc[ch];
#pragma unroll
          for (int c = 1; c < 4; ++c) prev = fmaf(wprev[c], acc[c * 3 + ch], prev);
#pragma unroll
          for (int o = 0; o < MAXO; ++o) {
            // for MAXO == 2 (ratio > 2) the second output cannot reach back into columns 0 and 1
            const int cfirst = (MAXO == 2 && o == 1) ? 2 : 0;
            out[o][ch] = w[o][cfirst] * acc[cfirst * 3 + ch];
#pragma unroll
            for (int c = cfirst + 1; c < 4; ++c) out[o][ch] = fmaf(w[o][c], acc[c * 3 + ch], out[o][ch]);
          }
          const float rcv = __shfl_down_sync(0xffffffffu, prev, 1);
#pragma unroll
          for (int o = 0; o < MAXO; ++o) out[o][ch] = fmaf(last[o], rcv, out[o][ch]);
        }
#pragma unroll
        for (int o = 0; o < MAXO; ++o) {
          if (st_ok[o]) {
            const uint32_t r = quant_bits(out[o][0]), gg = quant_bits(out[o][1]), b = quant_bits(out[o][2]);
            *reinterpret_cast<uint32_t*>(optr + o * xstep) =
                __byte_perm(__byte_perm(r, gg, 0x0040u), b, 0x0410u) | 0xff000000u;
          }
        }
        optr += ystep;
#pragma unroll
        for (int k = 0; k < 12; ++k) acc[k] = t.y * f[k];
      }
    }
  }
  // Translucent bitmaps (rare: photos are opaque): the unit is redone with the exact premultiplied
  // source-over evaluation.  Outside the hot loop so it costs the streaming path no registers.
  const bool need_fix = active && ((and_acc >> 24) != 0xffu);
  if (__any_sync(0xffffffffu, need_fix)) {
    const int ax = im.flip_x ? dw - ox1 : ox0, bx = im.flip_x ? dw - ox0 : ox1;
    const int ay = im.flip_y ? dh - oy1 : oy0, by = im.flip_y ? dh - oy0 : oy1;
    for (int y = Y0 + ay; y < Y0 + by; ++y) {
      uint32_t* row = reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch);
      for (int x = X0 + ax + lane; x < X0 + bx; x += 32) row[x] = shade_px(P, x, y);
    }
  }
}

}  // namespace stitch
