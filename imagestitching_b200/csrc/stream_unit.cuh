// The streaming area-downscale unit of the fused compose kernel (see compose_kernel.cuh).
//
// One warp, one unit = a band of 32 source groups (4 px = 16 B per lane) x tile_h output rows of
// one image.  Replaces ctx.drawImage (U:156-199) for integer-aligned downscales.
//
// Bands are laid out in GROUP space: unit tx owns the outputs that start inside groups
// [g_base + 31*tx, g_base + 31*tx + 31); lane 31 holds the next band's first group only to supply
// the tail of lane 30's last output, so adjacent bands overlap by exactly one group (3 % of the
// lanes) whatever the scale ratio is.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "stitch_plan.h"

namespace stitch {

#ifndef STREAM_NST
#define STREAM_NST 8  // ring depth: source rows in flight per lane (power of two, multiple of STREAM_UNR)
#endif
#define STREAM_UNR 4  // rows per iteration of the unrolled row loop
#define STREAM_STAGE_BYTES 512
#define STREAM_STRIDE 31  // groups whose outputs a unit stores

static_assert(STREAM_NST % STREAM_UNR == 0 && (STREAM_NST & (STREAM_NST - 1)) == 0 && STREAM_UNR % 2 == 0, "ring layout");

struct WarpSmem {
  alignas(128) uint8_t ring[STREAM_NST][STREAM_STAGE_BYTES];
  alignas(16) float2 tab[STREAM_TAB_MAX];
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
__device__ __forceinline__ float4 lds128f(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}

// ---- packed fp32x2 arithmetic (FFMA2 / FMUL2: two lanes of fp32 per issue slot) ----
// (dx, dy) += (ax, ay) * s
__device__ __forceinline__ void fma2_s(float& dx, float& dy, float ax, float ay, float s) {
  asm("{.reg .b64 ra, rb, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n mov.b64 rd, {%0,%1};\n"
      " fma.rn.f32x2 rd, ra, rb, rd;\n mov.b64 {%0,%1}, rd;}"
      : "+f"(dx), "+f"(dy)
      : "f"(ax), "f"(ay), "f"(s));
}
// (dx, dy) = (ax, ay) * s
__device__ __forceinline__ void mul2_s(float& dx, float& dy, float ax, float ay, float s) {
  asm("{.reg .b64 ra, rb, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n"
      " mul.rn.f32x2 rd, ra, rb;\n mov.b64 {%0,%1}, rd;}"
      : "=f"(dx), "=f"(dy)
      : "f"(ax), "f"(ay), "f"(s));
}
// (dx, dy) = floor-quantise: see quant_bits
__device__ __forceinline__ void quant2(uint32_t& qx, uint32_t& qy, float ax, float ay) {
  asm("{.reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n mov.b64 rc, {%5,%5};\n"
      " fma.rm.f32x2 rd, ra, rb, rc;\n mov.b64 {%0,%1}, rd;}"
      : "=r"(qx), "=r"(qy)
      : "f"(ax), "f"(ay), "f"(32768.f), "f"(8355840.5f));
}
// negated form, rounded towards +inf: bit pattern 0xCB0000kk, so byte 3 has its sign bit set
__device__ __forceinline__ void quant2_neg(uint32_t& qx, uint32_t& qy, float ax, float ay) {
  asm("{.reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n mov.b64 rc, {%5,%5};\n"
      " fma.rp.f32x2 rd, ra, rb, rc;\n mov.b64 {%0,%1}, rd;}"
      : "=r"(qx), "=r"(qy)
      : "f"(ax), "f"(ay), "f"(-32768.f), "f"(-8355840.5f));
}

// byte CH of px -> float 1 + b * 2^-15  (bit pattern 0x3F80bb00): one PRMT
template <int CH>
__device__ __forceinline__ float expand(uint32_t px) {
  return __uint_as_float(__byte_perm(px, 0x3F800000u, 0x7604u | (CH << 4)));
}
// float 1 + v * 2^-15 (v in [0,255]) -> 0x4B0000kk with kk = floor(v + 0.5): one FFMA.RM
__device__ __forceinline__ uint32_t quant_bits(float x) { return __float_as_uint(__fmaf_rd(x, 32768.f, 8355840.5f)); }
// same value negated (0xCB0000kk): FFMA.RP
__device__ __forceinline__ uint32_t quant_bits_neg(float x) { return __float_as_uint(__fmaf_ru(x, -32768.f, -8355840.5f)); }
// (r, g, b quantised as above, b in the negated form) -> 0xFFbbggrr with two PRMT: the alpha byte is
// the sign replication of b's top byte
__device__ __forceinline__ uint32_t pack_rgb(uint32_t r, uint32_t g, uint32_t bneg) {
  uint32_t v;  // __byte_perm() drops the selector's replicate-sign bit, prmt.b32 keeps it
  asm("prmt.b32 %0, %1, %2, 0xF410;" : "=r"(v) : "r"(__byte_perm(r, g, 0x0040u)), "r"(bneg));
  return v;
}

__device__ __forceinline__ uint32_t udiv_ceil(uint32_t n, uint32_t d) { return (n + d - 1) / d; }

// Builds the warp's row table for source rows [rs, rs + nrows): entry = (weight into the current
// output row, weight into the next one or -1 when the row does not complete an output row).
// Each lane fills a contiguous chunk, so it needs one division and then steps incrementally.
__device__ __forceinline__ void build_row_table(float2* tab, int lane, uint32_t rs, int nrows, int nrows_pad, uint32_t sh,
                                                uint32_t dh, uint32_t oy0, float inv_sh) {
  const int chunk = (nrows_pad + 31) >> 5;
  int k = lane * chunk;
  const int k_end = min(k + chunk, nrows_pad);
  if (k >= k_end) return;
  uint32_t lo = (rs + (uint32_t)k) * dh;  // start of row (rs+k) in source units scaled by dh
  uint32_t Ya = lo / sh;                  // output row that contains it
  uint32_t E = (Ya + 1) * sh;             // where that output row ends
  const float wfull = (float)dh * inv_sh;
  for (; k < k_end; ++k) {
    float wc, wn;
    if (k >= nrows) {  // padding up to a multiple of the unroll: contributes nothing
      wc = 0.f;
      wn = -1.f;
    } else {
      const uint32_t hi = lo + dh;
      if (E >= hi) {  // row entirely inside output row Ya
        wc = wfull;
        wn = (E == hi) ? 0.f : -1.f;  // boundary exactly at the row end: emit, the next row starts clean
      } else {
        wc = (float)(E - lo) * inv_sh;
        wn = (float)(hi - E) * inv_sh;
      }
      if (Ya < oy0) {  // the part that belongs to the previous unit's output row is dropped
        wc = (wn > 0.f) ? wn : 0.f;
        wn = -1.f;
      }
      lo = hi;
      if (E <= lo) {  // next row starts in the next output row
        ++Ya;
        E += sh;
      }
    }
    tab[k] = make_float2(wc, wn);
  }
}

// MAXO = max number of outputs that can START inside one 4-pixel source group = ceil(4 / ratio_x):
// 2 for ratio_x > 2 (the 9x12MP config), 3 for ratio_x > 4/3, 4 for ratio_x > 1.
// PARTIAL = some streamed bitmap has a width that is not a multiple of 4: the columns of its last
// group that lie in the row padding are masked out of the alpha check.
// All geometry is unsigned 32-bit; launch.cuh only admits images whose scaled coordinates fit.
template <int MAXO, bool PARTIAL>
__device__ __forceinline__ void stream_unit(const JobParam& P, const Region& R, int tx, int ty, WarpSmem& ws, int lane) {
  const ImgParam& im = P.img[R.img];
  const uint32_t sw = (uint32_t)im.sx_len, sh = (uint32_t)im.sy_len;
  const uint32_t dw = (uint32_t)im.W, dh = (uint32_t)im.H;
  const int X0 = (int)im.X0, Y0 = (int)im.Y0;
  const uint32_t oy0 = (uint32_t)(R.y0 + ty * R.tile_h), oy1 = min(oy0 + (uint32_t)R.tile_h, (uint32_t)R.y1);
  const float inv_sw = __frcp_rn((float)sw), inv_sh = __frcp_rn((float)sh);

  // ---- vertical: source rows [rs, re) ----
  const uint32_t rs = (oy0 * sh) / dh;
  const uint32_t re = min(sh, udiv_ceil(oy1 * sh, dh));
  const int nrows = (int)(re - rs);
  const int nrows_pad = (nrows + STREAM_UNR - 1) / STREAM_UNR * STREAM_UNR;
  __syncwarp();  // every lane is done reading the previous unit's table
  build_row_table(ws.tab, lane, rs, nrows, nrows_pad, sh, dh, oy0, inv_sh);

  // ---- horizontal: this lane's 4-pixel source group and the outputs that start in it ----
  const uint32_t g = (uint32_t)(R.g_base + tx * STREAM_STRIDE + lane);
  const uint32_t g_last = (sw + 3) / 4 - 1;
  const uint32_t c0 = 4u * g * dw;  // group start, in source units scaled by dw
  const int Xfirst = (int)udiv_ceil(c0, sw);
  int Xnext = __shfl_down_sync(0xffffffffu, Xfirst, 1);
  if (lane == 31) Xnext = Xfirst;  // lane 31 only feeds lane 30's tail
  const int cnt = min(Xnext - Xfirst, MAXO);
  float w[MAXO][4], wprev[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const uint32_t a = c0 + (uint32_t)c * dw, b = a + dw;  // column c of the group
    {
      const uint32_t e = (uint32_t)Xfirst * sw;  // end of output Xfirst-1 (it started in an earlier group)
      wprev[c] = (e > a) ? (float)(min(e, b) - a) * inv_sw : 0.f;
    }
#pragma unroll
    for (int o = 0; o < MAXO; ++o) {
      const uint32_t s = (uint32_t)(Xfirst + o) * sw, e = s + sw;
      const uint32_t lo = max(s, a), hi = min(e, b);
      w[o][c] = (hi > lo && (uint32_t)(Xfirst + o) < dw) ? (float)(hi - lo) * inv_sw : 0.f;
    }
  }
  bool st_ok[MAXO];
  float last[MAXO];  // 1 for the last output that starts in this group: it takes the right neighbour's tail
#pragma unroll
  for (int o = 0; o < MAXO; ++o) {
    const int X = Xfirst + o;
    st_ok[o] = (o < cnt) && (X >= R.x0) && (X < R.x1);
    last[o] = (o == cnt - 1) ? 1.f : 0.f;
  }
  uint32_t m1 = 0, m2 = 0, m3 = 0;
  if (PARTIAL) {  // columns past the bitmap edge (row padding) must not trip the alpha check
    m1 = (4 * g + 1 >= sw) ? 0xff000000u : 0u;
    m2 = (4 * g + 2 >= sw) ? 0xff000000u : 0u;
    m3 = (4 * g + 3 >= sw) ? 0xff000000u : 0u;
  }
  // output pointer walks with the rows: device position of this lane's output 0, +-4 bytes per output
  const int xstep = im.flip_x ? -4 : 4;
  const int64_t ystep = im.flip_y ? -P.out_pitch : P.out_pitch;
  uint8_t* optr = P.out_rgba + (int64_t)(Y0 + (im.flip_y ? (int)(dh - 1 - oy0) : (int)oy0)) * P.out_pitch +
                  (int64_t)(X0 + (im.flip_x ? (int)dw - 1 - Xfirst : Xfirst)) * 4;

  // ---- ring prologue: each lane stages its own 16 B per row; rows travel in pairs (one commit group
  // per pair, STREAM_NST rows deep).  Rows are addressed by a 32-bit offset clamped to the unit's last
  // row, so the refills need no predicate: past the end (and in lanes past the bitmap's last group)
  // they re-read bytes that are then weighted with zero. ----
  const uint32_t pitch = (uint32_t)im.pitch;
  const uint8_t* lbase = im.data + (int64_t)rs * im.pitch + (int64_t)min(g, g_last) * 16;
  const uint32_t off_max = (uint32_t)(nrows - 1) * pitch;
  uint32_t off = 0;
  const uint32_t lring = smem_u32(&ws.ring[0][lane * 16]);
#pragma unroll
  for (int u = 0; u < STREAM_NST; ++u) {
    cp_async16(lring + u * STREAM_STAGE_BYTES, lbase + off);
    off = min(off + pitch, off_max);
    if (u & 1) cp_async_commit();
  }
  __syncwarp();  // row table complete

  float acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = 0.f;
  uint32_t and_acc = 0xffffffffu;
  const uint32_t tab_s = smem_u32(ws.tab);
  uint32_t hoff = 0;  // byte offset of the ring part the current STREAM_UNR rows live in

  // One source row: vertical accumulate; when the row completes an output row (EMIT), the horizontal
  // pass runs in registers and the accumulators restart with the row's share of the next output row.
  auto row_step = [&](const uint4& px, const float wc, const float wn, const bool emit) {
    if (PARTIAL) {
      and_acc &= px.x;
      and_acc &= (px.y | m1);
      and_acc &= (px.z | m2);
      and_acc &= (px.w | m3);
    } else {
      and_acc &= px.x & px.y;
      and_acc &= px.z & px.w;
    }
    float f[12];
    f[0] = expand<0>(px.x); f[1] = expand<1>(px.x); f[2] = expand<2>(px.x);
    f[3] = expand<0>(px.y); f[4] = expand<1>(px.y); f[5] = expand<2>(px.y);
    f[6] = expand<0>(px.z); f[7] = expand<1>(px.z); f[8] = expand<2>(px.z);
    f[9] = expand<0>(px.w); f[10] = expand<1>(px.w); f[11] = expand<2>(px.w);
#pragma unroll
    for (int k = 0; k < 12; k += 2) fma2_s(acc[k], acc[k + 1], f[k], f[k + 1], wc);
    if (emit) {
      uint32_t q[MAXO][3];
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        float prev = wprev[0] * acc[ch];
#pragma unroll
        for (int c = 1; c < 4; ++c) prev = fmaf(wprev[c], acc[c * 3 + ch], prev);
        const float rcv = __shfl_down_sync(0xffffffffu, prev, 1);
        float out[MAXO];
        // outputs in pairs: (out[o], out[o+1]) = sum_c (w[o][c], w[o+1][c]) * acc[c]
#pragma unroll
        for (int o = 0; o + 1 < MAXO; o += 2) {
          mul2_s(out[o], out[o + 1], w[o][0], w[o + 1][0], acc[ch]);
#pragma unroll
          for (int c = 1; c < 4; ++c) fma2_s(out[o], out[o + 1], w[o][c], w[o + 1][c], acc[c * 3 + ch]);
          fma2_s(out[o], out[o + 1], last[o], last[o + 1], rcv);
          if (ch == 2) quant2_neg(q[o][ch], q[o + 1][ch], out[o], out[o + 1]);
          else quant2(q[o][ch], q[o + 1][ch], out[o], out[o + 1]);
        }
        if (MAXO & 1) {
          const int o = MAXO - 1;
          out[o] = w[o][0] * acc[ch];
#pragma unroll
          for (int c = 1; c < 4; ++c) out[o] = fmaf(w[o][c], acc[c * 3 + ch], out[o]);
          out[o] = fmaf(last[o], rcv, out[o]);
          q[o][ch] = (ch == 2) ? quant_bits_neg(out[o]) : quant_bits(out[o]);
        }
      }
#pragma unroll
      for (int o = 0; o < MAXO; ++o) {
        const uint32_t v = pack_rgb(q[o][0], q[o][1], q[o][2]);
        if (st_ok[o]) *reinterpret_cast<uint32_t*>(optr + o * xstep) = v;
      }
      optr += ystep;
#pragma unroll
      for (int k = 0; k < 12; k += 2) mul2_s(acc[k], acc[k + 1], f[k], f[k + 1], wn);
    }
  };

  for (int i0 = 0; i0 < nrows_pad; i0 += 32) {
    // which of the next 32 rows complete an output row: one ballot, so the per-row branch below is on
    // a warp-uniform mask and needs no reconvergence scaffolding around the shuffles
    const bool fl = (i0 + lane < nrows_pad) && (ws.tab[i0 + lane].y >= 0.f);
    uint32_t mask = __ballot_sync(0xffffffffu, fl);
    const int i1 = min(i0 + 32, nrows_pad);
    for (int i = i0; i < i1; i += STREAM_UNR) {
      const uint32_t ring_h = lring + hoff;
#pragma unroll
      for (int u = 0; u < STREAM_UNR; u += 2) {
        const float4 t2 = lds128f(tab_s + (uint32_t)(i + u) * 8u);  // table entries of rows i+u, i+u+1
        cp_async_wait<STREAM_NST / 2 - 1>();  // rows i+u, i+u+1 have landed (this lane's own copies)
        const uint4 pa = lds128(ring_h + u * STREAM_STAGE_BYTES);
        const uint4 pb = lds128(ring_h + (u + 1) * STREAM_STAGE_BYTES);
        cp_async16(ring_h + u * STREAM_STAGE_BYTES, lbase + off);  // rows i+u+NST, i+u+1+NST
        off = min(off + pitch, off_max);
        cp_async16(ring_h + (u + 1) * STREAM_STAGE_BYTES, lbase + off);
        off = min(off + pitch, off_max);
        cp_async_commit();
        row_step(pa, t2.x, t2.y, (mask >> u) & 1u);
        row_step(pb, t2.z, t2.w, (mask >> (u + 1)) & 1u);
      }
      mask >>= STREAM_UNR;
      hoff = (hoff + STREAM_UNR * STREAM_STAGE_BYTES) & (STREAM_NST * STREAM_STAGE_BYTES - 1);
    }
  }
  // Translucent bitmaps (rare: photos are opaque): the unit is redone with the exact premultiplied
  // source-over evaluation.  Outside the hot loop so it costs the streaming path no registers.
  const bool need_fix = (and_acc >> 24) != 0xffu;
  if (__any_sync(0xffffffffu, need_fix)) {
    const int xf0 = __shfl_sync(0xffffffffu, Xfirst, 0), xf1 = __shfl_sync(0xffffffffu, Xfirst, 31);
    const int ox0 = max(R.x0, xf0), ox1 = min(R.x1, xf1);
    const int ax = im.flip_x ? (int)dw - ox1 : ox0, bx = im.flip_x ? (int)dw - ox0 : ox1;
    const int ay = im.flip_y ? (int)(dh - oy1) : (int)oy0, by = im.flip_y ? (int)(dh - oy0) : (int)oy1;
    for (int y = Y0 + ay; y < Y0 + by; ++y) {
      uint32_t* row = reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch);
      for (int x = X0 + ax + lane; x < X0 + bx; x += 32) row[x] = shade_px(P, x, y);
    }
  }
}

}  // namespace stitch
