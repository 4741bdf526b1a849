// Fused resample + compose kernel (K1): ONE launch draws every image of a job into the canvas.
//
// Replaces the per-image ctx.drawImage loop (P:1439-1572 -> U:153-202) and the canvas
// pre-fill (P:1423-1424).  The canvas is partitioned on the host (launch.cuh) into regions,
// every region into warp-sized units; a persistent grid of warps pulls units from an atomic
// counter (heavy units first) until none are left:
//
//   REGION_STREAM  integer-aligned area downscale (1 < ratio_x <= 4, ratio_y > 1), EXIF 1-4.
//       The bandwidth path.  A warp owns a band of <=32 source groups (4 px = 16 B each) and
//       streams DOWN the source rows.  Row segments (<=512 B) are staged into a per-warp
//       shared-memory ring by the bulk-copy engine (cp.async.bulk + mbarrier complete_tx, one
//       elected lane issues), STREAM_NST rows deep, so the bytes in flight do not cost
//       registers; each lane then reads its own 16 B (conflict-free LDS.128).
//       Vertical area weights are warp-uniform (row table in smem); the horizontal pass runs in
//       registers when an output row completes: each lane reduces its 4 columns into the
//       <=MAXO outputs that start in its group plus the tail of its left neighbour's last
//       output, which travels by one shuffle.  Every source byte is fetched once per warp.
//       Arithmetic: a byte b is expanded with ONE PRMT to the float 1 + b*2^-15 (0x3F80bb00);
//       because the filter weights sum to 1 the accumulators stay in that form, and one
//       FFMA.RM against 2^23 - 2^15 + 0.5 leaves floor(value + 0.5) in the low mantissa byte.
//       Resolution of the form is 2^-8 LSB per rounding.  Alpha is not filtered: the pixels are
//       ANDed together and a unit that saw any alpha != 255 is redone by the exact
//       premultiplied gather path (photos are opaque, so this is the rare case).
//   REGION_FILL    white.
//   REGION_GATHER  per-pixel evaluation (gather_kernel.cuh shade()) for upscale / EXIF 5-8 /
//       ratios outside the streaming range.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "gather_kernel.cuh"
#include "stitch_plan.h"

namespace stitch {

#ifndef STREAM_NST
#define STREAM_NST 8  // ring depth: source rows in flight per warp
#endif
#ifndef STREAM_MINB
#define STREAM_MINB 3  // resident CTAs per SM the register allocation targets
#endif
#ifndef STREAM_USE_BULK
#define STREAM_USE_BULK 0  // 1: cp.async.bulk + mbarrier ring (one elected lane); 0: per-lane cp.async ring
#endif
#define STREAM_STAGE_BYTES 512

struct WarpSmem {
  alignas(128) uint8_t ring[STREAM_NST][STREAM_STAGE_BYTES];
  alignas(8) unsigned long long bar[STREAM_NST];
  alignas(8) float2 tab[STREAM_TAB_MAX];
};
static constexpr int kComposeSmemBytes = (int)sizeof(WarpSmem) * STREAM_WARPS;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk copy (TMA engine, UBLKCP); completion is signalled on the mbarrier.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// Per-lane 16-byte async copy global -> shared (LDGSTS).  A lane only ever refills the slot it has
// itself just read with LDS; both go through the lane's LSU queue in program order and the copy's
// shared-memory write lands a global-memory latency later, so no barrier is needed.
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global.L2::128B [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
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

// floor(n / d) for 0 <= n < 2^52, d > 0: one fp64 divide plus an exact fix-up (cheaper than the
// 64-bit integer division sequence).
__device__ __forceinline__ int64_t floor_div(int64_t n, int64_t d) {
  int64_t q = (int64_t)floor((double)n / (double)d);
  if (q * d > n) --q;
  if ((q + 1) * d <= n) ++q;
  return q;
}
__device__ __forceinline__ int64_t ceil_div(int64_t n, int64_t d) { return floor_div(n + d - 1, d); }

// `it` counts the rows this warp has pushed through its ring since kernel start (stage = it % NST,
// mbarrier parity = (it / NST) & 1), so the ring carries over from unit to unit.
template <int MAXO>
__device__ __forceinline__ void stream_unit(const JobParam& P, const Region& R, int tx, int ty, WarpSmem& ws, int lane,
                                            uint32_t& it) {
  const ImgParam& im = P.img[R.img];
  const int sw = im.sx_len, sh = im.sy_len;
  const int dw = (int)im.W, dh = (int)im.H;
  const int X0 = (int)im.X0, Y0 = (int)im.Y0;
  const int ox0 = R.x0 + tx * R.tile_w, ox1 = min(ox0 + R.tile_w, R.x1);
  const int oy0 = R.y0 + ty * R.tile_h, oy1 = min(oy0 + R.tile_h, R.y1);
  const float inv_sw = 1.0f / (float)sw, inv_sh = 1.0f / (float)sh;

  // ---- vertical: source rows [rs, re) and the warp's row table ----
  const int rs = (int)floor_div((int64_t)oy0 * sh, dh);
  const int re = (int)min((int64_t)sh, ceil_div((int64_t)oy1 * sh, dh));
  const int nrows = re - rs;
  float2* tab = ws.tab;
  for (int k = lane; k < nrows; k += 32) {
    const int64_t i = rs + k;
    const int64_t lo = i * dh, hi = lo + dh;           // row i, in source units scaled by dh
    const int64_t Ya = floor_div(lo, sh);              // output row containing the start of row i
    const int64_t E = (Ya + 1) * sh;                   // where output row Ya ends
    float wc, wn;
    if (E >= hi) {                                     // row entirely inside output Ya
      wc = (float)dh * inv_sh;
      wn = (E == hi) ? 0.f : -1.f;                     // boundary exactly at the row end: emit, next starts clean
    } else {
      wc = (float)(E - lo) * inv_sh;
      wn = (float)(hi - E) * inv_sh;
    }
    if (Ya < oy0) {                                    // the part that belongs to the previous unit's row is dropped
      wc = (wn > 0.f) ? wn : 0.f;
      wn = -1.f;
    }
    tab[k] = make_float2(wc, wn);
  }

  // ---- horizontal: this lane's 4-pixel source group and the outputs that start in it ----
  const int g0 = (int)(floor_div((int64_t)ox0 * sw, dw) >> 2);
  const int g = g0 + lane;
  const int64_t c0 = (int64_t)4 * g * dw;              // group start, in source units scaled by dw
  const int Xfirst = (int)ceil_div(c0, sw);
  const int Xnext = (int)ceil_div(c0 + (int64_t)4 * dw, sw);
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
  // groups the warp needs: up to the end of the tile's last output, inside the bitmap
  const int g_end = (int)min(ceil_div((int64_t)ox1 * sw, (int64_t)4 * dw), (int64_t)((sw + 3) >> 2));
  const int nact = min(32, g_end - g0);
  const bool active = lane < nact;
  const uint32_t row_bytes = (uint32_t)nact * 16u;
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

  const uint8_t* src = im.data + (int64_t)rs * im.pitch + (int64_t)g0 * 16;
  __syncwarp();  // row table visible; every lane is done with the previous unit's ring reads
#if STREAM_USE_BULK
  if (lane == 0) {
    const int pre = min(nrows, STREAM_NST);
    for (int u = 0; u < pre; ++u) {
      const uint32_t s = (it + u) % STREAM_NST;
      mbar_expect_tx(&ws.bar[s], row_bytes);
      bulk_g2s(ws.ring[s], src + (int64_t)u * im.pitch, row_bytes, &ws.bar[s]);
    }
  }
#else
  // each lane stages its own 16 B per row: STREAM_NST rows in flight, one commit group per row
  const uint8_t* lsrc = src + lane * 16;
  uint8_t* lring = &ws.ring[0][lane * 16];
  uint32_t stage = 0;
#pragma unroll
  for (int u = 0; u < STREAM_NST; ++u) {
    if (active && u < nrows) cp_async16(lring + u * STREAM_STAGE_BYTES, lsrc);
    lsrc += im.pitch;
    cp_async_commit();
  }
#endif

  float acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = 0.f;
  uint32_t and_acc = 0xffffffffu;

  for (int i = 0; i < nrows; ++i) {
#if STREAM_USE_BULK
    const uint32_t s = it % STREAM_NST;
    mbar_wait(&ws.bar[s], (it / STREAM_NST) & 1u);
    const uint4 px = *reinterpret_cast<const uint4*>(&ws.ring[s][lane * 16]);
    const float2 t = tab[i];
    __syncwarp();  // all lanes hold their 16 B: the stage can be refilled
    if (lane == 0 && i + STREAM_NST < nrows) {
      mbar_expect_tx(&ws.bar[s], row_bytes);
      bulk_g2s(ws.ring[s], src + (int64_t)(i + STREAM_NST) * im.pitch, row_bytes, &ws.bar[s]);
    }
    ++it;
#else
    cp_async_wait<STREAM_NST - 1>();   // row i has landed (this lane's own copy; no cross-lane hazard)
    uint8_t* slot = lring + stage * STREAM_STAGE_BYTES;
    const uint4 px = *reinterpret_cast<const uint4*>(slot);
    const float2 t = tab[i];
    if (active && i + STREAM_NST < nrows) cp_async16(slot, lsrc);
    lsrc += im.pitch;
    cp_async_commit();
    stage = (stage + 1 == STREAM_NST) ? 0u : stage + 1;
#endif
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
          out[o][ch] = w[o][0] * acc[ch];
#pragma unroll
          for (int c = 1; c < 4; ++c) out[o][ch] = fmaf(w[o][c], acc[c * 3 + ch], out[o][ch]);
        }
        const float rcv = __shfl_down_sync(0xffffffffu, prev, 1);
#pragma unroll
        for (int o = 0; o < MAXO; ++o) out[o][ch] = fmaf(last[o], rcv, out[o][ch]);
      }
#pragma unroll
      for (int o = 0; o < MAXO; ++o) {
        if (st_ok[o]) {
          const uint32_t r = __float_as_uint(__fmaf_rd(out[o][0], 32768.f, 8355840.5f));
          const uint32_t gg = __float_as_uint(__fmaf_rd(out[o][1], 32768.f, 8355840.5f));
          const uint32_t b = __float_as_uint(__fmaf_rd(out[o][2], 32768.f, 8355840.5f));
          *reinterpret_cast<uint32_t*>(optr + o * xstep) =
              __byte_perm(__byte_perm(r, gg, 0x0040u), b, 0x0410u) | 0xff000000u;
        }
      }
      optr += ystep;
#pragma unroll
      for (int k = 0; k < 12; ++k) acc[k] = t.y * f[k];
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
    for (int x = x0 + lane; x < x1; x += 32) row[x] = shade_px(P, x, y);
  }
}

// sched[0] = next unit to hand out, sched[1] = warps that have left the kernel; the last warp out
// resets both, so back-to-back launches on one stream need no memset in between.
//
// MAXO = max over the job's streaming regions of ceil(4 / ratio_x): 2 for ratio > 2 (the 9x12MP
// config), 3 for ratio > 4/3, 4 for ratio > 1.
template <int MAXO>
__global__ void __launch_bounds__(STREAM_WARPS * 32, STREAM_MINB)
compose_kernel(const __grid_constant__ JobParam P, const __grid_constant__ Plan plan, unsigned int* __restrict__ sched) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpSmem& ws = reinterpret_cast<WarpSmem*>(smem_raw)[warp];
  if (lane == 0) {
    for (int s = 0; s < STREAM_NST; ++s) mbar_init(&ws.bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  uint32_t it = 0;
  for (;;) {
    unsigned int unit = 0;
    if (lane == 0) unit = atomicAdd(&sched[0], 1u);
    unit = __shfl_sync(0xffffffffu, unit, 0);
    if (unit >= (unsigned int)plan.n_units) break;
    int ri = 0;
    while (ri + 1 < plan.n_regions && (int)unit >= plan.r[ri + 1].first_unit) ++ri;
    const Region& R = plan.r[ri];
    const int local = (int)unit - R.first_unit;
    const int ty = local / R.tiles_x, tx = local - ty * R.tiles_x;
    if (R.kind == REGION_STREAM) {
      stream_unit<MAXO>(P, R, tx, ty, ws, lane, it);
    } else if (R.kind == REGION_FILL) {
      fill_unit(P, R, tx, ty, lane);
    } else {
      gather_unit(P, R, tx, ty, lane);
    }
  }
  if (lane == 0) {
    __threadfence();
    const unsigned int total = gridDim.x * STREAM_WARPS;
    if (atomicAdd(&sched[1], 1u) == total - 1) {
      sched[0] = 0;
      sched[1] = 0;
      __threadfence();
    }
  }
}

}  // namespace stitch
