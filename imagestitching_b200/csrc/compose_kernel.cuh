// Fused resample + compose kernel (K1): ONE launch draws every image of a job into the canvas.
//
// Replaces the per-image ctx.drawImage loop (P:1439-1572 -> U:153-202) and the canvas
// pre-fill (P:1423-1424).  The canvas is partitioned on the host (launch.cuh) into regions,
// every region into warp-sized units; a persistent grid of warps pulls units from an atomic
// counter (heavy units first) until none are left:
//
//   REGION_STREAM  integer-aligned area downscale (1 < ratio_x <= 16, ratio_y > 1), EXIF 1-4.
//       The bandwidth path (stream_unit.cuh).  A warp owns a band of <=32 source groups
//       (4 px = 16 B per lane) and streams DOWN the source rows.  The rows are staged by TMA:
//       lane 0 requests one box of 128 px x 8 rows (cp.async.bulk.tensor.2d, one UTMALDG) per
//       stage into a per-warp shared-memory ring of 3 stages, each completing on its own mbarrier
//       (expect-tx); the warp that consumes a stage re-arms it, so no second warp and no block
//       barrier is involved.  (Round 1 staged 16 B per lane and row with cp.async / LDGSTS; the
//       first bulk attempt -- a 1-D cp.async.bulk per 512-byte row -- cost ~25 issue slots per row.)
//       Vertical pass: INTEGER.  Two rows at a time: 2 PRMT interleave their bytes channel by
//       channel, 3 IDP.2A apply both rows' 15-bit weights (a warp-uniform row-pair table in smem;
//       the weights of an output row sum to exactly 2^15).  The accumulators start at 0x4B000000, so
//       they are at the same time the floats 2^23 + v * 2^15 the horizontal pass works on.
//       Horizontal pass, in registers when an output row completes: each lane reduces its 4 columns
//       into the <=MAXO outputs that start in its group plus the tail of its left neighbour's last
//       output, which travels by one shuffle; one FFMA.RM against 2^23 - 256 + 0.5 leaves
//       floor(value + 0.5) in the low mantissa byte.  Every source byte is fetched once per warp.
//       Alpha is not filtered: the pixels are ANDed together and a unit that saw any alpha != 255 is
//       redone by the exact premultiplied gather path (photos are opaque, so this is the rare case).
//   REGION_UPSCALE bilinear enlargement of the fully covered pixels of one image (upscale_unit.cuh):
//       output-bound, 16-byte stores, source through L1.
//   REGION_FILL    white.
//   REGION_GATHER  per-pixel evaluation (gather_kernel.cuh shade()) for upscale / EXIF 5-8 /
//       ratios outside the streaming range.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "gather_kernel.cuh"
#include "stitch_plan.h"

namespace stitch {

#ifndef STREAM_MINB
#define STREAM_MINB 2  // resident CTAs per SM the register allocation targets (measured: 2 x 8 warps with 128 registers beats 3 x 8 with 80)
#endif

// Exact (gather) evaluation of one canvas pixel, kept out of line so the rare slow path does not
// inflate the streaming loop's register allocation.
__device__ __noinline__ void shade_rgb(const JobParam& P, int x, int y, float* rgba) {
  shade(P, x, y, rgba[0], rgba[1], rgba[2], rgba[3]);
}
__device__ __forceinline__ uint32_t shade_px(const JobParam& P, int x, int y) {
  float c[4];
  shade_rgb(P, x, y, c);
  return finish_px(P, c[0], c[1], c[2], c[3]);
}

// One 2x2 chroma block (bx, by) of the canvas, evaluated exactly: optional RGBA and Y for the
// block's pixels inside [x0,x1) x [y0,y1), and Cb/Cr with edge replication.
template <bool RGBA, bool LUMA, bool CHROMA>
__device__ __forceinline__ void shade_block(const JobParam& P, int bx, int by, int x0, int y0, int x1, int y1) {
  float cb = 0.f, cr = 0.f;
#pragma unroll
  for (int dy = 0; dy < 2; ++dy) {
#pragma unroll
    for (int dx = 0; dx < 2; ++dx) {
      const int xx = bx * 2 + dx, yy = by * 2 + dy;
      const int x = min(xx, P.cw - 1), y = min(yy, P.ch - 1);  // edge replication for odd sizes
      const bool inside = xx >= x0 && xx < x1 && yy >= y0 && yy < y1;
      if (!CHROMA && !inside) continue;
      float c[4];
      shade_rgb(P, x, y, c);
      if (inside) {
        if (LUMA) P.out_y[(int64_t)y * P.cw + x] = (uint8_t)q8(0.299f * c[0] + 0.587f * c[1] + 0.114f * c[2]);
        if (RGBA)
          *reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch + (int64_t)x * 4) =
              q8(c[0]) | (q8(c[1]) << 8) | (q8(c[2]) << 16) | 0xff000000u;
      }
      cb += 128.f - 0.168736f * c[0] - 0.331264f * c[1] + 0.5f * c[2];
      cr += 128.f + 0.5f * c[0] - 0.418688f * c[1] - 0.081312f * c[2];
    }
  }
  if (CHROMA) {
    const int cwh = (P.cw + 1) >> 1;
    P.out_cb[(int64_t)by * cwh + bx] = (uint8_t)q8(cb * 0.25f);
    P.out_cr[(int64_t)by * cwh + bx] = (uint8_t)q8(cr * 0.25f);
  }
}

// Exact re-evaluation of the canvas rect [x0,x1) x [y0,y1) by one warp (a streaming unit that met
// translucent pixels).  With chroma output y0/y1 are even and the unit owns the chroma blocks whose
// first pixel in source order -- even canvas x, odd when the image is mirrored -- lies in [x0,x1).
template <int MODE>
__device__ __noinline__ void redo_rect(const JobParam& P, int x0, int y0, int x1, int y1, int lane, int key_odd) {
  if (MODE == 0) {
    for (int y = y0; y < y1; ++y) {
      uint32_t* row = reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch);
      for (int x = x0 + lane; x < x1; x += 32) row[x] = shade_px(P, x, y);
    }
  } else {
    const int bx0 = (x0 - key_odd) >> 1, bx1 = (x1 - 1) >> 1;   // blocks that touch the rect
    for (int by = y0 >> 1; by < (y1 + 1) >> 1; ++by)
      for (int bx = bx0 + lane; bx <= bx1; bx += 32) {
        const int key = 2 * bx + key_odd;
        if (key >= x0 && key < x1) shade_block<MODE == 2, true, true>(P, bx, by, x0, y0, x1, y1);
        else shade_block<MODE == 2, true, false>(P, bx, by, x0, y0, x1, y1);
      }
  }
}

}  // namespace stitch

#include "stream_unit.cuh"
#include "upscale_unit.cuh"

namespace stitch {

static constexpr int kComposeSmemBytes = (int)sizeof(WarpSmem) * STREAM_WARPS;

// FILL / GATHER units.  MODE 0: the region is a device pixel rect.  With chroma output (MODE 1, 2)
// the region is a rect of 2x2 chroma BLOCKS, so no block is ever split between two units.
template <int MODE>
__device__ __forceinline__ void fill_unit(const JobParam& P, const Region& R, int tx, int ty, int lane) {
  const int x0 = R.x0 + tx * R.tile_w, x1 = min(x0 + R.tile_w, R.x1);
  const int y0 = R.y0 + ty * R.tile_h, y1 = min(y0 + R.tile_h, R.y1);
  if (MODE == 0) {
    const uint32_t bg = P.transparent ? 0u : 0xffffffffu;  // clearRect (P:1598) or '#ffffff' (P:1423-1424)
    for (int y = y0; y < y1; ++y) {
      uint32_t* row = reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch);
      for (int x = x0 + lane; x < x1; x += 32) row[x] = bg;
    }
  } else {
    const int cwh = (P.cw + 1) >> 1;
    for (int by = y0; by < y1; ++by)
      for (int bx = x0 + lane; bx < x1; bx += 32) {
        P.out_cb[(int64_t)by * cwh + bx] = 128;
        P.out_cr[(int64_t)by * cwh + bx] = 128;
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          const int x = 2 * bx + (d & 1), y = 2 * by + (d >> 1);
          if (x < P.cw && y < P.ch) {
            P.out_y[(int64_t)y * P.cw + x] = 255;
            if (MODE == 2) *reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch + (int64_t)x * 4) = 0xffffffffu;
          }
        }
      }
  }
}

template <int MODE>
__device__ __forceinline__ void gather_unit(const JobParam& P, const Region& R, int tx, int ty, int lane) {
  const int x0 = R.x0 + tx * R.tile_w, x1 = min(x0 + R.tile_w, R.x1);
  const int y0 = R.y0 + ty * R.tile_h, y1 = min(y0 + R.tile_h, R.y1);
  if (MODE == 0) {
    for (int y = y0; y < y1; ++y) {
      uint32_t* row = reinterpret_cast<uint32_t*>(P.out_rgba + (int64_t)y * P.out_pitch);
      for (int x = x0 + lane; x < x1; x += 32) row[x] = shade_px(P, x, y);
    }
  } else {
    for (int by = y0; by < y1; ++by)
      for (int bx = x0 + lane; bx < x1; bx += 32) shade_block<MODE == 2, true, true>(P, bx, by, 0, 0, P.cw, P.ch);
  }
}

// sched[0] = next unit to hand out, sched[1] = warps that have left the kernel; the last warp out
// resets both, so back-to-back launches on one stream need no memset in between.
//
// MAXO = max over the job's streaming regions of ceil(4 / ratio_x); HOPS = max of ceil(ratio_x / 4)
// (1, 2 or 4); MODE = STREAM_MODE_* (RGBA canvas, I420 planes, or both).
// The persistent unit loop shared by both kernels.  STREAM = false compiles the streaming unit out.
template <int MAXO, int HOPS, int MODE, bool STREAM>
__device__ __forceinline__ void compose_loop(const JobParam& P, const Plan& plan, const TmaMaps* maps, unsigned int* __restrict__ sched,
                                             WarpSmem* ws) {
  const int lane = threadIdx.x & 31;
  WarpState wst;
  // One ticket per unit, drawn when the previous unit is done.  (Drawing the next ticket ahead of a short
  // unit to hide the atomic's round trip was measured slower on every workload: a claimed unit waits.)
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
    if (STREAM && R.kind == REGION_STREAM) {
      stream_unit<MAXO, HOPS, MODE>(P, *maps, R, tx, ty, *ws, wst, lane);
    } else if (R.kind == REGION_UPSCALE) {
      // window of source pixels per lane and row (upscale_unit.cuh): launch.cuh picks 4 for scales >= 1.5; the
      // kernel with streaming units compiles the general width only
      if (!STREAM && R.g_base == 4) upscale_unit<MODE, 4>(P, R, tx, ty, lane);
      else upscale_unit<MODE, 5>(P, R, tx, ty, lane);
    } else if (R.kind == REGION_FILL) {
      fill_unit<MODE>(P, R, tx, ty, lane);
    } else {
      gather_unit<MODE>(P, R, tx, ty, lane);
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

template <int MAXO, int HOPS, int MODE>
__global__ void __launch_bounds__(STREAM_WARPS * 32, STREAM_MINB)
compose_kernel(const __grid_constant__ JobParam P, const __grid_constant__ Plan plan, const __grid_constant__ TmaMaps maps,
               unsigned int* __restrict__ sched) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpSmem& ws = reinterpret_cast<WarpSmem*>(smem_raw)[warp];
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STREAM_TMA_STAGES; ++s) mbar_init(smem_u32(&ws.full[s]), 1);
    fence_proxy_async();   // the barriers are visible to the TMA unit before the first box is issued
  }
  // Lanes that never copy (past the band) and padded rows read the ring as it is: give it defined,
  // opaque contents once so the alpha check cannot trip on uninitialised shared memory.
#pragma unroll
  for (int s = 0; s < STREAM_NST; ++s)
    *reinterpret_cast<uint4*>(&ws.ring[s][lane * 16]) = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
  __syncwarp();
  compose_loop<MAXO, HOPS, MODE, true>(P, plan, &maps, sched, &ws);
}

// Jobs without a streaming region (every image enlarged, transposed or unaligned: BASELINE configs[0], previews):
// no shared memory and a smaller register budget, so more warps are resident for the latency-bound units.
#ifndef LIGHT_MINB
#define LIGHT_MINB 4
#endif
template <int MODE>
__global__ void __launch_bounds__(STREAM_WARPS * 32, LIGHT_MINB)
compose_light_kernel(const __grid_constant__ JobParam P, const __grid_constant__ Plan plan, unsigned int* __restrict__ sched) {
  compose_loop<2, 1, MODE, false>(P, plan, nullptr, sched, nullptr);
}

}  // namespace stitch
