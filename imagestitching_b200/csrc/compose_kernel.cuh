// Fused resample + compose kernel (K1): ONE launch draws every image of a job into the canvas.
//
// Replaces the per-image ctx.drawImage loop (P:1439-1572 -> U:153-202) and the canvas
// pre-fill (P:1423-1424).  The canvas is partitioned on the host (launch.cuh) into regions,
// every region into warp-sized units; a persistent grid of warps pulls units from an atomic
// counter (heavy units first) until none are left:
//
//   REGION_STREAM  integer-aligned area downscale (1 < ratio_x <= 4, ratio_y > 1), EXIF 1-4.
//       The bandwidth path (stream_unit.cuh).  A warp owns a band of <=32 source groups
//       (4 px = 16 B per lane) and streams DOWN the source rows.  Each lane stages its own
//       16 B per row into a per-warp shared-memory ring with cp.async (LDGSTS, L1 bypassed),
//       STREAM_NST rows deep, so the bytes in flight cost no registers and no barrier.
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
//       (A cp.async.bulk + mbarrier ring fed by one elected lane was measured too: 0.164 ms vs
//       0.127 ms per 9x12MP job -- the elect/R2UR/UBLKCP sequence costs ~25 issue slots per
//       512-byte row, and the kernel is issue-bound, not latency-bound.  See DESIGN.md.)
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
#define STREAM_MINB 3  // resident CTAs per SM the register allocation targets
#endif

// Exact (gather) evaluation of one canvas pixel, kept out of line so the rare slow path does not
// inflate the streaming loop's register allocation.
__device__ __noinline__ uint32_t shade_px(const JobParam& P, int x, int y) {
  float Rr, Gg, Bb;
  shade(P, x, y, Rr, Gg, Bb);
  return q8(Rr) | (q8(Gg) << 8) | (q8(Bb) << 16) | 0xff000000u;
}

}  // namespace stitch

#include "stream_unit.cuh"

namespace stitch {

static constexpr int kComposeSmemBytes = (int)sizeof(WarpSmem) * STREAM_WARPS;

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
// MAXO = max over the job's streaming regions of ceil(4 / ratio_x); PARTIAL = some streamed bitmap's
// width is not a multiple of 4 (stream_unit.cuh).
template <int MAXO, bool PARTIAL>
__global__ void __launch_bounds__(STREAM_WARPS * 32, STREAM_MINB)
compose_kernel(const __grid_constant__ JobParam P, const __grid_constant__ Plan plan, unsigned int* __restrict__ sched) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpSmem& ws = reinterpret_cast<WarpSmem*>(smem_raw)[warp];
  // Lanes that never copy (past the band) and padded rows read the ring as it is: give it defined,
  // opaque contents once so the alpha check cannot trip on uninitialised shared memory.
#pragma unroll
  for (int s = 0; s < STREAM_NST; ++s)
    *reinterpret_cast<uint4*>(&ws.ring[s][lane * 16]) = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
  __syncwarp();
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
      stream_unit<MAXO, PARTIAL>(P, R, tx, ty, ws, lane);
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
