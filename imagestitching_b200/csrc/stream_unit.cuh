// The streaming area-downscale unit of the fused compose kernel (see compose_kernel.cuh).
//
// One warp, one unit = a band of 32 source groups (4 px = 16 B per lane) x tile_h output rows of
// one image.  Replaces ctx.drawImage (U:156-199) for integer-aligned downscales.
//
// Bands are laid out in GROUP space: unit tx owns the outputs that start inside groups
// [g_base + S*tx, g_base + S*tx + S), S = STREAM_STRIDE_OF(MODE, HOPS); the last 32 - S lanes hold the
// next band's first groups only to supply the tails of the outputs that reach into them (for the RGBA
// canvas and ratio_x <= 4: S = 31, one overlap group = 3 % of the lanes whatever the scale ratio is).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "stitch_plan.h"
#include "tma_maps.h"

namespace stitch {

// Staging by TMA: the warp's ring holds STREAM_TMA_STAGES stages of STREAM_TMA_ROWS source rows (512 B each); a stage is
// filled by ONE cp.async.bulk.tensor.2d box (UTMALDG) that lane 0 issues when the warp has finished reading the
// stage's previous contents, and completes on the stage's mbarrier (expect-tx bytes).  (Round 1 staged with per-lane
// cp.async / LDGSTS, 16 B per lane and row: 59.8 M warp instructions per 9x12 MP job against 57.3 M with the boxes,
// 89.3 against 87.6 us on the same box; profiles/r2_compose_staging_ab.md.)
#define STREAM_NST (STREAM_TMA_ROWS * STREAM_TMA_STAGES)
#define STREAM_UNR 4  // rows between two possible ends of a unit's row range (nrows is padded to a multiple)
#define STREAM_STAGE_BYTES 512
// Output modes of the fused launch
#define STREAM_MODE_RGBA 0  // RGBA8 canvas
#define STREAM_MODE_YUV 1   // planar I420 only (U:205-242 export feeding a JPEG encoder)
#define STREAM_MODE_BOTH 2  // RGBA8 canvas and I420
// HOPS = how many groups to the right an output can reach: 1 for ratio_x <= 4, 2 for <= 8, 4 for <= 16.
// Groups whose outputs a unit stores: 32 - HOPS (the last HOPS lanes only supply tails); 32 - 2*HOPS
// when chroma is produced, so that the chroma partner of the last stored output is complete too.
#define STREAM_STRIDE_OF(MODE, HOPS) ((MODE) == STREAM_MODE_RGBA ? 32 - (HOPS) : 32 - 2 * (HOPS))

static_assert(STREAM_TMA_ROWS == 8 && STREAM_UNR == 4, "the row loop is written for stages of 8 rows = 4 row pairs");

struct WarpSmem {
  alignas(128) uint8_t ring[STREAM_NST][STREAM_STAGE_BYTES];
  alignas(16) uint4 tab[STREAM_TAB_MAX / 2];      // row-PAIR table (build_pair_table)
  alignas(8) uint64_t full[STREAM_TMA_STAGES];   // one mbarrier per stage: completes when the stage's bytes have landed
};
// what a warp carries from unit to unit
struct WarpState {
  uint32_t slot = 0;     // ring slot the next stage goes to (slots are used round-robin across units)
  uint32_t parity = 0;   // parity of that slot's next mbarrier completion: flips when the slot index wraps
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

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

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Waits until the phase `parity` of the barrier has completed.  Every lane probes (try_wait suspends the thread for a
// bounded time when the phase is still open) and the loop exits on a VOTE, so the exit is warp-uniform by construction:
// with a per-lane exit branch the compiler has to assume the warp may have diverged and wraps every shuffle / vote of
// the row loop in divergence checks (BSSY / BRA.DIV: +4 instructions per source row, measured).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!__all_sync(0xffffffffu, ok != 0u));
}
// generic-proxy writes to shared memory (the ring scrub) ordered before later async-proxy (TMA) writes
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// One box of the tensor `map` at pixel column x, row y -> shared memory at dst; completes on `bar`.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int x, int y, uint32_t bar) {
#if STREAM_TMA_EVICT_FIRST
  // source rows are read exactly once: let them leave the L2 first, so that the canvas lines stay until they are full
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;"
               ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar), "l"(pol) : "memory");
#else
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar) : "memory");
#endif
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
// ---- the "1 + b * 2^-15" number form (upscale_unit.cuh): a byte becomes a float with ONE PRMT ----
// byte CH of px -> float 1 + b * 2^-15  (bit pattern 0x3F80bb00)
template <int CH>
__device__ __forceinline__ float expand(uint32_t px) {
  return __uint_as_float(__byte_perm(px, 0x3F800000u, 0x7604u | (CH << 4)));
}
// (qx, qy) = 0x4B0000kk with kk = floor(v + 0.5) for x = 1 + v * 2^-15: x * 2^15 + (2^23 - 2^15 + 0.5), rounded down
__device__ __forceinline__ void quant2_one(uint32_t& qx, uint32_t& qy, float ax, float ay) {
  asm("{.reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n mov.b64 rc, {%5,%5};\n"
      " fma.rm.f32x2 rd, ra, rb, rc;\n mov.b64 {%0,%1}, rd;}"
      : "=r"(qx), "=r"(qy)
      : "f"(ax), "f"(ay), "f"(32768.f), "f"(8355840.5f));
}
// negated form, rounded towards +inf: bit pattern 0xCB0000kk, so byte 3 has its sign bit set
__device__ __forceinline__ void quant2_neg_one(uint32_t& qx, uint32_t& qy, float ax, float ay) {
  asm("{.reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n mov.b64 rc, {%5,%5};\n"
      " fma.rp.f32x2 rd, ra, rb, rc;\n mov.b64 {%0,%1}, rd;}"
      : "=r"(qx), "=r"(qy)
      : "f"(ax), "f"(ay), "f"(-32768.f), "f"(-8355840.5f));
}
__device__ __forceinline__ uint32_t quant_bits_one(float x) { return __float_as_uint(__fmaf_rd(x, 32768.f, 8355840.5f)); }

// ---- number form of the streaming unit ----
// The vertical pass accumulates INTEGERS: pixel bytes times 15-bit row weights
// that sum to exactly 2^15 per output row (IDP.2A, two rows per instruction), into registers that start at
// 0x4B000000 -- so an accumulator is at the same time the bit pattern of the float 2^23 + s, s = v * 2^15 < 2^23 for
// a channel value v in [0, 255], and the horizontal pass (float weights that sum to 1) works on it without any
// conversion instruction.  Resolution 2^-15 LSB.
#define STREAM_ACC_BASE 0x4B000000u
// (qx, qy) = 0x4B0000kk with kk = floor(v + 0.5): one FFMA2.RM -- x * 2^-15 + (2^23 - 256 + 0.5) = 2^23 + v + 0.5, rounded down
__device__ __forceinline__ void quant2(uint32_t& qx, uint32_t& qy, float ax, float ay) {
  asm("{.reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n mov.b64 rc, {%5,%5};\n"
      " fma.rm.f32x2 rd, ra, rb, rc;\n mov.b64 {%0,%1}, rd;}"
      : "=r"(qx), "=r"(qy)
      : "f"(ax), "f"(ay), "f"(3.0517578125e-05f), "f"(8388352.5f));
}
// negated form, rounded towards +inf: bit pattern 0xCB0000kk, so byte 3 has its sign bit set
__device__ __forceinline__ void quant2_neg(uint32_t& qx, uint32_t& qy, float ax, float ay) {
  asm("{.reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n mov.b64 rc, {%5,%5};\n"
      " fma.rp.f32x2 rd, ra, rb, rc;\n mov.b64 {%0,%1}, rd;}"
      : "=r"(qx), "=r"(qy)
      : "f"(ax), "f"(ay), "f"(-3.0517578125e-05f), "f"(-8388352.5f));
}
// float 2^23 + v * 2^15 (v in [0,255]) -> 0x4B0000kk with kk = floor(v + 0.5): one FFMA.RM
__device__ __forceinline__ uint32_t quant_bits(float x) { return __float_as_uint(__fmaf_rd(x, 3.0517578125e-05f, 8388352.5f)); }
// same value negated (0xCB0000kk): FFMA.RP
__device__ __forceinline__ uint32_t quant_bits_neg(float x) { return __float_as_uint(__fmaf_ru(x, -3.0517578125e-05f, -8388352.5f)); }
// (r, g, b quantised as above, b in the negated form) -> 0xFFbbggrr with two PRMT: the alpha byte is
// the sign replication of b's top byte
__device__ __forceinline__ uint32_t pack_rgb(uint32_t r, uint32_t g, uint32_t bneg) {
  uint32_t v;  // __byte_perm() drops the selector's replicate-sign bit, prmt.b32 keeps it
  asm("prmt.b32 %0, %1, %2, 0xF410;" : "=r"(v) : "r"(__byte_perm(r, g, 0x0040u)), "r"(bneg));
  return v;
}

__device__ __forceinline__ uint32_t udiv_ceil(uint32_t n, uint32_t d) { return (n + d - 1) / d; }

// byte store under a predicate: keeps the value's arithmetic outside of any branch
__device__ __forceinline__ void st_u8_if(uint8_t* p, uint32_t v, bool ok) {
  asm volatile("{.reg .pred q;\n setp.ne.u32 q, %2, 0;\n @q st.global.u8 [%0], %1;}" ::"l"(p), "r"(v), "r"((uint32_t)ok) : "memory");
}

// Builds the warp's row-PAIR table for source rows [rs, rs + nrows), pair p = rows (2p, 2p+1) = (A, B).
// Entry = {W1, W2, W3, flags}; a W packs two 15-bit weights (row A in the low half, row B in the high half) for one
// IDP.2A per channel, and the pair is applied as
//     acc += W1 . (A, B);   if (flags != 0) { EMIT; acc = W2 . (A, B); }   if (flags == 3) { EMIT; acc = W3 . (A, B); }
// flags bit 0 / 1: row A / B completes an output row (its remainder, and every later row, belongs to the next one).
// The weights of an output row are differences of F(t) = round(t * 2^15 / sh) at the row boundaries inside it, so they
// sum to exactly 2^15 whatever the rounding does (a constant image stays exact), each within 2^-16 of the true area.
// Each lane fills a contiguous chunk, so it needs one division and then steps incrementally.
__device__ __forceinline__ void build_pair_table(uint4* tab, int lane, uint32_t rs, int nrows, int npairs, uint32_t sh, uint32_t dh,
                                                 uint32_t oy0, float q15) {
  const int chunk = (npairs + 31) >> 5;
  int p = lane * chunk;
  const int p_end = min(p + chunk, npairs);
  if (p >= p_end) return;
  uint32_t lo = (rs + 2u * (uint32_t)p) * dh;  // start of row (rs + 2p) in source units scaled by dh
  uint32_t Ya = lo / sh;                       // output row that contains it
  uint32_t base = Ya * sh;                     // where that output row starts
  auto F = [&](uint32_t t) { return (uint32_t)__float2int_rd(fmaf((float)t, q15, 0.5f)); };
  int k = 2 * p;
  auto row = [&](uint32_t& wc, uint32_t& wn, uint32_t& e) {
    wc = 0u; wn = 0u; e = 0u;
    if (k < nrows) {   // rows past the end are padding: no weight, no emit
      const uint32_t hi = lo + dh, E = base + sh;
      if (E >= hi) {   // row entirely inside output row Ya
        wc = F(hi - base) - F(lo - base);
        e = (E == hi) ? 1u : 0u;   // boundary exactly at the row end: emit, the next row starts clean
      } else {
        wc = 32768u - F(lo - base);
        wn = F(hi - E);
        e = 1u;
      }
      if (Ya < oy0) {  // the part that belongs to the previous unit's output row is dropped
        wc = wn;
        wn = 0u;
        e = 0u;
      }
      lo = hi;
      if (E <= lo) {   // next row starts in the next output row
        ++Ya;
        base = E;
      }
    }
    ++k;
  };
  for (; p < p_end; ++p) {
    uint32_t ca, na, ea, cb, nb, eb;
    row(ca, na, ea);
    row(cb, nb, eb);
    uint4 t;
    t.x = ca | ((ea ? 0u : cb) << 16);
    t.y = ea ? (na | (cb << 16)) : (nb << 16);
    t.z = nb << 16;
    t.w = ea | (eb << 1);
    tab[p] = t;
  }
}

// MAXO = max number of outputs that can START inside one 4-pixel source group = ceil(4 / ratio_x):
// 2 for ratio_x > 2 (the 9x12MP config), 3 for ratio_x > 4/3, 4 for ratio_x > 1.
// HOPS: see STREAM_STRIDE_OF; for ratio_x > 4 a group holds at most one output start and whole groups
// can lie inside one output, whose owner then collects the `prev` sums of up to HOPS lanes to its right.
// MODE = STREAM_MODE_*.  With chroma output the host aligns the region to even canvas coordinates and
// makes tile_h even, so every 2x2 chroma block lies inside one unit (launch.cuh).
// All geometry is unsigned 32-bit; launch.cuh only admits images whose scaled coordinates fit.
template <int MAXO, int HOPS, int MODE>
__device__ __forceinline__ void stream_unit(const JobParam& P, const TmaMaps& maps, const Region& R, int tx, int ty, WarpSmem& ws,
                                            WarpState& wst, int lane) {
  const ImgParam& im = P.img[R.img];
  const uint32_t sw = (uint32_t)im.sx_len, sh = (uint32_t)im.sy_len;
  const uint32_t dw = (uint32_t)im.W, dh = (uint32_t)im.H;
  const int X0 = (int)im.X0, Y0 = (int)im.Y0;
  const uint32_t oy0 = (uint32_t)(R.y0 + ty * R.tile_h), oy1 = min(oy0 + (uint32_t)R.tile_h, (uint32_t)R.y1);
  const float inv_sw = __frcp_rn((float)sw);

  // ---- vertical: source rows [rs, re) ----
  const uint32_t rs = (oy0 * sh) / dh;
  const uint32_t re = min(sh, udiv_ceil(oy1 * sh, dh));
  const int nrows = (int)(re - rs);
  const int nrows_pad = (nrows + STREAM_UNR - 1) / STREAM_UNR * STREAM_UNR;
  constexpr int STRIDE = STREAM_STRIDE_OF(MODE, HOPS);
  __syncwarp();  // every lane is done reading the previous unit's table and ring
  // ---- ring prologue, first thing in the unit: the first stages are in flight while the row table and the lane
  // weights are set up.  A stage = ROWS source rows x the band's 128 pixels.  The ring slots are used round-robin
  // across units (wst.slot), so the parity of a slot's mbarrier is one bit that flips when the slot index wraps.
  // A full stage is ONE box; the last, partial stage of a unit is fetched row by row, so nothing is read beyond
  // row re.  y_next / rows_left (the next stage to request) are kept by every lane so that the test stays uniform. ----
  constexpr int ROWS = STREAM_TMA_ROWS, STAGES = STREAM_TMA_STAGES;
  constexpr uint32_t SLOT_BYTES = ROWS * STREAM_STAGE_BYTES;
  const int box_x = 4 * (R.g_base + tx * STRIDE);
  const uint32_t ring_s = smem_u32(&ws.ring[0][0]), full_s = smem_u32(&ws.full[0]);
  const CUtensorMap* const map_box = &maps.box[R.img];
  int y_next = (int)rs, rows_left = nrows;
  auto issue = [&](uint32_t slot_) {   // lane 0 only: request rows [y_next, y_next + min(ROWS, rows_left)) into slot_
    const uint32_t bar = full_s + 8u * slot_, dst = ring_s + slot_ * SLOT_BYTES;
    if (rows_left >= ROWS) {
      mbar_expect_tx(bar, SLOT_BYTES);
      tma_load_2d(dst, map_box, box_x, y_next, bar);
    } else {
      mbar_expect_tx(bar, (uint32_t)rows_left * STREAM_STAGE_BYTES);
      const CUtensorMap* const map_row = map_box + STITCH_MAX_IMAGES;   // TmaMaps::row[img]
#pragma unroll 1
      for (int j = 0; j < rows_left; ++j) tma_load_2d(dst + j * STREAM_STAGE_BYTES, map_row, box_x, y_next + j, bar);
    }
  };
  {
    uint32_t ps = wst.slot;
#pragma unroll
    for (int s0 = 0; s0 < STAGES; ++s0) {
      if (rows_left > 0) {
        if (lane == 0) issue(ps);
        y_next += ROWS;
        rows_left -= ROWS;
        ps = (ps + 1 == STAGES) ? 0u : ps + 1;
      }
    }
  }
  build_pair_table(ws.tab, lane, rs, nrows, nrows_pad / 2, sh, dh, oy0, 32768.f * __frcp_rn((float)sh));

  // ---- horizontal: this lane's 4-pixel source group and the outputs that start in it ----
  constexpr bool RGBA = MODE != STREAM_MODE_YUV, YUV = MODE != STREAM_MODE_RGBA;
  const uint32_t g = (uint32_t)(R.g_base + tx * STRIDE + lane);
  const uint32_t g_last = (sw + 3) / 4 - 1;
  const uint32_t c0 = 4u * g * dw;  // group start, in source units scaled by dw
  const int Xfirst = (int)udiv_ceil(c0, sw);
  int Xnext = __shfl_down_sync(0xffffffffu, Xfirst, 1);
  if (lane == 31) Xnext = Xfirst;  // lane 31 only feeds tails
  const int cnt = min(Xnext - Xfirst, MAXO);
  // hop[k] = 1 when the tail of this lane's last output reaches k+1 groups to the right: it always
  // reaches the next group; it reaches further while the groups in between start no output
  float hop[HOPS];
  int nxt = 1;  // distance to the next lane that starts an output (chroma partner)
  {
    bool open = cnt > 0;
    bool seek = true;
#pragma unroll
    for (int k = 0; k < HOPS; ++k) {
      hop[k] = open ? 1.f : 0.f;
      const int c_k = __shfl_down_sync(0xffffffffu, cnt, k + 1);
      if (seek && c_k > 0) { nxt = k + 1; seek = false; }
      open = open && (c_k == 0);
    }
  }
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
    st_ok[o] = (lane < STRIDE) && (o < cnt) && (X >= R.x0) && (X < R.x1);
    last[o] = (o == cnt - 1) ? 1.f : 0.f;
  }
  // columns past the bitmap edge (row padding; zero-filled by TMA) must not trip the alpha check
  const uint32_t m1 = (4 * g + 1 >= sw) ? 0xff000000u : 0u, m2 = (4 * g + 2 >= sw) ? 0xff000000u : 0u,
                 m3 = (4 * g + 3 >= sw) ? 0xff000000u : 0u;
  const uint32_t m0 = (4 * g >= sw) ? 0xff000000u : 0u;
  // output pointers walk with the rows: device position of this lane's output 0, +-1 pixel per output
  const int xdir = im.flip_x ? -1 : 1;
  const int cx0 = X0 + (im.flip_x ? (int)dw - 1 - Xfirst : Xfirst);      // canvas x of output 0
  const int cy0 = Y0 + (im.flip_y ? (int)(dh - 1 - oy0) : (int)oy0);      // canvas y of the unit's first output row
  const int xstep = 4 * xdir;
  const int64_t ystep = im.flip_y ? -P.out_pitch : P.out_pitch;
  uint8_t* optr = RGBA ? P.out_rgba + (int64_t)cy0 * P.out_pitch + (int64_t)cx0 * 4 : nullptr;
  // I420: Y walks like the canvas; chroma is finished on every second emitted row.  Output o is the
  // FIRST (in source order) of its horizontal pair when its canvas x is even (odd when mirrored); the
  // lane that owns the first element stores the pair's chroma, its partner is output o+1 -- the next
  // one in this lane or the first one of the right neighbour.
  uint8_t* yptr = nullptr;
  uint8_t *cbptr = nullptr, *crptr = nullptr;
  int64_t ystep_y = 0, cstep = 0;
  bool c_own[MAXO], pair_local[MAXO];
  int c_off[MAXO];
  float ccb[MAXO], ccr[MAXO];  // vertical sums of the un-rounded chroma of the current block row
  bool second = false;          // the next emitted row completes a chroma row
  if (YUV) {
    const int cwh = (P.cw + 1) >> 1;
    yptr = P.out_y + (int64_t)cy0 * P.cw + cx0;
    ystep_y = im.flip_y ? -(int64_t)P.cw : (int64_t)P.cw;
    cbptr = P.out_cb + (int64_t)(cy0 >> 1) * cwh + (cx0 >> 1);
    crptr = P.out_cr + (int64_t)(cy0 >> 1) * cwh + (cx0 >> 1);
    cstep = im.flip_y ? -(int64_t)cwh : (int64_t)cwh;
#pragma unroll
    for (int o = 0; o < MAXO; ++o) {
      const int cx = cx0 + o * xdir;
      c_own[o] = st_ok[o] && ((cx & 1) == (im.flip_x ? 1 : 0));
      pair_local[o] = o + 1 < cnt;                 // the pair's second element is this lane's next output
      c_off[o] = (cx >> 1) - (cx0 >> 1);
      ccb[o] = ccr[o] = 0.f;
    }
  }
  // MAXO == 2: two adjacent outputs have different parity, so a lane owns at most ONE chroma sample per block row.
  // Which one, and where its horizontal partner lives, is fixed for the unit: the second-row step selects instead
  // of branching per output (the per-output form cost two divergent regions per chroma row).
  bool own1 = false, own_any = false, part_local = false;
  uint8_t *cb_own = nullptr, *cr_own = nullptr;
  if (YUV && MAXO == 2) {
    own1 = c_own[MAXO - 1];
    own_any = c_own[0] || c_own[MAXO - 1];
    part_local = !own1 && pair_local[0];
    cb_own = cbptr + (own1 ? c_off[MAXO - 1] : c_off[0]);
    cr_own = crptr + (own1 ? c_off[MAXO - 1] : c_off[0]);
  }

  const uint32_t lring = smem_u32(&ws.ring[0][lane * 16]);
  __syncwarp();  // row table complete

  // vertical accumulators: 4 pixels x (r, g, b), integer sums that are float bit patterns (STREAM_ACC_BASE)
  uint32_t acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = STREAM_ACC_BASE;
  uint32_t and_x = 0xffffffffu, and_y = 0xffffffffu, and_z = 0xffffffffu, and_w = 0xffffffffu;
  const uint32_t tab_s = smem_u32(ws.tab);

  // An output row is complete in acc: the horizontal pass runs in registers and the row is stored.
  auto emit_row = [&]() {
    {
      float out[3][MAXO];  // un-rounded outputs, still in the 2^23 + v * 2^15 form
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        float prev = wprev[0] * __uint_as_float(acc[ch]);
#pragma unroll
        for (int c = 1; c < 4; ++c) prev = fmaf(wprev[c], __uint_as_float(acc[c * 3 + ch]), prev);
        float rcv = __shfl_down_sync(0xffffffffu, prev, 1);
#pragma unroll
        for (int k = 1; k < HOPS; ++k) rcv = fmaf(hop[k], __shfl_down_sync(0xffffffffu, prev, k + 1), rcv);
        // outputs in pairs: (out[o], out[o+1]) = sum_c (w[o][c], w[o+1][c]) * acc[c]
#pragma unroll
        for (int o = 0; o + 1 < MAXO; o += 2) {
          mul2_s(out[ch][o], out[ch][o + 1], w[o][0], w[o + 1][0], __uint_as_float(acc[ch]));
#pragma unroll
          for (int c = 1; c < 4; ++c) fma2_s(out[ch][o], out[ch][o + 1], w[o][c], w[o + 1][c], __uint_as_float(acc[c * 3 + ch]));
          fma2_s(out[ch][o], out[ch][o + 1], last[o], last[o + 1], rcv);
        }
        if (MAXO & 1) {
          const int o = MAXO - 1;
          out[ch][o] = w[o][0] * __uint_as_float(acc[ch]);
#pragma unroll
          for (int c = 1; c < 4; ++c) out[ch][o] = fmaf(w[o][c], __uint_as_float(acc[c * 3 + ch]), out[ch][o]);
          out[ch][o] = fmaf(last[o], rcv, out[ch][o]);
        }
      }
      if (RGBA) {
        uint32_t q[3][MAXO];
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
#pragma unroll
          for (int o = 0; o + 1 < MAXO; o += 2) {
            if (ch == 2) quant2_neg(q[ch][o], q[ch][o + 1], out[ch][o], out[ch][o + 1]);
            else quant2(q[ch][o], q[ch][o + 1], out[ch][o], out[ch][o + 1]);
          }
          if (MAXO & 1) q[ch][MAXO - 1] = (ch == 2) ? quant_bits_neg(out[ch][MAXO - 1]) : quant_bits(out[ch][MAXO - 1]);
        }
#pragma unroll
        for (int o = 0; o < MAXO; ++o) {
          const uint32_t v = pack_rgb(q[0][o], q[1][o], q[2][o]);
          if (st_ok[o]) *reinterpret_cast<uint32_t*>(optr + o * xstep) = v;
        }
        optr += ystep;
      }
      if (YUV) {
        // JFIF full-range BT.601.  The luma weights sum to 1, so Y keeps the biased
        // form; the chroma weights sum to 0, so the bias cancels and cb/cr are plain values * 2^-15.
        uint32_t yq[MAXO];
#pragma unroll
        for (int o = 0; o + 1 < MAXO; o += 2) {
          float y0_, y1_;
          mul2_s(y0_, y1_, out[2][o], out[2][o + 1], 0.114f);
          fma2_s(y0_, y1_, out[1][o], out[1][o + 1], 0.587f);
          fma2_s(y0_, y1_, out[0][o], out[0][o + 1], 0.299f);
          quant2(yq[o], yq[o + 1], y0_, y1_);
          fma2_s(ccb[o], ccb[o + 1], out[2][o], out[2][o + 1], 0.5f);
          fma2_s(ccb[o], ccb[o + 1], out[1][o], out[1][o + 1], -0.331264f);
          fma2_s(ccb[o], ccb[o + 1], out[0][o], out[0][o + 1], -0.168736f);
          fma2_s(ccr[o], ccr[o + 1], out[2][o], out[2][o + 1], -0.081312f);
          fma2_s(ccr[o], ccr[o + 1], out[1][o], out[1][o + 1], -0.418688f);
          fma2_s(ccr[o], ccr[o + 1], out[0][o], out[0][o + 1], 0.5f);
        }
        if (MAXO & 1) {
          const int o = MAXO - 1;
          const float R_ = out[0][o], G_ = out[1][o], B_ = out[2][o];
          yq[o] = quant_bits(fmaf(0.299f, R_, fmaf(0.587f, G_, 0.114f * B_)));
          ccb[o] = fmaf(-0.168736f, R_, fmaf(-0.331264f, G_, fmaf(0.5f, B_, ccb[o])));
          ccr[o] = fmaf(0.5f, R_, fmaf(-0.418688f, G_, fmaf(-0.081312f, B_, ccr[o])));
        }
        {
          uint8_t* const yq_ptr = yptr;
#pragma unroll
          for (int o = 0; o < MAXO; ++o) {
            uint8_t* const p = yq_ptr + o * xdir;
            if (st_ok[o]) *p = (uint8_t)yq[o];
          }
        }
        yptr += ystep_y;
        if (__any_sync(0xffffffffu, second)) {   // warp-uniform by construction; the vote tells the compiler
          const float nb_cb = (HOPS == 1) ? __shfl_down_sync(0xffffffffu, ccb[0], 1) : __shfl_sync(0xffffffffu, ccb[0], lane + nxt);
          const float nb_cr = (HOPS == 1) ? __shfl_down_sync(0xffffffffu, ccr[0], 1) : __shfl_sync(0xffffffffu, ccr[0], lane + nxt);
          const float bias = 8388608.f + 128.f * 32768.f;   // 128 in the accumulator form
          if (MAXO == 2) {
            const float mcb = own1 ? ccb[MAXO - 1] : ccb[0], mcr = own1 ? ccr[MAXO - 1] : ccr[0];
            const float qcb = part_local ? ccb[MAXO - 1] : nb_cb, qcr = part_local ? ccr[MAXO - 1] : nb_cr;
            st_u8_if(cb_own, quant_bits(fmaf(0.25f, mcb + qcb, bias)), own_any);
            st_u8_if(cr_own, quant_bits(fmaf(0.25f, mcr + qcr, bias)), own_any);
            cb_own += cstep;
            cr_own += cstep;
          }
#pragma unroll
          for (int o = 0; o < (MAXO == 2 ? 0 : MAXO); ++o) {
            float pcb = nb_cb, pcr = nb_cr;  // partner = output o+1
            if (o + 1 < MAXO) {
              pcb = pair_local[o] ? ccb[o + 1] : nb_cb;
              pcr = pair_local[o] ? ccr[o + 1] : nb_cr;
            }
            const uint32_t cbq = quant_bits(fmaf(0.25f, ccb[o] + pcb, bias));
            const uint32_t crq = quant_bits(fmaf(0.25f, ccr[o] + pcr, bias));
            uint8_t* const pb = cbptr + c_off[o];
            uint8_t* const pr = crptr + c_off[o];
            if (c_own[o]) *pb = (uint8_t)cbq;
            if (c_own[o]) *pr = (uint8_t)crq;
          }
#pragma unroll
          for (int o = 0; o < MAXO; ++o) ccb[o] = ccr[o] = 0.f;
          if (MAXO != 2) {
            cbptr += cstep;
            crptr += cstep;
          }
        }
        second = !second;
      }
    }
  };
  // One row pair (A, B) of the lane's group: see build_pair_table.  e1: one of the rows completes an output row,
  // e2: both do (ratio_y < 2).  Per pixel 2 PRMT interleave the two rows' bytes channel by channel and 3 IDP.2A
  // apply both row weights at once (12 PRMT + 6 FFMA2 per ROW in the float form this replaces).
  auto pair_step = [&](const uint4& pa, const uint4& pb, const uint4& t, const bool e1, const bool e2) {
    uint32_t rg[4], ba[4];
    rg[0] = __byte_perm(pa.x, pb.x, 0x5140); ba[0] = __byte_perm(pa.x, pb.x, 0x7362);
    rg[1] = __byte_perm(pa.y, pb.y, 0x5140); ba[1] = __byte_perm(pa.y, pb.y, 0x7362);
    rg[2] = __byte_perm(pa.z, pb.z, 0x5140); ba[2] = __byte_perm(pa.z, pb.z, 0x7362);
    rg[3] = __byte_perm(pa.w, pb.w, 0x5140); ba[3] = __byte_perm(pa.w, pb.w, 0x7362);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      acc[3 * c] = __dp2a_lo(t.x, rg[c], acc[3 * c]);
      acc[3 * c + 1] = __dp2a_hi(t.x, rg[c], acc[3 * c + 1]);
      acc[3 * c + 2] = __dp2a_lo(t.x, ba[c], acc[3 * c + 2]);
    }
    if (e1) {
      emit_row();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        acc[3 * c] = __dp2a_lo(t.y, rg[c], STREAM_ACC_BASE);
        acc[3 * c + 1] = __dp2a_hi(t.y, rg[c], STREAM_ACC_BASE);
        acc[3 * c + 2] = __dp2a_lo(t.y, ba[c], STREAM_ACC_BASE);
      }
      if (e2) {
        emit_row();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          acc[3 * c] = __dp2a_lo(t.z, rg[c], STREAM_ACC_BASE);
          acc[3 * c + 1] = __dp2a_hi(t.z, rg[c], STREAM_ACC_BASE);
          acc[3 * c + 2] = __dp2a_lo(t.z, ba[c], STREAM_ACC_BASE);
        }
      }
    }
  };

  uint32_t slot = wst.slot, parity = wst.parity;
  const int npairs = nrows_pad / 2;   // even: nrows_pad is a multiple of STREAM_UNR = 4
  for (int p0 = 0; p0 < npairs; p0 += 32) {
    // which of the next 32 row pairs complete output rows: two ballots over the table flags, so the branches in
    // pair_step are on warp-uniform masks and need no reconvergence scaffolding around the shuffles
    const uint32_t fl = (p0 + lane < npairs) ? ws.tab[p0 + lane].w : 0u;
    uint32_t mask1 = __ballot_sync(0xffffffffu, fl != 0u), mask2 = __ballot_sync(0xffffffffu, fl == 3u);
    const int p1 = min(p0 + 32, npairs);
#pragma unroll 1
    for (int ps = p0; ps < p1; ps += ROWS / 2) {   // one stage = 4 row pairs (the unit's last one may hold 2)
      mbar_wait(full_s + 8u * slot, parity);       // the stage's rows have landed
      const uint32_t ring_h = lring + slot * SLOT_BYTES;
      const uint32_t tab_h = tab_s + (uint32_t)ps * 16u;
#pragma unroll
      for (int u = 0; u < ROWS / 2; ++u) {
        if (u == ROWS / 4 && ps + u >= p1) break;
        const uint4 t = lds128(tab_h + u * 16);
        const uint4 pa = lds128(ring_h + (2 * u) * STREAM_STAGE_BYTES);
        const uint4 pb = lds128(ring_h + (2 * u + 1) * STREAM_STAGE_BYTES);
        // alpha check: per column and per row pair, one three-input LOP3 each; the edge masks come in at the end
        and_x &= pa.x & pb.x;
        and_y &= pa.y & pb.y;
        and_z &= pa.z & pb.z;
        and_w &= pa.w & pb.w;
        pair_step(pa, pb, t, (mask1 >> u) & 1u, (mask2 >> u) & 1u);
      }
      mask1 >>= ROWS / 2;
      mask2 >>= ROWS / 2;
      // every lane holds the stage's rows in registers: lane 0 re-arms the slot and requests the unit's next rows
      __syncwarp();
      if (rows_left > 0) {
        if (lane == 0) issue(slot);
        y_next += ROWS;
        rows_left -= ROWS;
      }
      if (++slot == STAGES) {
        slot = 0;
        parity ^= 1u;
      }
    }
  }
  wst.slot = slot;
  wst.parity = parity;
  // Translucent bitmaps (rare: photos are opaque): the unit is redone with the exact premultiplied
  // source-over evaluation.  Outside the hot loop so it costs the streaming path no registers.
  const uint32_t and_acc = (and_x | m0) & (and_y | m1) & (and_z | m2) & (and_w | m3);
  const bool need_fix = (and_acc >> 24) != 0xffu;
  if (__any_sync(0xffffffffu, need_fix)) {
    const int xf0 = __shfl_sync(0xffffffffu, Xfirst, 0), xf1 = __shfl_sync(0xffffffffu, Xfirst, STRIDE);
    const int ox0 = max(R.x0, xf0), ox1 = min(R.x1, xf1);
    const int ax = im.flip_x ? (int)dw - ox1 : ox0, bx = im.flip_x ? (int)dw - ox0 : ox1;
    const int ay = im.flip_y ? (int)(dh - oy1) : (int)oy0, by = im.flip_y ? (int)(dh - oy0) : (int)oy1;
    redo_rect<MODE>(P, X0 + ax, Y0 + ay, X0 + bx, Y0 + by, lane, im.flip_x ? 1 : 0);
  }
  // A lane whose group reaches past the bitmap edge received zero-filled (transparent) pixels: a later unit of this
  // warp reads ring rows it has not fetched as padding (nrows_pad > nrows), and those must look opaque.  Rare
  // (right-edge units of bitmaps whose width is no multiple of 128 px), outside the hot loop.  The branch is taken on
  // a vote and the stores are predicated: a per-lane branch here leaves the compiler unable to prove that the warp
  // is converged when the unit loop comes round, and it then guards every shuffle of the row loop (BRA.DIV).
  if (__any_sync(0xffffffffu, m3 != 0u)) {
#pragma unroll 1
    for (int r = 0; r < STREAM_NST; ++r)
      asm volatile("{\n .reg .pred q;\n setp.ne.u32 q, %2, 0;\n @q st.shared.v4.u32 [%0], {%1,%1,%1,%1};\n}"
                   ::"r"(lring + r * STREAM_STAGE_BYTES), "r"(0xffffffffu), "r"(m3) : "memory");
    fence_proxy_async();
  }
}

}  // namespace stitch
