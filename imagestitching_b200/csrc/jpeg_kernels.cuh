// Baseline JPEG export of the stitched canvas on the GPU.
//
// Replaces the encoder behind safeCanvasToTempFilePath (U:205-242, wx.canvasToTempFilePath with fileType
// 'jpg', quality 1) for the I420 planes the compose kernel writes.  Format: ITU-T T.81 baseline, 4:2:0,
// Annex K tables, JFIF container, restart markers every `ri` MCUs -- the restart interval is what makes
// the entropy coder parallel: every segment starts byte-aligned with its DC predictors at zero.
//
//   J1 jpeg_dct_kernel      one thread per 8x8 block: level shift, integer LL&M forward DCT in registers,
//                           quantise, zig-zag -> int16 coefficients in MCU order (Y00 Y01 Y10 Y11 Cb Cr)
//   J2a jpeg_blockbits_kernel  one THREAD per block: Huffman-codes the block on its own (DC difference against
//                           the previous block of its component, ffs walk over the nonzero AC positions) into
//                           <= 53 words of a lane-interleaved scratch, and records the bit count
//   J2b jpeg_pack_kernel    one WARP per restart segment, a lane per block: warp scan of the bit counts, the
//                           lanes funnel-shift their words into a shared-memory window, which is flushed as
//                           bytes with 0xFF stuffing into the segment's slot of the scratch stream
//   J3 jpeg_offsets_kernel  exclusive scan of the segment lengths (+2 per restart marker) and the header
//      jpeg_gather_kernel   one warp per segment copies it to its final place and appends RSTn / EOI
//
// All arithmetic is int32, so the CPU checker of the test suite reproduces the file bit for bit.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "jpeg_tables.h"

namespace stitch {
namespace jpeg {

struct HeaderBytes {
  uint8_t b[704];
  int32_t n;
};

// worst case of one entropy-coded block: DC 9 + 11 bits, 63 x (16 + 10) bits = 1658 bits = 208 bytes, every
// byte stuffed
#define JPEG_BLOCK_MAX_BYTES 432
#define JPEG_BLK_WORDS 64       // words a coded block may take in the block-bits scratch (worst case 1665 bits = 53)
#define JPEG_WIN_WORDS 512      // per-warp bit window of the packer: 2 KB

#define JDESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))

// one 8-point pass of the LL&M integer DCT (13-bit constants); FIRST = row pass (results scaled up by 4)
template <bool FIRST>
__device__ __forceinline__ void fdct8(int32_t& d0, int32_t& d1, int32_t& d2, int32_t& d3, int32_t& d4, int32_t& d5,
                                      int32_t& d6, int32_t& d7) {
  int32_t t0 = d0 + d7, t7 = d0 - d7, t1 = d1 + d6, t6 = d1 - d6;
  int32_t t2 = d2 + d5, t5 = d2 - d5, t3 = d3 + d4, t4 = d3 - d4;
  const int32_t t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
  constexpr int SH = FIRST ? 11 : 15;
  if (FIRST) {
    d0 = (t10 + t11) << 2;
    d4 = (t10 - t11) << 2;
  } else {
    d0 = JDESCALE(t10 + t11, 2);
    d4 = JDESCALE(t10 - t11, 2);
  }
  int32_t z1 = (t12 + t13) * 4433;
  d2 = JDESCALE(z1 + t13 * 6270, SH);
  d6 = JDESCALE(z1 - t12 * 15137, SH);
  z1 = t4 + t7;
  int32_t z2 = t5 + t6, z3 = t4 + t6, z4 = t5 + t7;
  const int32_t z5 = (z3 + z4) * 9633;
  t4 *= 2446;
  t5 *= 16819;
  t6 *= 25172;
  t7 *= 12299;
  z1 *= -7373;
  z2 *= -20995;
  z3 = z3 * -16069 + z5;
  z4 = z4 * -3196 + z5;
  d7 = JDESCALE(t4 + z1 + z3, SH);
  d5 = JDESCALE(t5 + z2 + z4, SH);
  d3 = JDESCALE(t6 + z2 + z3, SH);
  d1 = JDESCALE(t7 + z1 + z4, SH);
}

// round-half-up division of the magnitude by qv = q * 8; the quotient comes from a multiply-high with
// ceil(2^32 / qv), exact for every numerator an 8-bit block can produce (< 2^16)
__device__ __forceinline__ int32_t jquant(int32_t v, int32_t qv, uint32_t magic) {
  const bool neg = v < 0;
  const uint32_t n = (uint32_t)(neg ? -v : v) + (uint32_t)(qv >> 1);
  const int32_t r = (int32_t)__umulhi(n, magic);
  return neg ? -r : r;
}

// J1.  Threads [0, 4M) are the luma blocks in raster order of the (2*mx) x (2*my) luma block grid (so a
// warp reads 256 contiguous bytes of each pixel row), [4M, 5M) the Cb blocks, [5M, 6M) the Cr blocks.
__global__ void __launch_bounds__(128) jpeg_dct_kernel(const uint8_t* __restrict__ py, const uint8_t* __restrict__ pcb,
                                                       const uint8_t* __restrict__ pcr, int w, int h, int mx, int my,
                                                       const __grid_constant__ Tables tab, int16_t* __restrict__ coef) {
  const int M = mx * my;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 6 * M) return;
  const uint8_t* plane;
  int pw, ph, bx, by, slot, comp;
  if (t < 4 * M) {
    const int gx = t % (2 * mx), gy = t / (2 * mx);
    plane = py; pw = w; ph = h; bx = gx; by = gy; comp = 0;
    slot = ((gy >> 1) * mx + (gx >> 1)) * 6 + ((gy & 1) << 1) + (gx & 1);
  } else {
    const int c = (t - 4 * M) / M, m = (t - 4 * M) % M;
    plane = c ? pcr : pcb; pw = (w + 1) >> 1; ph = (h + 1) >> 1; bx = m % mx; by = m / mx; comp = 1;
    slot = m * 6 + 4 + c;
  }
  int32_t d[64];
  const int x0 = bx * 8, y0 = by * 8;
  const bool fast = (x0 + 8 <= pw) && (((pw | (int)(uintptr_t)plane) & 3) == 0);
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const int yy = min(y0 + r, ph - 1);
    const uint8_t* row = plane + (size_t)yy * pw;
    if (fast) {
      const uint32_t a = *reinterpret_cast<const uint32_t*>(row + x0), b = *reinterpret_cast<const uint32_t*>(row + x0 + 4);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        d[r * 8 + c] = (int32_t)((a >> (8 * c)) & 0xff) - 128;
        d[r * 8 + 4 + c] = (int32_t)((b >> (8 * c)) & 0xff) - 128;
      }
    } else {
#pragma unroll
      for (int c = 0; c < 8; ++c) d[r * 8 + c] = (int32_t)row[min(x0 + c, pw - 1)] - 128;  // edge replication
    }
  }
#pragma unroll
  for (int r = 0; r < 8; ++r)
    fdct8<true>(d[r * 8], d[r * 8 + 1], d[r * 8 + 2], d[r * 8 + 3], d[r * 8 + 4], d[r * 8 + 5], d[r * 8 + 6], d[r * 8 + 7]);
#pragma unroll
  for (int c = 0; c < 8; ++c)
    fdct8<false>(d[c], d[8 + c], d[16 + c], d[24 + c], d[32 + c], d[40 + c], d[48 + c], d[56 + c]);
  constexpr uint8_t ZZ[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                              41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                              30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
  uint32_t packed[32];
#pragma unroll
  for (int k = 0; k < 64; k += 2) {
    const int32_t a = jquant(d[ZZ[k]], tab.qdiv[comp][ZZ[k]], tab.qmagic[comp][ZZ[k]]);
    const int32_t b = jquant(d[ZZ[k + 1]], tab.qdiv[comp][ZZ[k + 1]], tab.qmagic[comp][ZZ[k + 1]]);
    packed[k >> 1] = ((uint32_t)a & 0xffffu) | ((uint32_t)b << 16);
  }
  uint4* o = reinterpret_cast<uint4*>(coef + (size_t)slot * 64);
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] = make_uint4(packed[4 * i], packed[4 * i + 1], packed[4 * i + 2], packed[4 * i + 3]);
}

// ---- J2: entropy coding in two steps: code every block on its own, then pack the segments ----

__device__ __forceinline__ int hibit(uint32_t x) { return 31 - __clz(x); }  // -1 for x == 0

// value bits of a coefficient / DC difference of magnitude category s (T.81 F.1.2.1)
__device__ __forceinline__ uint32_t vbits(int v, int s) { return (uint32_t)(v < 0 ? v - 1 : v) & ((1u << s) - 1u); }

// Word j of block `slot` in the block-bits scratch.  Blocks are grouped by 32 (one warp of J2a); inside a group
// the words are interleaved -- word j of the 32 blocks is one 128-byte line -- so that both J2a (lane = block,
// all lanes near the same j) and J2b (lane = block of a segment) touch whole lines.
__device__ __forceinline__ size_t blockbits_index(int slot, int j) {
  return ((size_t)(slot >> 5) * JPEG_BLK_WORDS + (size_t)j) * 32u + (size_t)(slot & 31);
}

// J2a.  One THREAD per 8x8 block, blocks in MCU order (the order of `coef`).  The thread parks its 32 coefficient
// words in its own shared-memory column (bank = lane: conflict-free for any index), walks the nonzero AC
// positions with ffs over a 64-bit map, and shifts DC difference, (ZRL,) run/size codes + value bits and EOB
// into a 64-bit register that is drained one 32-bit word at a time, MSB first.  blen[slot] = bits coded; the
// unused low bits of the last word are zero.  The DC predictor is the previous block of the same component
// (zero at the start of a restart segment), read from `coef`.
__global__ void __launch_bounds__(128) jpeg_blockbits_kernel(const int16_t* __restrict__ coef, int n_blocks, int ri,
                                                             const __grid_constant__ Tables tab, uint32_t* __restrict__ bits,
                                                             uint32_t* __restrict__ blen) {
  __shared__ uint32_t s_huff[4][256];
  __shared__ uint32_t s_c[32][128];
  const int tid = threadIdx.x;
  for (int i = tid; i < 4 * 256; i += 128) s_huff[i >> 8][i & 255] = tab.huff[i >> 8][i & 255];
  const int slot = blockIdx.x * 128 + tid;
  const bool live = slot < n_blocks;
  uint32_t nzlo = 0, nzhi = 0;   // nonzero maps of zig-zag positions 0..31 / 32..63
  if (live) {
    const uint4* src = reinterpret_cast<const uint4*>(coef + (size_t)slot * 64);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint4 v = src[i];
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int wi = 4 * i + k;   // holds positions 2*wi, 2*wi + 1
        s_c[wi][tid] = w[k];
        const uint32_t m = ((w[k] & 0xffffu) ? 1u : 0u) | ((w[k] >> 16) ? 2u : 0u);
        if (wi < 16) nzlo |= m << (2 * wi);
        else nzhi |= m << (2 * wi - 32);
      }
    }
  }
  __syncthreads();   // the code tables; s_c columns are private to their thread
  if (!live) return;
  const int b = slot % 6, m_idx = slot / 6;
  const int comp = b < 4 ? 0 : b - 3;
  const uint32_t* dc = s_huff[comp ? 1 : 0];
  const uint32_t* ac = s_huff[comp ? 3 : 2];
  uint32_t* out = bits + blockbits_index(slot, 0);
  uint64_t acc = 0;
  int nb = 0, nw = 0;
  auto put = [&](uint32_t val, int len) {   // len <= 27, nb < 32 on entry
    acc = (acc << len) | val;
    nb += len;
    if (nb >= 32) {
      nb -= 32;
      out[(size_t)nw * 32u] = (uint32_t)(acc >> nb);
      ++nw;
    }
  };
  {
    int pred = 0;
    const bool seg_first = (m_idx % ri) == 0;
    if ((b >= 1 && b <= 3) || !seg_first) {
      const int prev_slot = b == 0 ? slot - 3 : (b < 4 ? slot - 1 : slot - 6);
      pred = coef[(size_t)prev_slot * 64];
    }
    const int diff = (int)(int16_t)(s_c[0][tid] & 0xffffu) - pred;
    const int s = 32 - __clz((uint32_t)(diff < 0 ? -diff : diff));
    const uint32_t e = dc[s];
    put(((e & 0xffffu) << s) | vbits(diff, s), (int)(e >> 16) + s);
  }
  const uint32_t zrl = ac[0xf0];
  int prev = 0;
  nzlo &= ~1u;   // position 0 is the DC coefficient
  // (fetching the next nonzero position's word while the current one is coded was measured: 67 -> 70 us)
#pragma unroll 1
  for (int half = 0; half < 2; ++half) {
    uint32_t mask = half ? nzhi : nzlo;
    while (mask) {
      const int k = __ffs(mask) - 1 + 32 * half;
      mask &= mask - 1;
      const uint32_t w = s_c[k >> 1][tid];
      const int v = (int)(int16_t)((k & 1) ? (w >> 16) : (w & 0xffffu));
      int run = k - prev - 1;
      prev = k;
      while (run >= 16) {
        put(zrl & 0xffffu, (int)(zrl >> 16));
        run -= 16;
      }
      const int s = 32 - __clz((uint32_t)(v < 0 ? -v : v));
      const uint32_t e = ac[(run << 4) | s];
      put(((e & 0xffffu) << s) | vbits(v, s), (int)(e >> 16) + s);
    }
  }
  if (prev != 63) {
    const uint32_t e = ac[0];
    put(e & 0xffffu, (int)(e >> 16));
  }
  if (nb) out[(size_t)nw * 32u] = (uint32_t)(acc << (32 - nb));
  blen[slot] = (uint32_t)(nw * 32 + nb);
}

// J2b.  One WARP per restart segment: concatenates the segment's coded blocks and writes its bytes with 0xFF
// stuffing into the segment's slot of the scratch stream; seg_len[s] = byte count (stuffing included).
// Per round a lane takes one block: a warp scan of the bit lengths gives every block its bit offset, the
// round takes as many blocks as fit the shared-memory window (at least one: the worst block is 1665 bits),
// every lane funnel-shifts its block's words to the offset and ORs them into the window -- neighbouring
// blocks share a word, hence the atomics -- and the complete bytes are flushed.
__global__ void __launch_bounds__(128) jpeg_pack_kernel(const uint32_t* __restrict__ bits, const uint32_t* __restrict__ blen,
                                                        int n_blocks, int ri, int n_seg, uint8_t* __restrict__ scratch,
                                                        int64_t seg_cap, uint32_t* __restrict__ seg_len) {
  __shared__ uint32_t s_win[4][JPEG_WIN_WORDS + 4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t* win = s_win[warp];
  for (int i = lane; i < JPEG_WIN_WORDS + 4; i += 32) win[i] = 0;
  __syncwarp();
  const int seg = blockIdx.x * 4 + warp;
  if (seg >= n_seg) return;
  uint8_t* out = scratch + (int64_t)seg * seg_cap;
  const uint32_t lt = (1u << lane) - 1u;
  int outpos = 0;   // bytes written to the segment so far
  int carry = 0;    // bits of the last, incomplete byte waiting in win[0] (< 8)
  int s0 = (int)min((int64_t)seg * ri * 6, (int64_t)n_blocks);
  const int s1 = (int)min((int64_t)s0 + (int64_t)ri * 6, (int64_t)n_blocks);
  while (s0 < s1) {
    const int slot = s0 + lane;
    int len = slot < s1 ? (int)blen[slot] : 0;
    int incl = len;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    const bool fits = slot < s1 && carry + incl <= JPEG_WIN_WORDS * 32;
    const int n_take = __popc(__ballot_sync(0xffffffffu, fits));   // a prefix of the lanes, >= 1
    const int total = carry + __shfl_sync(0xffffffffu, incl, n_take - 1);
    if (lane >= n_take) len = 0;
    const int pos = carry + incl - len;
    const int nw = (len + 31) >> 5;
    const int maxw = __reduce_max_sync(0xffffffffu, nw);
    const uint32_t* src = bits + blockbits_index(min(slot, n_blocks - 1), 0);
    const int wi = pos >> 5, bo = pos & 31;
    uint32_t prevw = 0;
    for (int j = 0; j <= maxw; ++j) {
      const uint32_t w = j < nw ? src[(size_t)j * 32u] : 0u;
      const uint32_t piece = __funnelshift_r(w, prevw, bo);   // low word of (prevw:w) >> bo
      if (j <= nw && piece) atomicOr(&win[wi + j], piece);
      prevw = w;
    }
    __syncwarp();
    s0 += n_take;
    // flush the complete bytes with 0xFF stuffing
    const int nbytes = total >> 3;
    int extra = 0;
    for (int j0 = 0; j0 < nbytes; j0 += 32) {
      const int j = j0 + lane;
      const bool valid = j < nbytes;
      const uint32_t by = valid ? (win[j >> 2] >> (24 - 8 * (j & 3))) & 0xffu : 0u;
      const uint32_t ffm = __ballot_sync(0xffffffffu, valid && by == 0xffu);
      if (valid) {
        const int p = outpos + j + extra + __popc(ffm & lt);
        out[p] = (uint8_t)by;
        if (by == 0xffu) out[p + 1] = 0;
      }
      extra += __popc(ffm);
    }
    outpos += nbytes + extra;
    // keep the partial byte, clear the rest of the window
    carry = total & 7;
    const uint32_t part = carry ? ((win[nbytes >> 2] >> (24 - 8 * (nbytes & 3))) & 0xffu) << 24 : 0u;
    __syncwarp();
    for (int i = lane; i <= (total >> 5) + 1; i += 32) win[i] = 0;
    __syncwarp();
    if (lane == 0) win[0] = part;
    __syncwarp();
  }
  // pad the last byte with 1-bits (stuffed like any other byte)
  if (lane == 0) {
    if (carry) {
      const uint32_t by = ((win[0] >> 24) & 0xffu) | ((1u << (8 - carry)) - 1u);
      out[outpos++] = (uint8_t)by;
      if (by == 0xffu) out[outpos++] = 0;
    }
    seg_len[seg] = (uint32_t)outpos;
  }
}

// ---- J3: place the segments ----

// One block: exclusive scan of (seg_len + 2 bytes of RSTn, none after the last segment) starting behind the
// header; writes the header, the segment offsets and the total file size (EOI included).  info[0] = total
// bytes, info[1] = 1 when the file does not fit `cap`.
__global__ void __launch_bounds__(1024) jpeg_offsets_kernel(const uint32_t* __restrict__ seg_len, int n_seg,
                                                            const __grid_constant__ HeaderBytes hdr, uint8_t* __restrict__ out,
                                                            int64_t cap, uint64_t* __restrict__ seg_off, uint64_t* __restrict__ info) {
  __shared__ uint64_t s_warp[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // thread t owns the segments [t * per, t * per + per): its loads are independent, then ONE block scan
  const int per = (n_seg + 1023) / 1024;
  const int i0 = min(tid * per, n_seg), i1 = min(i0 + per, n_seg);
  uint64_t mine = 0;
  for (int i = i0; i < i1; ++i) mine += (uint64_t)seg_len[i] + (i + 1 < n_seg ? 2u : 0u);
  uint64_t incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint64_t up = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += up;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint64_t v = s_warp[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint64_t up = __shfl_up_sync(0xffffffffu, v, d);
      if (lane >= d) v += up;
    }
    s_warp[lane] = v;
  }
  __syncthreads();
  uint64_t run = (uint64_t)hdr.n + (warp ? s_warp[warp - 1] : 0) + incl - mine;
  for (int i = i0; i < i1; ++i) {
    seg_off[i] = run;
    run += (uint64_t)seg_len[i] + (i + 1 < n_seg ? 2u : 0u);
  }
  const uint64_t total = (uint64_t)hdr.n + s_warp[31] + 2;   // EOI
  if (tid == 0) {
    info[0] = total;
    info[1] = total > (uint64_t)cap ? 1u : 0u;
  }
  if (total <= (uint64_t)cap)
    for (int i = tid; i < hdr.n; i += 1024) out[i] = hdr.b[i];
}

// One warp per segment: copy it behind its predecessors, then RSTn (n = segment index mod 8) or, after the
// last one, EOI.
__global__ void __launch_bounds__(256) jpeg_gather_kernel(const uint8_t* __restrict__ scratch, int64_t seg_cap,
                                                          const uint32_t* __restrict__ seg_len,
                                                          const uint64_t* __restrict__ seg_off, int n_seg,
                                                          const uint64_t* __restrict__ info, uint8_t* __restrict__ out) {
  const int seg = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (seg >= n_seg || info[1]) return;
  const uint8_t* src = scratch + (int64_t)seg * seg_cap;
  uint8_t* dst = out + seg_off[seg];
  const uint32_t n = seg_len[seg];
  for (uint32_t j = lane; j < n; j += 32) dst[j] = src[j];
  if (lane == 0) {
    dst[n] = 0xff;
    dst[n + 1] = (uint8_t)(seg + 1 < n_seg ? 0xd0 + (seg & 7) : 0xd9);
  }
}

}  // namespace jpeg
}  // namespace stitch
