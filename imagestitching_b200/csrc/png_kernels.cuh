// PNG export on the GPU (STITCH_OUT_PNG): what the page asks canvasToTempFilePath for (fileType 'png',
// U:205-242, called at P:1577-1579).  ISO/IEC 15948 colour type 6, 8 bit; IDAT = zlib (RFC 1950) / deflate
// (RFC 1951).  The encoder's free choices are made for parallelism:
//   * Paeth filter on every row (the predictor only looks at the canvas, never at compressed state);
//   * the rows are cut into bands of PNG_BAND_ROWS rows, one CTA each; a band is ONE dynamic-Huffman block of
//     literals and run-length matches (distance 1; see round_tokens) + -- except for the last band -- an empty
//     stored block that byte-aligns the stream, and is its own IDAT chunk, so bands are coded, CRC'd and placed
//     independently;
//   * the Adler-32 of the whole stream is combined from per-band sums.
// P1 png_hist_kernel    filter + tokens (kept for P2), token histogram + Adler partial sums per band, then the length-
//                       limited Huffman code (merge the two smallest (frequency, index); too deep -> halve the
//                       frequencies and rebuild) and its canonical, bit-reversed codes
// P2 png_encode_kernel  block header + codes: per round every thread codes the tokens of 2 pixels into registers, a
//                       CTA scan of the bit counts places them in a shared-memory bit window flushed as words
// P3 png_offsets_kernel scan of the band sizes, Adler combine, signature + IHDR + IEND
// P4 png_gather_kernel  chunk length, 'IDAT', data, CRC-32 (32 lanes over 1/32 of the chunk each, combined
//                       with x^(8n) mod P multiplications)
// Every choice is restated by the sequential writer of the test tree; the files are compared byte for byte and
// the test tree's writer is pinned by decoding with libpng.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace stitch {
namespace png {

#define PNG_BAND_ROWS 8
#define PNG_NSYM 286           // 256 literals + end of block + 29 length codes
#define PNG_ROUND_PX 512        // pixels per tokenisation round (2 per thread); runs never cross a round
#define PNG_THREADS 256
#define PNG_TAB_STRIDE 320     // per-band stride of the length / code tables
#define PNG_CRC_POLY 0xEDB88320u

struct HeaderBytes {
  uint8_t b[48];               // signature + IHDR chunk (33 bytes)
  int32_t n;
};

struct Params {
  const uint8_t* rgba;
  int64_t pitch;
  int32_t w, h, nb;
  uint32_t w_magic;            // ceil(2^32 / w) for 2 <= w <= 16384, else 0 (see div_w)
  uint8_t* lens;               // [nb][PNG_TAB_STRIDE]
  uint16_t* codes;             // [nb][PNG_TAB_STRIDE]
  uint32_t* adler;             // [nb][2]: Adler pair of the band's bytes started from (1, 0)
  uint32_t* nmatch;            // [nb] match tokens in the band (0: the distance alphabet is empty)
  uint32_t* fpx;               // [h][w] filtered pixels (P1 -> P2)
  uint64_t* desc;              // [nb][rounds][PNG_THREADS] token descriptors (P1 -> P2), see pack_tokens
  int32_t rounds;              // rounds per band = ceil(PNG_BAND_ROWS * w / PNG_ROUND_PX)
  uint8_t* scratch;            // [nb][band_cap] coded bands
  int64_t band_cap;
  uint32_t* blen;              // [nb] bytes of zlib data per band (the last one includes the 4 Adler bytes)
  uint64_t* off;               // [nb] file offset of the band's chunk
  uint64_t* info;              // {file size, overflow flag}
  uint8_t* file;
  int64_t cap;
  uint32_t x2n[32];            // x^(2^k) mod P, reflected (CRC combine)
  uint32_t xinv2n[32];         // x^(-2^k) mod P
};

__device__ __forceinline__ int paeth(int a, int b, int c) {
  const int p = a + b - c;
  const int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// The four filtered bytes of pixel (j, y), packed little-endian.
__device__ __forceinline__ uint32_t filter_px(const Params& P, int j, int y) {
  const uint8_t* row = P.rgba + (int64_t)y * P.pitch;
  const uint32_t x = *reinterpret_cast<const uint32_t*>(row + (int64_t)j * 4);
  const uint32_t a = j > 0 ? *reinterpret_cast<const uint32_t*>(row + (int64_t)(j - 1) * 4) : 0u;
  uint32_t b = 0, c = 0;
  if (y > 0) {
    const uint8_t* prev = row - P.pitch;
    b = *reinterpret_cast<const uint32_t*>(prev + (int64_t)j * 4);
    c = j > 0 ? *reinterpret_cast<const uint32_t*>(prev + (int64_t)(j - 1) * 4) : 0u;
  }
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int s = 8 * k;
    const int pr = paeth((int)((a >> s) & 255u), (int)((b >> s) & 255u), (int)((c >> s) & 255u));
    out |= ((((x >> s) & 255u) - (uint32_t)pr) & 255u) << s;
  }
  return out;
}


// ---- run-length tokens of one round ----
// LZ77 restricted to runs: inside a round (PNG_ROUND_PX pixels in raster order, the filter-type byte in front
// of a row's first pixel) a maximal run of equal bytes is its first byte + matches of distance 1: chunks of up
// to 258 bytes, a chunk shorter than 3 stays literal.  A token depends on its round only; the sequential writer
// of the test tree derives the same tokens.
//
// A thread owns 2 pixels = ten SLOTS, statically indexed: slot 5e is the filter-type byte in front of pixel e
// (present when the pixel starts a row), slots 5e+1..5e+4 its four bytes; `valid` marks the slots that exist.
// Almost every warp is PLAIN -- all 32 lanes have both pixels and no row start -- and runs the instantiation in
// which `valid` is the constant 0x3de, so every validity test and position count folds away.
#define PNG_PLAIN_VALID 0x3deu
struct RoundSmem {   // per-warp partials, double-buffered by round parity (one barrier per round)
  int warp_a[2][PNG_THREADS / 32];      // last run start inside the warp (its very first byte not counted), or -1
  int warp_b[2][PNG_THREADS / 32];      // first run start inside the warp (same exclusion), or INT_MAX
  int warp_g0[2][PNG_THREADS / 32];     // round index of the warp's first byte, -1: the warp has no bytes
  uint8_t warp_first[2][PNG_THREADS / 32], warp_last[2][PNG_THREADS / 32];   // its first and last byte values
};
struct ThreadTokens {
  uint32_t f0, f1;    // the two filtered pixels (0 when absent)
  uint32_t valid;     // slots that exist
  int off;            // round index of the thread's first byte
  uint32_t startmask; // slots that start a run
  int a, b;           // warp scans (inclusive) of the thread's last (max, from the left) / first (min, from the right) start
  uint32_t cls;       // 2 bits per slot: 0 nothing to emit, 1 literal, 2 match of distance 1
  uint32_t lens;      // 8 bits per match, in slot order: length - 3
};

// q / w for a pixel index q inside a band (q < (PNG_BAND_ROWS + 1) * w): one multiply-high.  Exact for 2 <= w <= 16384:
// with m = ceil(2^32 / w), m * w - 2^32 < w and q * w < 9 * 16384^2 < 2^32.  Other widths: w_magic = 0, plain division.
__device__ __forceinline__ int div_w(const Params& P, int q) { return P.w_magic ? (int)__umulhi((uint32_t)q, P.w_magic) : q / P.w; }
__device__ __forceinline__ int multiples_before(const Params& P, int q) { return div_w(P, q + P.w - 1); }   // multiples of w in [0, q)

__device__ __forceinline__ uint32_t slot_byte(const ThreadTokens& tt, int i) {
  if (i == 0 || i == 5) return 4u;                  // the row's filter-type byte (Paeth)
  return ((i < 5 ? tt.f0 : tt.f1) >> (8 * ((i % 5) - 1))) & 255u;
}
template <bool PLAIN>
__device__ __forceinline__ int slot_pos(const ThreadTokens& tt, int i) {   // round index of slot i's byte
  const uint32_t valid = PLAIN ? PNG_PLAIN_VALID : tt.valid;
  return tt.off + __popc(valid & ((1u << i) - 1u));
}

// Pixels, slots and position of the thread in the round; returns whether the whole warp is PLAIN.
__device__ __forceinline__ bool round_load(const Params& P, int y0, int npx, int base, int mb_base, ThreadTokens& tt) {
  const int t = threadIdx.x;
  const int q_end = min(base + PNG_ROUND_PX, npx);
  const int q0 = base + 2 * t;
  tt.valid = 0;
  tt.f0 = tt.f1 = 0;
  tt.off = 4 * (min(q0, q_end) - base) + multiples_before(P, min(q0, q_end)) - mb_base;
  if (q0 < q_end) {
    const int r = div_w(P, q0), j = q0 - r * P.w;
    tt.valid = j == 0 ? 0x1fu : 0x1eu;
    tt.f0 = filter_px(P, j, y0 + r);
    if (q0 + 1 < q_end) {
      const bool wrap = j + 1 == P.w;
      tt.valid |= wrap ? 0x3e0u : 0x3c0u;
      tt.f1 = filter_px(P, wrap ? 0 : j + 1, y0 + r + (wrap ? 1 : 0));
    }
  }
  return __all_sync(0xffffffffu, tt.valid == PNG_PLAIN_VALID);
}

// Run starts of the thread (the first byte of lane 0 is left out: its predecessor lives in another warp), warp
// scans, and the warp's partials for the other warps.
template <bool PLAIN>
__device__ __forceinline__ void round_starts(ThreadTokens& tt, RoundSmem& rs, int buf) {
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  const uint32_t valid = PLAIN ? PNG_PLAIN_VALID : tt.valid;
  const uint32_t mylast = ((valid >> 6 & 1u) ? tt.f1 : tt.f0) >> 24;
  uint32_t prev = __shfl_up_sync(0xffffffffu, mylast, 1);
  bool have_prev = lane > 0;
  uint32_t sm = 0;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    if (valid >> i & 1u) {
      const uint32_t v = slot_byte(tt, i);
      if (have_prev && v != prev) sm |= 1u << i;
      prev = v;
      have_prev = true;
    }
  }
  tt.startmask = sm;
  int a = -1, b = 0x7fffffff;
  if (sm) {
    a = slot_pos<PLAIN>(tt, 31 - __clz(sm));
    b = slot_pos<PLAIN>(tt, __ffs(sm) - 1);
  }
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int ya = __shfl_up_sync(0xffffffffu, a, d), yb = __shfl_down_sync(0xffffffffu, b, d);
    if (lane >= d) a = max(a, ya);
    if (lane + d < 32) b = min(b, yb);
  }
  tt.a = a;
  tt.b = b;
  if (lane == 31) { rs.warp_a[buf][wp] = a; rs.warp_last[buf][wp] = (uint8_t)mylast; }
  if (lane == 0) {
    rs.warp_b[buf][wp] = b;
    rs.warp_g0[buf][wp] = valid ? tt.off : -1;
    rs.warp_first[buf][wp] = (uint8_t)((valid & 1u) ? 4u : (tt.f0 & 255u));
  }
}

// After the barrier: the run around every byte of the thread -> cls / lens.
template <bool PLAIN>
__device__ __forceinline__ void round_finish(ThreadTokens& tt, const RoundSmem& rs, int buf, int n_round) {
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  const uint32_t valid = PLAIN ? PNG_PLAIN_VALID : tt.valid;
  int before = __shfl_up_sync(0xffffffffu, tt.a, 1), after = __shfl_down_sync(0xffffffffu, tt.b, 1);
  if (lane == 0) before = -1;
  if (lane == 31) after = 0x7fffffff;
  uint32_t sm = tt.startmask;
  {
    // the other warps' partials: lane k (k < 8) looks at warp k, two warp reductions give every lane the last run
    // start before its warp and the first one after it (every thread walking all 8 entries was 12 % of P1)
    static_assert(PNG_THREADS / 32 <= 32, "one lane per warp of the CTA");
    const int k = lane < PNG_THREADS / 32 ? lane : 0;
    const int g0 = rs.warp_g0[buf][k];
    // does warp k's first byte start a run?  (the round's first byte always does)
    const bool real = g0 >= 0 && (k == 0 || rs.warp_first[buf][k] != rs.warp_last[buf][k > 0 ? k - 1 : 0]);
    const int ak = max(rs.warp_a[buf][k], real ? g0 : -1);
    const int bk = real ? g0 : rs.warp_b[buf][k];
    const bool mine = lane < PNG_THREADS / 32;
    before = max(before, __reduce_max_sync(0xffffffffu, (mine && k < wp) ? ak : -1));
    after = min(after, __reduce_min_sync(0xffffffffu, (mine && k > wp) ? bk : 0x7fffffff));
    const int real_wp = __shfl_sync(0xffffffffu, real ? 1 : 0, wp), g0_wp = __shfl_sync(0xffffffffu, g0, wp);
    if (real_wp) {
      if (lane == 0) sm |= (valid & 1u) ? 1u : 2u;
      else before = max(before, g0_wp);
    }
  }
  after = min(after, n_round);
  int e_of[10];
  {
    int cur = after;
#pragma unroll
    for (int i = 9; i >= 0; --i) {
      e_of[i] = cur;
      if (sm >> i & 1u) cur = slot_pos<PLAIN>(tt, i);
    }
  }
  int cur_s = before;
  uint32_t cls = 0, lens = 0;
  int mi = 0;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    if (valid >> i & 1u) {
      const int g = slot_pos<PLAIN>(tt, i);
      if (sm >> i & 1u) cur_s = g;
      const int k = g - cur_s, L = e_of[i] - cur_s;
      uint32_t c = 1u;
      if (k >= 1) {
        const int ch = (k - 1) / 258, o = (k - 1) - 258 * ch;
        const int clen = min(258, (L - 1) - 258 * ch);
        if (clen >= 3) {
          c = o == 0 ? 2u : 0u;
          if (o == 0) { lens |= (uint32_t)(clen - 3) << (8 * mi); ++mi; }
        }
      }
      cls |= c << (2 * i);
    }
  }
  tt.cls = cls;
  tt.lens = lens;
}

// length 3..258 -> literal/length symbol 257..285 and its extra bits (RFC 1951 3.2.5)
__device__ __forceinline__ int len_symbol(int len, int* extra, int* nextra) {
  int c, e = 0, ne = 0;
  if (len < 11) c = 254 + len;
  else if (len == 258) c = 285;
  else {
    const int n = len - 3;
    ne = (31 - __clz(n)) - 2;
    c = 257 + 4 * ne + (n >> ne);
    e = n & ((1 << ne) - 1);
  }
  *extra = e;
  *nextra = ne;
  return c;
}

// P1 leaves a thread's tokens of a round to P2 as d = cls | lens << 20 (44 bits) next to the filtered pixels, so
// that P2 neither filters nor tokenises again.

// P1's share of a round after the barrier: tokens, their histogram, and the Adler sums of the thread's bytes
// (s2 weights a byte by the number of stream bytes from it to the end of the band).
template <bool PLAIN>
__device__ __forceinline__ void hist_round(ThreadTokens& tt, const RoundSmem& rs, int buf, int n_round, int64_t left,
                                           uint32_t* hist, uint32_t& zeros, uint32_t& matches,
                                           unsigned long long& s1, unsigned long long& s2) {
  round_finish<PLAIN>(tt, rs, buf, n_round);
  const uint32_t valid = PLAIN ? PNG_PLAIN_VALID : tt.valid;
  {
    // a byte sits at round index (off - c0) + r: r = slot for slots 0..5, slot - c5 for the second pixel
    const int c0 = (valid & 1u) ? 0 : 1, c5 = (valid >> 5 & 1u) ? 0 : 1;
    const uint32_t sum0 = __dp4a(tt.f0, 0x01010101u, 0u), sum1 = __dp4a(tt.f1, 0x01010101u, 0u);
    const uint32_t S = sum0 + sum1 + 4u * (uint32_t)((1 - c0) + (1 - c5));
    const uint32_t W = __dp4a(tt.f0, 0x04030201u, 0u) + __dp4a(tt.f1, 0x09080706u, 0u) + (c5 ? 0u : 20u) - (uint32_t)c5 * sum1;
    s1 += S;
    s2 += (unsigned long long)(left - (tt.off - c0)) * S - W;
  }
  int mi = 0;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    if (!(valid >> i & 1u)) continue;
    const uint32_t c = tt.cls >> (2 * i) & 3u;
    if (c == 1u) {
      const uint32_t v = slot_byte(tt, i);
      if (v == 0) ++zeros;
      else atomicAdd(&hist[v], 1u);
    } else if (c == 2u) {
      int ex, nex;
      atomicAdd(&hist[len_symbol(3 + (int)(tt.lens >> (8 * mi) & 255u), &ex, &nex)], 1u);
      ++mi;
      ++matches;
    }
  }
}

// ---- P1 ----
__global__ void __launch_bounds__(PNG_THREADS) png_hist_kernel(const __grid_constant__ Params P) {
  __shared__ uint32_t hist[PNG_THREADS / 32][PNG_NSYM + 3];
  __shared__ uint32_t freq[2 * PNG_NSYM];      // node weights (leaves, then internal nodes)
  __shared__ uint16_t parent[2 * PNG_NSYM];
  __shared__ uint8_t depth[2 * PNG_NSYM];
  __shared__ unsigned long long red[2][PNG_THREADS / 32];
  const int band = blockIdx.x, t = threadIdx.x, lane = t & 31, wp = t >> 5;
  const int y0 = band * PNG_BAND_ROWS, rows = min(PNG_BAND_ROWS, P.h - y0);
  const int64_t rowlen = (int64_t)P.w * 4 + 1, n = rowlen * rows;
  for (int i = t; i < (PNG_THREADS / 32) * (PNG_NSYM + 3); i += PNG_THREADS) (&hist[0][0])[i] = 0;
  __syncthreads();
  __shared__ RoundSmem rs;
  __shared__ uint32_t s_nmatch;
  if (t == 0) s_nmatch = 0;
  uint32_t zeros = 0, matches = 0;
  unsigned long long s1 = 0, s2 = 0;
  const int npx = rows * P.w;
  for (int base = 0, round = 0; base < npx; base += PNG_ROUND_PX, ++round) {
    ThreadTokens tt;
    const int q_end = min(base + PNG_ROUND_PX, npx);
    const int mb_base = multiples_before(P, base);
    const int n_round = 4 * (q_end - base) + multiples_before(P, q_end) - mb_base;
    const bool plain = round_load(P, y0, npx, base, mb_base, tt);
    if (plain) round_starts<true>(tt, rs, round & 1);
    else round_starts<false>(tt, rs, round & 1);
    __syncthreads();
    const int64_t left = n - (4 * (int64_t)base + mb_base);     // stream bytes from the round's first byte on
    if (plain) hist_round<true>(tt, rs, round & 1, n_round, left, hist[wp], zeros, matches, s1, s2);
    else hist_round<false>(tt, rs, round & 1, n_round, left, hist[wp], zeros, matches, s1, s2);
    {
      const int q0 = base + 2 * t;
      uint32_t* fp = P.fpx + (int64_t)y0 * P.w + q0;
      if (q0 < npx) fp[0] = tt.f0;
      if (q0 + 1 < npx) fp[1] = tt.f1;
      P.desc[((int64_t)band * P.rounds + round) * PNG_THREADS + t] = (uint64_t)tt.cls | ((uint64_t)tt.lens << 20);
    }
  }
  if (matches) atomicAdd(&s_nmatch, matches);
  atomicAdd(&hist[wp][0], zeros);
  // block reduction of the Adler sums
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, d);
    s2 += __shfl_xor_sync(0xffffffffu, s2, d);
  }
  if (lane == 0) { red[0][wp] = s1; red[1][wp] = s2; }
  __syncthreads();
  for (int i = t; i < PNG_NSYM; i += PNG_THREADS) {
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < PNG_THREADS / 32; ++k) v += hist[k][i];
    if (i == 256) v = 1;                        // end of block
    freq[i] = v;
  }
  if (t == 0) {
    unsigned long long a = 0, b = 0;
    for (int k = 0; k < PNG_THREADS / 32; ++k) { a += red[0][k]; b += red[1][k]; }
    P.adler[2 * band] = (uint32_t)((1 + a) % 65521ull);
    P.adler[2 * band + 1] = (uint32_t)(((unsigned long long)n + b) % 65521ull);
    P.nmatch[band] = s_nmatch;
  }
  __syncthreads();
  // ---- length-limited Huffman code of freq[0..256] ----
  // Same tree as "merge the two smallest (weight, index) until one node is left": the leaves are ranked by
  // (frequency, index) by counting, then one thread runs the two-queue merge (internal nodes are created in
  // non-decreasing weight order, so the two queue heads are always the two smallest; on a tie the leaf -- lower
  // index -- goes first), every thread walks its leaf up to the root for the depth, and the canonical code of a
  // symbol is next[len] + the number of lower symbols of the same length.
  uint8_t* lens = P.lens + (int64_t)band * PNG_TAB_STRIDE;
  uint16_t* codes = P.codes + (int64_t)band * PNG_TAB_STRIDE;
  __shared__ uint16_t order[PNG_NSYM];
  __shared__ int s_nleaf, s_maxd;
  __shared__ int bl_count[16], next_code[16];
  for (;;) {
    if (t == 0) { s_nleaf = 0; s_maxd = 0; }
    for (int i = t; i < 2 * PNG_NSYM; i += PNG_THREADS) parent[i] = 0xffff;
    __syncthreads();
    for (int i = t; i < PNG_NSYM; i += PNG_THREADS) {
      const uint32_t f = freq[i];
      if (f == 0) continue;
      int rank = 0;
      for (int k = 0; k < PNG_NSYM; ++k) {
        const uint32_t g = freq[k];
        rank += (g != 0 && (g < f || (g == f && k < i))) ? 1 : 0;
      }
      order[rank] = (uint16_t)i;
      atomicAdd(&s_nleaf, 1);
    }
    __syncthreads();
    if (t == 0) {
      const int nl = s_nleaf;
      int qa = 0, qb = PNG_NSYM, nn = PNG_NSYM;     // heads of the leaf queue (order[]) and the internal-node queue
      for (int m = 0; m < nl - 1; ++m) {
        int pick[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const bool has_a = qa < nl, has_b = qb < nn;
          const uint32_t wa = has_a ? freq[order[qa]] : 0xffffffffu, wb = has_b ? freq[qb] : 0xffffffffu;
          if (has_a && (!has_b || wa <= wb)) pick[e] = order[qa++];
          else pick[e] = qb++;
        }
        freq[nn] = freq[pick[0]] + freq[pick[1]];
        parent[pick[0]] = (uint16_t)nn;
        parent[pick[1]] = (uint16_t)nn;
        ++nn;
      }
    }
    __syncthreads();
    for (int i = t; i < PNG_NSYM; i += PNG_THREADS) {
      int d = 0;
      if (freq[i] != 0) {
        int p = parent[i];
        while (p != 0xffff) { ++d; p = parent[p]; }
        if (s_nleaf == 1) d = 1;
      }
      depth[i] = (uint8_t)min(d, 255);
      if (d > 15) atomicMax(&s_maxd, d);
    }
    __syncthreads();
    if (s_maxd <= 15) break;
    for (int i = t; i < PNG_NSYM; i += PNG_THREADS)
      if (freq[i] > 0) freq[i] = (freq[i] + 1) >> 1;
    __syncthreads();
  }
  if (t < 16) bl_count[t] = 0;
  __syncthreads();
  for (int i = t; i < PNG_NSYM; i += PNG_THREADS) {
    lens[i] = depth[i];
    if (depth[i]) atomicAdd(&bl_count[depth[i]], 1);
  }
  __syncthreads();
  if (t == 0) {   // RFC 1951 3.2.2
    int c = 0;
    next_code[0] = 0;
    for (int bb = 1; bb <= 15; ++bb) { c = (c + (bb > 1 ? bl_count[bb - 1] : 0)) << 1; next_code[bb] = c; }
  }
  __syncthreads();
  for (int i = t; i < PNG_NSYM; i += PNG_THREADS) {
    const int l = depth[i];
    uint32_t v = 0;
    if (l) {
      int lower = 0;
      for (int k = 0; k < i; ++k) lower += depth[k] == l ? 1 : 0;
      v = __brev((uint32_t)(next_code[l] + lower)) >> (32 - l);
    }
    codes[i] = (uint16_t)v;
  }
}

// ---- P2 ----
// A thread's bits for one round: up to 10 literals of <= 15 bits (2 pixels, each possibly behind a filter byte;
// a match token is <= 21 bits but covers >= 3 bytes) = 150 bits = 5 words.  Codes are shifted into a 64-bit
// register, LSB first as deflate packs them, and drained one word at a time into the thread's own column of a
// shared-memory staging array (bank = lane, no conflicts; the first version selected one of five registers
// with predicates on every put: half of the kernel's instructions).
struct LocalBits {
  unsigned long long acc;
  int nb;          // bits waiting in acc (< 32 between puts)
  int nw;          // words drained
  int n;           // bits of the round so far
  uint32_t* col;   // word k of this thread at col[k * PNG_THREADS]
};
__device__ __forceinline__ void lb_reset(LocalBits& b, uint32_t* col) {
  b.acc = 0;
  b.nb = b.nw = b.n = 0;
  b.col = col;
}
__device__ __forceinline__ void lb_put(LocalBits& b, uint32_t v, int len) {   // len <= 32
  b.acc |= (unsigned long long)v << b.nb;
  b.nb += len;
  b.n += len;
  if (b.nb >= 32) {
    b.col[b.nw * PNG_THREADS] = (uint32_t)b.acc;
    b.acc >>= 32;
    b.nb -= 32;
    ++b.nw;
  }
}

#define PNG_WIN_WORDS 1280   // 256 threads x 150 bits = 1200 words + carry

__global__ void __launch_bounds__(PNG_THREADS) png_encode_kernel(const __grid_constant__ Params P) {
  __shared__ uint32_t win[PNG_WIN_WORDS];
  __shared__ uint16_t s_code[PNG_NSYM + 7];
  __shared__ uint8_t s_len[PNG_NSYM + 7];
  __shared__ int warp_bits[PNG_THREADS / 32];
  __shared__ uint32_t s_lb[5][PNG_THREADS];
  const int band = blockIdx.x, t = threadIdx.x, lane = t & 31, wp = t >> 5;
  const int y0 = band * PNG_BAND_ROWS, rows = min(PNG_BAND_ROWS, P.h - y0);
  const bool last = band == P.nb - 1;
  uint32_t* const col = &s_lb[0][t];
  for (int i = t; i < PNG_NSYM; i += PNG_THREADS) {
    s_code[i] = P.codes[(int64_t)band * PNG_TAB_STRIDE + i];
    s_len[i] = P.lens[(int64_t)band * PNG_TAB_STRIDE + i];
  }
  for (int i = t; i < PNG_WIN_WORDS; i += PNG_THREADS) win[i] = 0;
  __syncthreads();
  // scratch band = 'IDAT' + zlib data: exactly the bytes the chunk CRC covers
  uint32_t* out = reinterpret_cast<uint32_t*>(P.scratch + (int64_t)band * P.band_cap) + 1;
  if (t == 0) out[-1] = 0x54414449u;
  int64_t words_out = 0;     // words already flushed
  int carry = 0;             // bits waiting at the start of the window

  // One round: every thread contributes `lb`; bits are placed in thread order after `carry`.
  auto place = [&](const LocalBits& lb) {
    if (lb.nb > 0) lb.col[lb.nw * PNG_THREADS] = (uint32_t)lb.acc;   // the partial last word
    int x = lb.n;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) warp_bits[wp] = x;
    __syncthreads();
    int before = carry + x - lb.n, total = carry;
#pragma unroll
    for (int k = 0; k < PNG_THREADS / 32; ++k) {
      const int wb = warp_bits[k];
      if (k < wp) before += wb;
      total += wb;
    }
    if (lb.n > 0) {
      const int wi = before >> 5, bo = before & 31;
      const int nw = (lb.n + 31) >> 5;
      for (int k = 0; k < nw; ++k) {
        const uint32_t w = lb.col[k * PNG_THREADS];
        atomicOr(&win[wi + k], w << bo);
        if (bo) atomicOr(&win[wi + k + 1], w >> (32 - bo));
      }
    }
    __syncthreads();
    // flush the complete words, keep the partial one
    const int full = total >> 5;
    for (int i = t; i < full; i += PNG_THREADS) out[words_out + i] = win[i];
    const uint32_t tailw = win[full];
    __syncthreads();
    for (int i = t; i <= full + 1 && i < PNG_WIN_WORDS; i += PNG_THREADS) win[i] = 0;
    __syncthreads();
    if (t == 0) win[0] = tailw;
    __syncthreads();
    words_out += full;
    carry = total & 31;
  };

  // ---- zlib header (first band) + block header: BFINAL, BTYPE = 10, HLIT = 29, HDIST = 0, HCLEN = 15, the
  // 19 code-length-code lengths (16, 17, 18 unused; 0..15 four bits each), then 286 literal/length code lengths
  // and the single distance code length (1 when the band has a match) as 4-bit codes ----
  {
    LocalBits lb;
    lb_reset(lb, col);
    if (t == 0) {
      if (band == 0) { lb_put(lb, 0x78u, 8); lb_put(lb, 0x01u, 8); }
      lb_put(lb, last ? 1u : 0u, 1);
      lb_put(lb, 2u, 2);
      lb_put(lb, (uint32_t)(PNG_NSYM - 257), 5);
      lb_put(lb, 0u, 5);
      lb_put(lb, 15u, 4);
      lb_put(lb, 0u, 9);                          // 16, 17, 18: length 0
      for (int i = 0; i < 16; ++i) lb_put(lb, 4u, 3);
    }
    place(lb);
    const int dist_len = P.nmatch[band] ? 1 : 0;
#pragma unroll
    for (int part = 0; part < 2; ++part) {
      lb_reset(lb, col);
      const int i = part * PNG_THREADS + t;
      if (i <= PNG_NSYM) {
        const int s = i < PNG_NSYM ? s_len[i] : dist_len;
        lb_put(lb, (uint32_t)(__brev((uint32_t)s) >> 28), 4);
      }
      place(lb);
    }
  }
  // ---- the band's tokens: 2 pixels per thread and round, as P1 left them ----
  const int npx = rows * P.w;
  for (int base = 0; base < npx; base += PNG_ROUND_PX) {
    const int q0 = base + 2 * t;
    const uint64_t d = P.desc[((int64_t)band * P.rounds + base / PNG_ROUND_PX) * PNG_THREADS + t];
    const uint32_t* fp = P.fpx + (int64_t)y0 * P.w + q0;
    const uint32_t f0 = q0 < npx ? fp[0] : 0u, f1 = q0 + 1 < npx ? fp[1] : 0u;
    LocalBits lb;
    lb_reset(lb, col);
    int mi = 0;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      const uint32_t cls = (uint32_t)(d >> (2 * i)) & 3u;
      if (cls == 1u) {
        const uint32_t v = (i % 5 == 0) ? 4u : ((i < 5 ? f0 : f1) >> (8 * ((i % 5) - 1))) & 255u;
        lb_put(lb, s_code[v], s_len[v]);
      } else if (cls == 2u) {
        int ex, nex;
        const int c = len_symbol(3 + (int)((d >> (20 + 8 * mi)) & 255ull), &ex, &nex);
        ++mi;
        // length code, its extra bits, then the 1-bit distance code (0): one field of <= 21 bits
        lb_put(lb, (uint32_t)s_code[c] | ((uint32_t)ex << s_len[c]), s_len[c] + nex + 1);
      }
    }
    place(lb);
  }
  // ---- end of block; byte alignment: empty stored block, or (last band) zero padding + room for the Adler-32 ----
  {
    LocalBits lb;
    lb_reset(lb, col);
    if (t == 0) {
      lb_put(lb, s_code[256], s_len[256]);
      if (!last) lb_put(lb, 0u, 3);
      const int used = (carry + lb.n) & 7;
      if (used) lb_put(lb, 0u, 8 - used);
      if (!last) { lb_put(lb, 0x0000u, 16); lb_put(lb, 0xFFFFu, 16); }
    }
    place(lb);
  }
  // the stream is byte-aligned now; write the partial word and the length
  if (t == 0) {
    out[words_out] = win[0];
    P.blen[band] = (uint32_t)(words_out * 4 + (carry >> 3)) + (last ? 4u : 0u);
  }
}

// ---- P3 ----
__global__ void __launch_bounds__(1024) png_offsets_kernel(const __grid_constant__ Params P, const __grid_constant__ HeaderBytes H) {
  __shared__ unsigned long long warp_sum[32];
  __shared__ unsigned long long carry_s;
  __shared__ uint32_t s_ad[3][1024];            // per-thread Adler pair + length (mod 65521) of a slice of bands
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  if (t == 0) carry_s = (unsigned long long)H.n;
  {
    const int per = (P.nb + 1023) / 1024;
    unsigned long long a = 1, b = 0, nn = 0;
    for (int k = t * per; k < min((t + 1) * per, P.nb); ++k) {
      const int rows = min(PNG_BAND_ROWS, P.h - k * PNG_BAND_ROWS);
      const unsigned long long n = (((unsigned long long)P.w * 4 + 1) * rows) % 65521ull;
      const unsigned long long a2 = P.adler[2 * k], b2 = P.adler[2 * k + 1];
      b = (b + b2 + n * ((a + 65520ull) % 65521ull)) % 65521ull;
      a = (a + a2 + 65520ull) % 65521ull;
      nn = (nn + n) % 65521ull;
    }
    s_ad[0][t] = (uint32_t)a;
    s_ad[1][t] = (uint32_t)b;
    s_ad[2][t] = (uint32_t)nn;
  }
  __syncthreads();
  for (int base = 0; base < P.nb; base += 1024) {
    const int i = base + t;
    const unsigned long long v = i < P.nb ? (unsigned long long)P.blen[i] + 12ull : 0ull;
    unsigned long long x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) warp_sum[w] = x;
    __syncthreads();
    if (w == 0) {
      unsigned long long s = warp_sum[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, s, d);
        if (lane >= d) s += y;
      }
      warp_sum[lane] = s;
    }
    __syncthreads();
    const unsigned long long before = carry_s + (w ? warp_sum[w - 1] : 0ull) + x - v;
    if (i < P.nb) P.off[i] = before;
    __syncthreads();
    if (t == 1023) carry_s = before + v;
    __syncthreads();
  }
  // Adler-32 of the whole stream: thread t combined its slice of bands above; combine the slices in order with a
  // butterfly (the operation is associative; (1, 0, 0) is its identity)
  {
    unsigned long long a = s_ad[0][t], b = s_ad[1][t], nn = s_ad[2][t];
    auto level = [&](unsigned long long oa, unsigned long long ob, unsigned long long on, bool i_am_right) {
      const unsigned long long la = i_am_right ? oa : a, lb = i_am_right ? ob : b, ln = i_am_right ? on : nn;
      const unsigned long long ra = i_am_right ? a : oa, rb = i_am_right ? b : ob, rn = i_am_right ? nn : on;
      b = (lb + rb + rn * ((la + 65520ull) % 65521ull)) % 65521ull;   // + n_right * (a_left - 1)
      a = (la + ra + 65520ull) % 65521ull;                            // a_left + a_right - 1
      nn = (ln + rn) % 65521ull;
    };
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long oa = __shfl_xor_sync(0xffffffffu, a, d), ob = __shfl_xor_sync(0xffffffffu, b, d),
                               on = __shfl_xor_sync(0xffffffffu, nn, d);
      level(oa, ob, on, (lane & d) != 0);
    }
    __syncthreads();
    if (lane == 0) { s_ad[0][w] = (uint32_t)a; s_ad[1][w] = (uint32_t)b; s_ad[2][w] = (uint32_t)nn; }
    __syncthreads();
    if (w == 0) {
      a = s_ad[0][lane]; b = s_ad[1][lane]; nn = s_ad[2][lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long oa = __shfl_xor_sync(0xffffffffu, a, d), ob = __shfl_xor_sync(0xffffffffu, b, d),
                                 on = __shfl_xor_sync(0xffffffffu, nn, d);
        level(oa, ob, on, (lane & d) != 0);
      }
      if (lane == 0) P.info[2] = (b << 16) | a;
    }
  }
  if (t == 0) {
    const unsigned long long total = carry_s + 12ull;
    const bool fits = total <= (unsigned long long)P.cap;
    P.info[0] = total;
    P.info[1] = fits ? 0ull : 1ull;
    if (fits) {
      for (int k = 0; k < H.n; ++k) P.file[k] = H.b[k];
      const uint8_t iend[12] = {0, 0, 0, 0, 'I', 'E', 'N', 'D', 0xAE, 0x42, 0x60, 0x82};
      for (int k = 0; k < 12; ++k) P.file[total - 12 + k] = iend[k];
    }
  }
}

// ---- P4 ----
__device__ __forceinline__ uint32_t gf2_mul(uint32_t a, uint32_t b) {   // a * b mod P, reflected bit order
  uint32_t m = 1u << 31, p = 0;
  for (;;) {
    if (a & m) {
      p ^= b;
      if ((a & (m - 1)) == 0) break;
    }
    m >>= 1;
    b = (b & 1u) ? (b >> 1) ^ PNG_CRC_POLY : b >> 1;
  }
  return p;
}
// The chunk = length, then 'IDAT' + data (the band in scratch, contiguous), then the CRC-32 of those bytes.
// Copy: aligned words of the file, each funnel-shifted from two scratch words.  CRC: the message is padded with
// zero words at the END to whole tiles of PNG_THREADS x 16 words; per tile a thread runs slicing-by-4 over its 16
// consecutive words (initial value 0), `acc = acc * x^(8 * tile bytes) + c` keeps its share of the tiles, a
// butterfly over the threads (`left * x^(8 * 64 * 2^l) + right`) gives the CRC of the padded message with a zero
// initial value, a multiplication by x^(-8 * pad) takes the padding back, and complementing the first four bytes
// and the result stands for the 0xFFFFFFFF initial and final values.
__global__ void __launch_bounds__(PNG_THREADS) png_gather_kernel(const __grid_constant__ Params P) {
  __shared__ uint32_t tab[4][256];
  __shared__ uint32_t s_part[PNG_THREADS / 32];
  if (P.info[1]) return;
  const int band = blockIdx.x, t = threadIdx.x, lane = t & 31, wp = t >> 5;
  const uint32_t dl = P.blen[band];
  uint32_t* srcw = reinterpret_cast<uint32_t*>(P.scratch + (int64_t)band * P.band_cap);
  uint8_t* dst = P.file + P.off[band];
  const bool last = band == P.nb - 1;
  const uint32_t total = dl + 4;                  // bytes under the CRC
  const uint32_t nw = (total + 3) >> 2;
  uint32_t c0 = (uint32_t)t;
#pragma unroll
  for (int k = 0; k < 8; ++k) c0 = (c0 & 1u) ? PNG_CRC_POLY ^ (c0 >> 1) : c0 >> 1;
  tab[0][t] = c0;
  if (last && t == 0) {                           // the stream's Adler-32 closes the last band
    uint8_t* sb = reinterpret_cast<uint8_t*>(srcw);
    for (int k = 0; k < 4; ++k) sb[total - 4 + k] = (uint8_t)(P.info[2] >> (24 - 8 * k));
  }
  __syncthreads();
  {
    uint32_t c = c0;
#pragma unroll
    for (int k = 1; k < 4; ++k) { c = tab[0][c & 255u] ^ (c >> 8); tab[k][t] = c; }
  }
  // ---- copy ----
  if (t < 4) dst[t] = (uint8_t)(dl >> (24 - 8 * t));
  {
    uint8_t* D = dst + 4;
    const uint8_t* sb = reinterpret_cast<const uint8_t*>(srcw);
    const uint32_t h = min((uint32_t)((4 - (reinterpret_cast<uintptr_t>(D) & 3)) & 3), total);   // bytes before the first aligned word
    const uint32_t nfw = (total - h) >> 2;
    uint32_t* Dw = reinterpret_cast<uint32_t*>(D + h);
    for (uint32_t j = t; j < nfw; j += PNG_THREADS) {
      const uint32_t lo = srcw[j], hi = h ? srcw[j + 1] : 0u;
      Dw[j] = __funnelshift_r(lo, hi, 8 * h);
    }
    if ((uint32_t)t < h) D[t] = sb[t];
    const uint32_t tail0 = h + 4 * nfw;
    if (tail0 + t < total && t < 4) D[tail0 + t] = sb[tail0 + t];
  }
  __syncthreads();                                // tables ready
  // ---- CRC ----
  const uint32_t tile_words = PNG_THREADS * 16;
  const uint32_t ntiles = (nw + tile_words - 1) / tile_words;
  const uint32_t tail_mask = (total & 3u) ? (1u << (8 * (total & 3u))) - 1u : 0xFFFFFFFFu;
  uint32_t acc = 0;
  for (uint32_t tile = 0; tile < ntiles; ++tile) {
    const uint32_t g0 = tile * tile_words + 16u * (uint32_t)t;
    uint32_t c = 0;
    if (g0 < nw) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (g0 + 4 * q < nw) v = *reinterpret_cast<const uint4*>(srcw + g0 + 4 * q);
        const uint32_t wv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const uint32_t g = g0 + 4 * q + r;
          uint32_t x = g < nw ? wv[r] : 0u;
          if (g == nw - 1) x &= tail_mask;
          if (g == 0) x = ~x;
          c ^= x;
          c = tab[3][c & 255u] ^ tab[2][(c >> 8) & 255u] ^ tab[1][(c >> 16) & 255u] ^ tab[0][c >> 24];
        }
      }
    }
    acc = (tile ? gf2_mul(P.x2n[17], acc) : 0u) ^ c;      // tile = 2^14 bytes: x^(8 * 2^14) = x^(2^17)
  }
  // butterfly: thread groups of 2^l neighbours, 64 * 2^l bytes each: x^(8 * 64 * 2^l) = x^(2^(9 + l))
#pragma unroll
  for (int l = 0; l < 5; ++l) {
    const uint32_t other = __shfl_xor_sync(0xffffffffu, acc, 1 << l);
    const bool right = (lane >> l) & 1;
    acc = gf2_mul(P.x2n[9 + l], right ? other : acc) ^ (right ? acc : other);
  }
  if (lane == 0) s_part[wp] = acc;
  __syncthreads();
  if (wp != 0) return;
  acc = lane < PNG_THREADS / 32 ? s_part[lane] : 0u;
#pragma unroll
  for (int l = 0; l < 3; ++l) {
    const uint32_t other = __shfl_xor_sync(0xffffffffu, acc, 1 << l);
    const bool right = (lane >> l) & 1;
    acc = gf2_mul(P.x2n[14 + l], right ? other : acc) ^ (right ? acc : other);
  }
  // x^(-8 * pad): one factor per bit of 8 * pad, multiplied up by a butterfly
  const uint32_t pad_bits = 8u * (ntiles * tile_words * 4u - total);
  uint32_t inv = (pad_bits >> lane & 1u) ? P.xinv2n[lane] : 0x80000000u;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) inv = gf2_mul(inv, __shfl_xor_sync(0xffffffffu, inv, d));
  if (lane == 0) {
    const uint32_t crc = gf2_mul(inv, acc) ^ 0xFFFFFFFFu;
    uint8_t* e = dst + 8 + dl;
    e[0] = (uint8_t)(crc >> 24); e[1] = (uint8_t)(crc >> 16); e[2] = (uint8_t)(crc >> 8); e[3] = (uint8_t)crc;
  }
}

}  // namespace png
}  // namespace stitch
