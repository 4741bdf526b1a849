// PNG export on the GPU (STITCH_OUT_PNG): what the page asks canvasToTempFilePath for (fileType 'png',
// U:205-242, called at P:1577-1579).  ISO/IEC 15948 colour type 6, 8 bit; IDAT = zlib (RFC 1950) / deflate
// (RFC 1951).  The encoder's free choices are made for parallelism:
//   * Paeth filter on every row (the predictor only looks at the canvas, never at compressed state);
//   * the rows are cut into bands of PNG_BAND_ROWS rows, one CTA each; a band is ONE dynamic-Huffman block of
//     literals (no LZ77 matches) + -- except for the last band -- an empty stored block that byte-aligns the
//     stream, and is its own IDAT chunk, so bands are coded, CRC'd and placed independently;
//   * the Adler-32 of the whole stream is combined from per-band sums.
// P1 png_hist_kernel    filtered-byte histogram + Adler partial sums per band, then warp 0 builds the length-
//                       limited Huffman code (merge the two smallest (frequency, index); too deep -> halve the
//                       frequencies and rebuild) and its canonical, bit-reversed codes
// P2 png_encode_kernel  block header + codes: per chunk every thread codes 2 pixels into registers, a CTA scan of
//                       the bit counts places them in a shared-memory bit window that is flushed as words
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
#define PNG_NSYM 257
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
  uint8_t* lens;               // [nb][PNG_TAB_STRIDE]
  uint16_t* codes;             // [nb][PNG_TAB_STRIDE]
  uint32_t* adler;             // [nb][2]: Adler pair of the band's bytes started from (1, 0)
  uint8_t* scratch;            // [nb][band_cap] coded bands
  int64_t band_cap;
  uint32_t* blen;              // [nb] bytes of zlib data per band (the last one includes the 4 Adler bytes)
  uint64_t* off;               // [nb] file offset of the band's chunk
  uint64_t* info;              // {file size, overflow flag}
  uint8_t* file;
  int64_t cap;
  uint32_t x2n[32];            // x^(2^k) mod P, reflected (CRC combine)
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
  uint32_t zeros = 0;
  unsigned long long s1 = 0, s2 = 0;
  const int npx = rows * P.w;
  int r = t / P.w, j = t - r * P.w;
  for (int q = t; q < npx; q += PNG_THREADS) {
    const uint32_t f = filter_px(P, j, y0 + r);
    const int64_t pos = r * rowlen + 1 + (int64_t)j * 4;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t d = (f >> (8 * k)) & 255u;
      if (d == 0) ++zeros;
      else {
        atomicAdd(&hist[wp][d], 1u);
        s1 += d;
        s2 += (unsigned long long)(n - (pos + k)) * d;
      }
    }
    if (j == 0) {                               // the row's filter-type byte (4 = Paeth)
      s1 += 4;
      s2 += (unsigned long long)(n - r * rowlen) * 4;
    }
    j += PNG_THREADS;
    while (j >= P.w) { j -= P.w; ++r; }
  }
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
    if (i == 4) v += (uint32_t)rows;
    if (i == 256) v = 1;                        // end of block
    freq[i] = v;
  }
  if (t == 0) {
    unsigned long long a = 0, b = 0;
    for (int k = 0; k < PNG_THREADS / 32; ++k) { a += red[0][k]; b += red[1][k]; }
    P.adler[2 * band] = (uint32_t)((1 + a) % 65521ull);
    P.adler[2 * band + 1] = (uint32_t)(((unsigned long long)n + b) % 65521ull);
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
// A thread's bits for one chunk: up to 9 symbols of <= 15 bits (2 pixels + a filter byte) = 135 bits.
struct LocalBits {
  uint32_t w[5];
  int n;
};
__device__ __forceinline__ void lb_put(LocalBits& b, uint32_t v, int len) {
  const int wi = b.n >> 5, bo = b.n & 31;
  // static indexing only: select the word with predicates
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    if (k == wi) b.w[k] |= v << bo;
    if (k == wi + 1 && bo + len > 32) b.w[k] |= v >> (32 - bo);
  }
  b.n += len;
}

#define PNG_WIN_WORDS 1152   // 256 threads x 135 bits = 1080 words + carry

__global__ void __launch_bounds__(PNG_THREADS) png_encode_kernel(const __grid_constant__ Params P) {
  __shared__ uint32_t win[PNG_WIN_WORDS];
  __shared__ uint16_t s_code[PNG_NSYM + 7];
  __shared__ uint8_t s_len[PNG_NSYM + 7];
  __shared__ int warp_bits[PNG_THREADS / 32];
  const int band = blockIdx.x, t = threadIdx.x, lane = t & 31, wp = t >> 5;
  const int y0 = band * PNG_BAND_ROWS, rows = min(PNG_BAND_ROWS, P.h - y0);
  const bool last = band == P.nb - 1;
  for (int i = t; i < PNG_NSYM; i += PNG_THREADS) {
    s_code[i] = P.codes[(int64_t)band * PNG_TAB_STRIDE + i];
    s_len[i] = P.lens[(int64_t)band * PNG_TAB_STRIDE + i];
  }
  for (int i = t; i < PNG_WIN_WORDS; i += PNG_THREADS) win[i] = 0;
  __syncthreads();
  uint32_t* out = reinterpret_cast<uint32_t*>(P.scratch + (int64_t)band * P.band_cap);
  int64_t words_out = 0;     // words already flushed
  int carry = 0;             // bits waiting at the start of the window

  // One round: every thread contributes `lb`; bits are placed in thread order after `carry`.
  auto place = [&](const LocalBits& lb) {
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
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        if (k < nw) {
          atomicOr(&win[wi + k], lb.w[k] << bo);
          if (bo) atomicOr(&win[wi + k + 1], lb.w[k] >> (32 - bo));
        }
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

  // ---- zlib header (first band) + block header: BFINAL, BTYPE = 10, HLIT = 0, HDIST = 0, HCLEN = 15, the
  // 19 code-length-code lengths (16, 17, 18 unused; 0..15 four bits each), then 258 lengths as 4-bit codes ----
  {
    LocalBits lb;
#pragma unroll
    for (int k = 0; k < 5; ++k) lb.w[k] = 0;
    lb.n = 0;
    if (t == 0) {
      if (band == 0) { lb_put(lb, 0x78u, 8); lb_put(lb, 0x01u, 8); }
      lb_put(lb, last ? 1u : 0u, 1);
      lb_put(lb, 2u, 2);
      lb_put(lb, 0u, 5);
      lb_put(lb, 0u, 5);
      lb_put(lb, 15u, 4);
      lb_put(lb, 0u, 9);                          // 16, 17, 18: length 0
      for (int i = 0; i < 16; ++i) lb_put(lb, 4u, 3);
    }
    place(lb);
#pragma unroll
    for (int k = 0; k < 5; ++k) lb.w[k] = 0;
    lb.n = 0;
    {
      const int s = t < PNG_NSYM ? s_len[t] : 0;
      lb_put(lb, (uint32_t)(__brev((uint32_t)s) >> 28), 4);
      if (t == PNG_THREADS - 1) {                 // symbol 256 and the single distance code (length 0)
        const int s2 = s_len[256];
        lb_put(lb, (uint32_t)(__brev((uint32_t)s2) >> 28), 4);
        lb_put(lb, 0u, 4);
      }
    }
    place(lb);
  }
  // ---- the band's bytes: 2 pixels per thread and round ----
  const int npx = rows * P.w;
  for (int base = 0; base < npx; base += 2 * PNG_THREADS) {
    LocalBits lb;
#pragma unroll
    for (int k = 0; k < 5; ++k) lb.w[k] = 0;
    lb.n = 0;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int q = base + 2 * t + e;
      if (q < npx) {
        const int r = q / P.w, j = q - r * P.w;
        if (j == 0) lb_put(lb, s_code[4], s_len[4]);
        const uint32_t f = filter_px(P, j, y0 + r);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t d = (f >> (8 * k)) & 255u;
          lb_put(lb, s_code[d], s_len[d]);
        }
      }
    }
    place(lb);
  }
  // ---- end of block; byte alignment: empty stored block, or (last band) zero padding + room for the Adler-32 ----
  {
    LocalBits lb;
#pragma unroll
    for (int k = 0; k < 5; ++k) lb.w[k] = 0;
    lb.n = 0;
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
  if (t == 0) {
    const unsigned long long total = carry_s + 12ull;
    const bool fits = total <= (unsigned long long)P.cap;
    P.info[0] = total;
    P.info[1] = fits ? 0ull : 1ull;
    // Adler-32 of the whole stream: thread t combined its slice of bands above (s_ad); fold the slices in order
    unsigned long long a = 1, b = 0;
    for (int k = 0; k < 1024; ++k) {
      const unsigned long long a2 = s_ad[0][k], b2 = s_ad[1][k], n2 = s_ad[2][k];
      b = (b + b2 + n2 * ((a + 65520ull) % 65521ull)) % 65521ull;   // + n * (a - 1)
      a = (a + a2 + 65520ull) % 65521ull;                          // a + a2 - 1
    }
    P.info[2] = (b << 16) | a;
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
__device__ __forceinline__ uint32_t x_pow_8n(const Params& P, uint32_t nbytes) {  // x^(8 n) mod P
  uint32_t p = 1u << 31;
  uint32_t k = 3;
  while (nbytes) {
    if (nbytes & 1u) p = gf2_mul(P.x2n[k & 31], p);
    nbytes >>= 1;
    ++k;
  }
  return p;
}

__global__ void __launch_bounds__(PNG_THREADS) png_gather_kernel(const __grid_constant__ Params P) {
  __shared__ uint32_t tab[256];
  __shared__ uint32_t part_crc[32], part_len[32];
  if (P.info[1]) return;
  const int band = blockIdx.x, t = threadIdx.x, lane = t & 31;
  {
    uint32_t c = (uint32_t)t;
#pragma unroll
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? PNG_CRC_POLY ^ (c >> 1) : c >> 1;
    tab[t] = c;
  }
  const uint32_t dl = P.blen[band];
  const uint8_t* src = P.scratch + (int64_t)band * P.band_cap;
  uint8_t* dst = P.file + P.off[band];
  const bool last = band == P.nb - 1;
  const uint32_t copy_n = last ? dl - 4 : dl;
  for (uint32_t i = t; i < copy_n; i += PNG_THREADS) dst[8 + i] = src[i];
  if (t < 4) {
    dst[t] = (uint8_t)(dl >> (24 - 8 * t));
    dst[4 + t] = (uint8_t)("IDAT"[t]);
    if (last) dst[8 + copy_n + t] = (uint8_t)(P.info[2] >> (24 - 8 * t));
  }
  __syncthreads();
  // CRC-32 over 'IDAT' + data = dst[4 .. 8 + dl): warp 0, each lane a contiguous slice, then combine in order
  if (t >= 32) return;
  const uint32_t total = dl + 4;
  const uint32_t per = (total + 31) / 32;
  const uint32_t b0 = min(lane * per, total), b1 = min(b0 + per, total);
  uint32_t c = 0xFFFFFFFFu;
  for (uint32_t i = b0; i < b1; ++i) c = tab[(c ^ dst[4 + i]) & 255u] ^ (c >> 8);
  part_crc[lane] = c ^ 0xFFFFFFFFu;
  part_len[lane] = b1 - b0;
  __syncwarp();
  if (lane == 0) {
    uint32_t crc = part_crc[0];
    const uint32_t xp = x_pow_8n(P, per);          // all slices but the last have `per` bytes
    for (int k = 1; k < 32; ++k) {
      if (part_len[k] == 0) continue;
      crc = gf2_mul(part_len[k] == per ? xp : x_pow_8n(P, part_len[k]), crc) ^ part_crc[k];
    }
    uint8_t* e = dst + 8 + dl;
    e[0] = (uint8_t)(crc >> 24); e[1] = (uint8_t)(crc >> 16); e[2] = (uint8_t)(crc >> 8); e[3] = (uint8_t)crc;
  }
}

}  // namespace png
}  // namespace stitch
