// Baseline JPEG decoding on the GPU: replaces the bitmap decode inside loadImageFrom
// (U:27-121, P:1441-1520) so a job can take the FILES instead of 48.8 MB RGBA bitmaps per image.
//
// Huffman-coded data has no random access, so the entropy decoder is the self-synchronising one
// (Klein & Wiseman 2003; Weissenberger & Schmidt 2018): the scan is cut into subsequences of DecBatch::sub (128 ... 512) raw
// bytes, one THREAD each.
//   D1 first pass   every thread decodes its subsequence from its first bit as if a block started there and
//                   records where it ended: (byte, bit, block-in-MCU, zig-zag index) -- the state the next
//                   subsequence has to start from IF this one started right.
//   D2 sync         repeated: a thread restarts from its predecessor's end state; when its own end state stops
//                   changing the chain from the (known) start of the scan is consistent.  Huffman codes
//                   re-synchronise after a few symbols, so this converges in a few rounds; restart markers
//                   are hard synchronisation points.
//   D3 offsets      exclusive scan of the number of blocks each subsequence starts.
//   D4 write        decode once more from the true start states, coefficients into scan-order blocks.
//   D5 dc           DC differences -> values: segmented scan over MCUs (segments = restart intervals).
//   D6 idct         one thread per block: dequantise, integer LL&M inverse DCT (13-bit constants), planes.
//   D7 colour       triangle-filter chroma upsampling + 16-bit fixed-point YCbCr -> RGB, RGBA8 out.
// All arithmetic is integer and follows libjpeg's default pipeline; tests hold the bitmap byte-identical to
// the sequential CPU decoder of the test tree, which is itself pinned to libjpeg.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "jpeg_parse.h"

namespace stitch {
namespace jpegdec {

// Raw bytes per subsequence (DecBatch::sub), chosen per batch by jdec_pick_sub: 512 when that still leaves >= 32 k
// threads, else 256, else 128.  Long subsequences mean fewer links in the synchronisation chains (the sync kernel's
// in-block iterations) and fewer wasted symbols after each guessed start: 9 x 12 MP files decode + stitch + export in
// 3.2 ms/job with 512 against 4.0 with 128 (profiles/r2_decode_ab.md); a single small file keeps 128 for parallelism.
#define JDEC_SUB_MIN 128
#define JDEC_SUB_MAX 512
inline uint32_t jdec_pick_sub(size_t total_scan_bytes) {
  if (const char* e = std::getenv("STITCH_JDEC_SUB")) {   // tests force every size on small files
    const int v = std::atoi(e);
    if (v == 128 || v == 256 || v == 512) return (uint32_t)v;
  }
  for (uint32_t sub = JDEC_SUB_MAX; sub > JDEC_SUB_MIN; sub >>= 1)
    if (total_scan_bytes / sub >= 32768) return sub;
  return JDEC_SUB_MIN;
}
#ifndef JDEC_THREADS
#define JDEC_THREADS 256
#endif
#define JDEC_MAX_IMAGES 9
#define JDEC_SEG 128          // bytes between the checkpoints inside a subsequence (see redecode)

struct DecImage {
  const uint8_t* scan;        // entropy-coded segment (device)
  const Tables* tab;          // device copy of the header tables
  uint32_t scan_len;
  int32_t w, h, ncomp, hs, vs, mcux, mcuy, ri, bpm;   // bpm = blocks per MCU
  int32_t n_sub, n_blocks;
  int32_t td[3], ta[3], tq[3];
  uint4* state;               // [n_sub] end state {byte, bit | b << 8 | z << 16 | done << 24, blocks started, -}
  uint32_t* prefix;           // [n_sub] exclusive scan of blocks started
  int16_t* coef;              // [n_blocks][64] scan order, zig-zag order inside a block
  int32_t* dcsum;             // [3][n_mcu] per-MCU DC sums, then their segmented scan
  uint2* consumed;            // [blocks of the sync grid] predecessor state each block last started from
  uint4* cp;                  // [n_sub][nseg] checkpoints: {byte, packed state, blocks started} at the end of every 128-byte
                              // segment of a subsequence, as its LAST decode saw them (nullptr: no checkpoints)
  uint8_t* plane[3];          // component planes, MCU-padded
  int32_t pw[3], ph[3];
  uint8_t* rgba;              // output bitmap
  int64_t pitch;
};
struct DecBatch {
  DecImage img[JDEC_MAX_IMAGES];
  int32_t n;
  uint32_t* changed;          // device flags: a sync round raises changed[flag_cur] when a block's last end state moved
  int32_t round;              // sync round (0 = first)
  int32_t flag_cur;           // flag this round raises
  int32_t flag_prev;          // flag of the round before (-1: none, the host decides whether to launch)
  uint32_t sub;               // raw bytes per subsequence (jdec_pick_sub)
  uint32_t nseg;              // checkpointed segments per subsequence (sub / JDEC_SEG; 1 = none)
  __device__ __forceinline__ const DecImage& item(int i) const { return img[i]; }
};

// ---- Huffman tables in shared memory (SmemTables: jpeg_parse.h) ----

__device__ __forceinline__ void build_tables(const Tables* __restrict__ g, SmemTables& s) {
  for (int i = threadIdx.x; i < 4 * 256; i += blockDim.x) {
    s.vals[i >> 8][i & 255] = g->vals[i >> 8][i & 255];
    s.lut[i >> 8][i & 255] = 0;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    const int t = threadIdx.x;
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
      const int nb = g->bits[t][l - 1];
      s.valptr[t][l] = k;
      s.mincode[t][l] = code;
      if (l <= 8)
        for (int j = 0; j < nb; ++j) {
          const int first = (code + j) << (8 - l);
          for (int f = 0; f < (1 << (8 - l)); ++f) s.lut[t][(first + f) & 255] = (uint16_t)(l << 8 | g->vals[t][(k + j) & 255]);
        }
      code += nb;
      k += nb;
      s.maxcode[t][l] = nb ? code - 1 : -1;
      code <<= 1;
    }
  }
  __syncthreads();
}

// ---- bit reader over the raw (byte-stuffed) scan ----
struct Reader {
  const uint8_t* p;
  uint32_t n;      // scan length
  uint32_t pos;    // next raw byte to fetch
  uint64_t acc;    // unread bits, top aligned
  int cnt;         // bits in acc (real + fake)
  int fake;        // zero bits fed after a marker / the end of the data
  bool mark;       // fetching has reached a marker (p[pos] == 0xFF, p[pos+1] != 0) or the end
};

__device__ __forceinline__ void rd_start(Reader& r, uint32_t byte, int bit) {
  r.pos = byte;
  r.acc = 0;
  r.cnt = 0;
  r.fake = 0;
  r.mark = false;
  (void)bit;
}
__device__ __forceinline__ void rd_fill(Reader& r) {
  if (r.cnt > 32) return;                        // a symbol needs at most 16 + 11 bits
  if (!r.mark && r.pos + 8 <= r.n) {
    // fast path: four raw bytes at once (two aligned words, big-endian, funnel-shifted) when none is 0xFF
    const uint32_t* p32 = reinterpret_cast<const uint32_t*>(r.p);
    const uint32_t w0 = __byte_perm(p32[r.pos >> 2], 0, 0x0123), w1 = __byte_perm(p32[(r.pos >> 2) + 1], 0, 0x0123);
    const uint32_t v = __funnelshift_l(w1, w0, 8 * (r.pos & 3));
    const uint32_t t = ~v;
    if (((t - 0x01010101u) & ~t & 0x80808080u) == 0) {
      r.acc |= (uint64_t)v << (32 - r.cnt);
      r.cnt += 32;
      r.pos += 4;
      return;
    }
  }
  while (r.cnt <= 56) {
    uint32_t v = 0;
    if (!r.mark) {
      if (r.pos < r.n) {
        v = r.p[r.pos];
        if (v == 0xFFu) {
          const uint32_t nx = (r.pos + 1 < r.n) ? r.p[r.pos + 1] : 0xD9u;
          if (nx == 0) r.pos += 2;               // stuffed zero
          else { r.mark = true; v = 0; }         // marker: stay on it
        } else {
          r.pos += 1;
        }
      } else {
        r.mark = true;
      }
    }
    if (r.mark) r.fake += 8;
    r.acc |= (uint64_t)v << (56 - r.cnt);
    r.cnt += 8;
  }
}
__device__ __forceinline__ void rd_skip(Reader& r, int nbits) {
  r.acc <<= nbits;
  r.cnt -= nbits;
}
// Raw byte index + bit offset of the next unread bit (the canonical form states are compared in).
__device__ __forceinline__ void rd_where(const Reader& r, uint32_t* byte, int* bit) {
  const int real = r.cnt - r.fake;               // unread real bits (>= 0 at symbol boundaries)
  const int nb = (real + 7) >> 3;
  uint32_t c = r.pos;
  for (int k = 0; k < nb; ++k) {
    --c;
    if (r.p[c] == 0 && c > 0 && r.p[c - 1] == 0xFFu) --c;   // the data byte is the 0xFF of an FF 00 pair
  }
  *byte = c;
  *bit = (8 - (real & 7)) & 7;
}

struct State {
  uint32_t byte;
  int bit, b, z;
  bool done;       // reached the end of the scan (EOI or any non-RST marker, or the end of the data)
};
__device__ __forceinline__ uint32_t pack_state(const State& s) {
  return (uint32_t)s.bit | (uint32_t)s.b << 8 | (uint32_t)s.z << 16 | (s.done ? 1u << 24 : 0u);
}
__device__ __forceinline__ State unpack_state(uint32_t byte, uint32_t v) {
  State s;
  s.byte = byte;
  s.bit = v & 255;
  s.b = (v >> 8) & 255;
  s.z = (v >> 16) & 255;
  s.done = (v >> 24) & 1;
  return s;
}

__device__ __forceinline__ int huff_decode(const SmemTables& T, int t, Reader& r) {
  const uint32_t v = (uint32_t)(r.acc >> 48);    // 16 bits
  const uint32_t e = T.lut[t][v >> 8];
  if (e) {
    rd_skip(r, e >> 8);
    return e & 255;
  }
  for (int l = 9; l <= 16; ++l) {
    const int code = (int)(v >> (16 - l));
    if (code <= T.maxcode[t][l] && code >= T.mincode[t][l]) {
      rd_skip(r, l);
      return T.vals[t][(T.valptr[t][l] + code - T.mincode[t][l]) & 255];
    }
  }
  rd_skip(r, 16);                                // no such code (only on a wrong path or in a corrupt file)
  return 0;
}

// Decodes the symbols whose first bit lies in raw bytes [.., end_byte), starting from `s`.  Returns the
// state after the last of them and the number of blocks STARTED (DC symbols).  WRITE: also stores the
// coefficients; blk is the scan-order index of the block in progress (or of the next one when s.z == 0).
// WRITE: the coefficients of the block in progress are collected in shared memory (sblk: this thread's 64 int16, laid
// out [e / 8][thread][8]) and go to global memory as whole 16-byte pieces when the block ends -- 8 vector stores per
// block instead of ~20 scattered 2-byte ones.  A block that straddles two subsequences is written by both threads,
// each the zig-zag positions it decoded: [.., z) by the first, [z, 64) by the second.
template <bool WRITE>
__device__ __forceinline__ State decode_range(const DecImage& I, const SmemTables& T, State s, uint32_t end_byte,
                                              uint32_t* started, int64_t blk, int16_t* sblk = nullptr) {
  auto st_at = [&](int e) -> int16_t* { return sblk + (e >> 3) * (JDEC_THREADS * 8) + (e & 7); };
  auto st_zero = [&]() {
#pragma unroll
    for (int g8 = 0; g8 < 8; ++g8) *reinterpret_cast<uint4*>(sblk + g8 * (JDEC_THREADS * 8)) = make_uint4(0u, 0u, 0u, 0u);
  };
  auto st_flush = [&](int16_t* dst, int lo, int hi) {
    if (!dst) return;
    if (lo == 0 && hi == 64) {
#pragma unroll
      for (int g8 = 0; g8 < 8; ++g8)
        reinterpret_cast<uint4*>(dst)[g8] = *reinterpret_cast<const uint4*>(sblk + g8 * (JDEC_THREADS * 8));
    } else {
      for (int e = lo; e < hi; ++e) dst[e] = *st_at(e);
    }
  };
  int lo = s.z;   // first zig-zag position of the block in progress that this thread owns
  Reader r;
  r.p = I.scan;
  r.n = I.scan_len;
  rd_start(r, s.byte, s.bit);
  uint32_t count = 0;
  const int ny = I.hs * I.vs;
  bool first = true;
  int b = s.b, z = s.z;
  State out = s;
  if (s.done) { *started = 0; return s; }
  int16_t* cur = nullptr;
  if (WRITE && z > 0) {
    --blk;
    cur = (blk >= 0 && blk < I.n_blocks) ? I.coef + blk * 64 : nullptr;
    ++blk;
    st_zero();
  }
  for (;;) {
    rd_fill(r);
    if (first) { rd_skip(r, s.bit); first = false; }
    // Position of the next symbol.  With `real` unread bits in nb (possibly partial) data bytes the raw index c
    // of its first byte lies in [pos - 2 nb, pos - nb] (stuffed zeros); only when the upper bound reaches the
    // boundary is the exact index needed, and rd_where walks back over the last bytes to get it.
    {
      const int real = r.cnt - r.fake;
      const uint32_t nb = (uint32_t)((real + 7) >> 3);
      if (r.pos - nb >= end_byte) {
        uint32_t c;
        int cbit;
        rd_where(r, &c, &cbit);
        if (c >= end_byte) {
          out.byte = c; out.bit = cbit; out.b = b; out.z = z; out.done = false;
          if (WRITE && z > 0) st_flush(cur, lo, z);   // the next subsequence's thread writes the rest of this block
          break;
        }
      }
    }
    // ---- one symbol.  DC and AC symbols take the SAME instruction path (table, run and the state update are
    // selected, not branched on): lanes of a warp sit at different places of their blocks, and a branch here
    // would run both sides on almost every iteration. ----
    const int comp = b < ny ? 0 : b - ny + 1;
    const bool is_dc = (z == 0);
    const int rs = huff_decode(T, is_dc ? I.td[comp] : 2 + I.ta[comp], r);
    const int sz = rs & 15, run = is_dc ? 0 : rs >> 4;
    int value = 0;
    if (sz) {
      const uint32_t bits = (uint32_t)(r.acc >> (64 - sz));
      rd_skip(r, sz);
      value = (bits < (1u << (sz - 1))) ? (int)bits - (1 << sz) + 1 : (int)bits;
    }
    // AC with size 0: ZRL (run 15) skips 16 zeros, anything else ends the block; otherwise the coefficient
    // lands at z + run and the next one is expected behind it
    const bool coef = is_dc || sz != 0;
    const int zpos = coef ? z + run : -1;
    z = is_dc ? 1 : (sz != 0 ? z + run + 1 : (run == 15 ? z + 16 : 64));
    const bool block_done = z > 63;
    if (r.cnt < r.fake) {
      // the symbol ran into a marker (it was padding) or past the end: drop it
      uint32_t mpos = r.pos;
      while (mpos + 2 < r.n && r.p[mpos + 1] == 0xFFu) ++mpos;   // fill bytes in front of the marker (T.81 B.1.1.2)
      const uint32_t m = (mpos + 1 < r.n) ? r.p[mpos + 1] : 0xD9u;
      if (mpos + 1 < r.n && m >= 0xD0u && m <= 0xD7u) {   // RSTn: byte-aligned restart, predictors and position reset
        b = 0;
        z = 0;
        rd_start(r, mpos + 2, 0);
        if (mpos + 2 >= end_byte) {
          out.byte = mpos + 2; out.bit = 0; out.b = 0; out.z = 0; out.done = false;
          break;
        }
        continue;
      }
      out.byte = r.n; out.bit = 0; out.b = 0; out.z = 0; out.done = true;
      break;
    }
    if (is_dc) {
      ++count;
      if (WRITE) {
        cur = (blk >= 0 && blk < I.n_blocks) ? I.coef + blk * 64 : nullptr;
        ++blk;
        st_zero();
        lo = 0;
        *st_at(0) = (int16_t)value;
      }
    } else if (WRITE && zpos > 0 && zpos < 64) {
      *st_at(zpos) = (int16_t)value;
    }
    if (block_done) {
      if (WRITE) st_flush(cur, lo, 64);
      z = 0;
      b = (b + 1 == I.bpm) ? 0 : b + 1;
    }
  }
  *started = count;
  return out;
}


// Start state of subsequence i when nothing is known: the first data bit of raw byte i * sub, taken as
// the DC symbol of the first block of an MCU.  (A stuffed zero or the second byte of an RSTn is stepped over.)
__device__ __forceinline__ State guess_state(const DecImage& I, uint32_t i, uint32_t sub) {
  State s;
  s.byte = i * sub;
  s.bit = 0;
  s.b = 0;
  s.z = 0;
  s.done = false;
  if (i > 0 && s.byte < I.scan_len && I.scan[s.byte - 1] == 0xFFu) {
    const uint32_t v = I.scan[s.byte];
    if (v == 0 || (v >= 0xD0u && v <= 0xD7u)) s.byte += 1;
  }
  return s;
}

// What the shared synchronisation / prefix kernels ask of a batch item (progressive scans answer in jpeg_prog.cuh).
__device__ __forceinline__ State decode_nowrite(const DecImage& I, const SmemTables& T, State s, uint32_t end_byte, uint32_t* started) {
  return decode_range<false>(I, T, s, end_byte, started, 0);
}
__device__ __forceinline__ uint32_t scan_start_word(const DecImage&) { return 0u; }   // packed state at byte 0: block 0, DC

// A subsequence decoded again from a new entry state.  With checkpoints (I.cp) the decode stops at the first segment
// boundary where it arrives in the state the PREVIOUS decode of this subsequence had there: from that point on the two
// are the same decode, so the old end state and the old block counts of the remaining segments stand.  Entry states
// that differ only re-synchronise a few symbols in (that is what the whole scheme rests on), so a re-decode costs one
// segment or two instead of the subsequence -- and the synchronisation rounds are chains of re-decodes.
// EARLY = false (the first pass): nothing to compare with yet, the checkpoints are only written.
template <bool EARLY, class BATCH>
__device__ __forceinline__ State redecode(const BATCH& B, const DecImage& I, const SmemTables& T, uint32_t idx, State in, uint32_t* started) {
  const uint32_t nseg = B.nseg;
  if (!I.cp || nseg <= 1) return decode_nowrite(I, T, in, (idx + 1) * B.sub, started);
  uint4* cp = I.cp + (size_t)idx * nseg;
  uint32_t total = 0;
  State cur = in;
  for (uint32_t c = 0; c < nseg; ++c) {
    uint32_t n = 0;
    const State e = decode_nowrite(I, T, cur, idx * B.sub + (c + 1) * JDEC_SEG, &n);
    const uint32_t pk = pack_state(e);
    const uint4 old = cp[c];
    cp[c] = make_uint4(e.byte, pk, n, 0u);
    total += n;
    cur = e;
    if (EARLY && c + 1 < nseg && old.x == e.byte && old.y == pk) {
      for (uint32_t k = c + 1; k < nseg; ++k) total += cp[k].z;
      const uint4 last = cp[nseg - 1];
      cur = unpack_state(last.x, last.y);
      break;
    }
  }
  *started = total;
  return cur;
}

// D1: first pass from guessed states.
__global__ void __launch_bounds__(JDEC_THREADS) jdec_first_kernel(const __grid_constant__ DecBatch B) {
  __shared__ SmemTables T;
  const DecImage& I = B.img[blockIdx.y];
  if (blockIdx.x * JDEC_THREADS >= (unsigned)I.n_sub) return;
  build_tables(I.tab, T);
  const uint32_t i = blockIdx.x * JDEC_THREADS + threadIdx.x;
  if (i >= (uint32_t)I.n_sub) return;
  uint32_t started = 0;
  const State e = redecode<false>(B, I, T, i, guess_state(I, i, B.sub), &started);
  I.state[i] = make_uint4(e.byte, pack_state(e), started, 0u);
}

// D2: one synchronisation round.  Inside a block the subsequences iterate until their chain is consistent with
// the state the first one got from the previous block (as it was when the kernel started); the host relaunches
// until no block's last end state moves any more.  In every iteration the subsequences that have to be decoded
// again (their input state moved) are COMPACTED onto the first threads, so a few stragglers cost a few warps,
// not the whole block.
//
// Rounds are enqueued back to back without a host round trip: round r raises changed[r], and round r + 1 returns at
// once when changed[r] is still zero (the chains have converged: 2 rounds for photos), so a fixed number of launches
// costs a few empty grids; the host looks at the LAST flag when the job is waited for.
template <class BATCH>
__global__ void __launch_bounds__(JDEC_THREADS) jdec_sync_kernel(const __grid_constant__ BATCH B) {
  if (B.flag_prev >= 0 && B.changed[B.flag_prev] == 0u) return;   // grid-uniform: converged in an earlier round
  __shared__ SmemTables T;
  __shared__ uint2 sh_state[JDEC_THREADS + 1];   // [s + 1] = end state of subsequence s; [0] = predecessor block's last
  __shared__ uint32_t sh_cnt[JDEC_THREADS];      // blocks started by subsequence s
  __shared__ uint8_t sh_dirty[2][JDEC_THREADS + 1];   // [s]: the INPUT state of subsequence s moved
  __shared__ uint16_t sh_list[JDEC_THREADS];
  __shared__ int sh_warp_n[JDEC_THREADS / 32];
  __shared__ int sh_last_moved;
  const auto& I = B.item(blockIdx.y);
  if (blockIdx.x * JDEC_THREADS >= (unsigned)I.n_sub) return;
  build_tables(I.tab, T);
  const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
  const uint32_t base = blockIdx.x * JDEC_THREADS;
  const int n_here = min(JDEC_THREADS, I.n_sub - (int)base);
  {
    const uint4 mine = t < n_here ? I.state[base + t] : make_uint4(0, 0, 0, 0);
    sh_state[t + 1] = make_uint2(mine.x, mine.y);
    sh_cnt[t] = mine.z;
    // first round: everybody re-decodes from the predecessor's state; later rounds: only the chain that
    // starts at a first subsequence whose input (the previous block's last end state) has moved since it was used
    sh_dirty[0][t + 1] = (B.round == 0 && t + 1 < n_here) ? 1 : 0;
    sh_dirty[1][t + 1] = 0;
  }
  if (t == 0) {
    uint2 in = make_uint2(0u, scan_start_word(I));   // the scan starts at byte 0, bit 0, block 0, DC
    if (base > 0) {
      const uint4 p = I.state[base - 1];
      in = make_uint2(p.x, p.y);
    }
    sh_state[0] = in;
    const uint2 was = I.consumed[blockIdx.x];
    sh_dirty[0][0] = (B.round == 0 || was.x != in.x || was.y != in.y) ? 1 : 0;
    sh_dirty[1][0] = 0;
    I.consumed[blockIdx.x] = in;
    sh_last_moved = 0;
  }
  __syncthreads();
  int cur = 0;
  bool settled = false;
  for (int iter = 0; iter < JDEC_THREADS + 1; ++iter) {
    // compact the dirty subsequences: warp ballots + a prefix over the 8 warps
    const bool work = t < n_here && sh_dirty[cur][t] != 0;
    const uint32_t m = __ballot_sync(0xffffffffu, work);
    if (lane == 0) sh_warp_n[wp] = __popc(m);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int k = 0; k < JDEC_THREADS / 32; ++k) {
      const int c = sh_warp_n[k];
      if (k < wp) before += c;
      total += c;
    }
    if (total == 0) { settled = true; break; }
    if (work) sh_list[before + __popc(m & ((1u << lane) - 1u))] = (uint16_t)t;
    sh_dirty[cur ^ 1][t] = 0;
    if (t == 0) sh_dirty[cur ^ 1][JDEC_THREADS] = 0;
    __syncthreads();
    uint2 fresh = make_uint2(0u, 0u);
    uint32_t started = 0;
    int s = -1;
    if (t < total) {
      s = sh_list[t];
      const uint2 ps = sh_state[s];
      const State e = redecode<true>(B, I, T, base + s, unpack_state(ps.x, ps.y), &started);
      fresh = make_uint2(e.byte, pack_state(e));
    }
    __syncthreads();
    if (s >= 0) {
      const uint2 old = sh_state[s + 1];
      const bool moved = fresh.x != old.x || fresh.y != old.y;
      sh_state[s + 1] = fresh;
      sh_cnt[s] = started;
      if (moved) {
        sh_dirty[cur ^ 1][s + 1] = 1;
        if (s == n_here - 1) sh_last_moved = 1;
      }
    }
    __syncthreads();
    cur ^= 1;
  }
  if (t < n_here) {
    const uint2 st = sh_state[t + 1];
    I.state[base + t] = make_uint4(st.x, st.y, sh_cnt[t], 0u);
  }
  // the LAST subsequence's end state feeds the next block: if it moved, another round is needed
  if (t == 0 && (sh_last_moved || !settled) && base + n_here < (uint32_t)I.n_sub) atomicOr(&B.changed[B.flag_cur], 1u);
}

// D3: exclusive scan of the blocks-started counts (one block per image).
template <class BATCH>
__global__ void __launch_bounds__(1024) jdec_scan_kernel(const __grid_constant__ BATCH B) {
  __shared__ uint32_t warp_sum[32];
  __shared__ uint32_t carry_s;
  const auto& I = B.item(blockIdx.x);
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  if (t == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < I.n_sub; base += 1024) {
    const int i = base + t;
    const uint32_t v = i < I.n_sub ? I.state[i].z : 0u;
    uint32_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) warp_sum[w] = x;
    __syncthreads();
    if (w == 0) {
      uint32_t s = warp_sum[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, s, d);
        if (lane >= d) s += y;
      }
      warp_sum[lane] = s;
    }
    __syncthreads();
    const uint32_t before = carry_s + (w ? warp_sum[w - 1] : 0u) + x - v;
    if (i < I.n_sub) I.prefix[i] = before;
    __syncthreads();
    if (t == 1023) carry_s = before + v;
    __syncthreads();
  }
}

// D4: write pass from the synchronised states.
__global__ void __launch_bounds__(JDEC_THREADS) jdec_write_kernel(const __grid_constant__ DecBatch B) {
  __shared__ SmemTables T;
  __shared__ __align__(16) int16_t stage[JDEC_THREADS * 64];
  const DecImage& I = B.img[blockIdx.y];
  if (blockIdx.x * JDEC_THREADS >= (unsigned)I.n_sub) return;
  build_tables(I.tab, T);
  const uint32_t i = blockIdx.x * JDEC_THREADS + threadIdx.x;
  if (i >= (uint32_t)I.n_sub) return;
  State s;
  if (i == 0) { s.byte = 0; s.bit = 0; s.b = 0; s.z = 0; s.done = false; }
  else { const uint4 p = I.state[i - 1]; s = unpack_state(p.x, p.y); }
  uint32_t started = 0;
  decode_range<true>(I, T, s, (i + 1) * B.sub, &started, (int64_t)I.prefix[i], stage + threadIdx.x * 8);
}

// D5: DC differences -> DC values, in three steps: (a) per-MCU sums of the differences per component,
// (b) one block per image: segmented inclusive scan of those sums over the MCUs (a restart interval
// boundary -- MCU index multiple of ri -- zeroes the predictors), in place, (c) per MCU: predictor entering
// it = scanned value minus its own sum; rewrite the DC slots of its blocks.
__global__ void __launch_bounds__(256) jdec_dc_sum_kernel(const __grid_constant__ DecBatch B) {
  const DecImage& I = B.img[blockIdx.y];
  const int m = blockIdx.x * 256 + threadIdx.x, n_mcu = I.mcux * I.mcuy, ny = I.hs * I.vs;
  if (m >= n_mcu) return;
  const int16_t* blk = I.coef + (int64_t)m * I.bpm * 64;
  int s0 = 0;
  for (int b = 0; b < ny; ++b) s0 += blk[b * 64];
  I.dcsum[m] = s0;
  if (I.ncomp == 3) {
    I.dcsum[n_mcu + m] = blk[ny * 64];
    I.dcsum[2 * n_mcu + m] = blk[(ny + 1) * 64];
  }
}

__global__ void __launch_bounds__(1024) jdec_dc_scan_kernel(const __grid_constant__ DecBatch B) {
  __shared__ int wsum[32];
  __shared__ int wflag[32];
  __shared__ int carry;
  const DecImage& I = B.img[blockIdx.x];
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  const int n_mcu = I.mcux * I.mcuy;
  int* v = I.dcsum + (int64_t)blockIdx.y * n_mcu;     // blockIdx.y = component
  if ((int)blockIdx.y >= I.ncomp) return;
  if (t == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n_mcu; base += 1024) {
    const int m = base + t;
    const bool live = m < n_mcu;
    const int own = live ? v[m] : 0;
    int x = own, f = (live && I.ri > 0 && (m % I.ri) == 0) ? 1 : 0;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int f2 = __shfl_up_sync(0xffffffffu, f, d);
      const int y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) {
        if (!f) x += y;
        f |= f2;
      }
    }
    if (lane == 31) { wsum[w] = x; wflag[w] = f; }
    __syncthreads();
    if (w == 0) {
      int xs = wsum[lane], fs = wflag[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int f2 = __shfl_up_sync(0xffffffffu, fs, d);
        const int y = __shfl_up_sync(0xffffffffu, xs, d);
        if (lane >= d) {
          if (!fs) xs += y;
          fs |= f2;
        }
      }
      wsum[lane] = xs;
      wflag[lane] = fs;
    }
    __syncthreads();
    int r = x;                                          // segmented inclusive value
    if (!f) {                                           // no restart in my warp at or before me
      if (w > 0) r += wsum[w - 1];
      if (!(w > 0 && wflag[w - 1])) r += carry;         // nor in the earlier warps of this chunk
    }
    if (live) v[m] = r;
    __syncthreads();
    if (t == 1023) carry = r;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) jdec_dc_apply_kernel(const __grid_constant__ DecBatch B) {
  const DecImage& I = B.img[blockIdx.y];
  const int m = blockIdx.x * 256 + threadIdx.x, n_mcu = I.mcux * I.mcuy, ny = I.hs * I.vs;
  if (m >= n_mcu) return;
  int16_t* blk = I.coef + (int64_t)m * I.bpm * 64;
  int s0 = 0;
  for (int b = 0; b < ny; ++b) s0 += blk[b * 64];
  int p = I.dcsum[m] - s0;                              // luma predictor entering this MCU
  for (int b = 0; b < ny; ++b) { p += blk[b * 64]; blk[b * 64] = (int16_t)p; }
  if (I.ncomp == 3) {
    blk[ny * 64] = (int16_t)I.dcsum[n_mcu + m];
    blk[(ny + 1) * 64] = (int16_t)I.dcsum[2 * n_mcu + m];
  }
}

// ---- D6: dequantise + inverse DCT (LL&M, 13-bit constants; same arithmetic as libjpeg's default) ----
__device__ __forceinline__ void idct8(const int* in, int* out, int shift) {
  int z2 = in[2], z3 = in[6];
  int z1 = (z2 + z3) * 4433;
  int tmp2 = z1 + z3 * (-15137), tmp3 = z1 + z2 * 6270;
  z2 = in[0];
  z3 = in[4];
  int tmp0 = (z2 + z3) * 8192, tmp1 = (z2 - z3) * 8192;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  tmp0 = in[7];
  tmp1 = in[5];
  tmp2 = in[3];
  tmp3 = in[1];
  z1 = tmp0 + tmp3;
  z2 = tmp1 + tmp2;
  z3 = tmp0 + tmp2;
  int z4 = tmp1 + tmp3;
  const int z5 = (z3 + z4) * 9633;
  tmp0 *= 2446;
  tmp1 *= 16819;
  tmp2 *= 25172;
  tmp3 *= 12299;
  z1 *= -7373;
  z2 *= -20995;
  z3 *= -16069;
  z4 *= -3196;
  z3 += z5;
  z4 += z5;
  tmp0 += z1 + z3;
  tmp1 += z2 + z4;
  tmp2 += z2 + z3;
  tmp3 += z1 + z4;
  const int r = 1 << (shift - 1);
  out[0] = (tmp10 + tmp3 + r) >> shift;
  out[7] = (tmp10 - tmp3 + r) >> shift;
  out[1] = (tmp11 + tmp2 + r) >> shift;
  out[6] = (tmp11 - tmp2 + r) >> shift;
  out[2] = (tmp12 + tmp1 + r) >> shift;
  out[5] = (tmp12 - tmp1 + r) >> shift;
  out[3] = (tmp13 + tmp0 + r) >> shift;
  out[4] = (tmp13 - tmp0 + r) >> shift;
}

__global__ void __launch_bounds__(128) jdec_idct_kernel(const __grid_constant__ DecBatch B) {
  __shared__ uint16_t sq[4][64];     // zig-zag order
  const DecImage& I = B.img[blockIdx.y];
  if (blockIdx.x * 128 >= (unsigned)I.n_blocks) return;
  constexpr uint8_t ZZ[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                              41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                              30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
  for (int k = threadIdx.x; k < 256; k += 128) sq[k >> 6][k & 63] = I.tab->q[k >> 6][k & 63];   // zig-zag order, as stored
  __syncthreads();
  const int blk = blockIdx.x * 128 + threadIdx.x;
  if (blk >= I.n_blocks) return;
  const int m = blk / I.bpm, b = blk - m * I.bpm, ny = I.hs * I.vs;
  const int comp = b < ny ? 0 : b - ny + 1;
  const int mx = m % I.mcux, my = m / I.mcux;
  int bx, by;
  if (comp == 0) { bx = mx * I.hs + (b % I.hs); by = my * I.vs + (b / I.hs); }
  else { bx = mx; by = my; }
  const uint16_t* q = sq[I.tq[comp]];
  // load (zig-zag order, 128 contiguous bytes) and dequantise into natural order
  int v[64];
  const int4* src = reinterpret_cast<const int4*>(I.coef + (int64_t)blk * 64);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int4 a = src[j];
    const int w4[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k0 = j * 8 + e * 2;
      v[ZZ[k0]] = (int)(int16_t)(w4[e] & 0xffff) * (int)q[k0];
      v[ZZ[k0 + 1]] = (int)(int16_t)(w4[e] >> 16) * (int)q[k0 + 1];
    }
  }
  // columns, then rows
  int ws[64];
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    int in[8], o[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) in[r] = v[r * 8 + c];
    idct8(in, o, 11);
#pragma unroll
    for (int r = 0; r < 8; ++r) ws[r * 8 + c] = o[r];
  }
  uint8_t* dst = I.plane[comp] + (int64_t)by * 8 * I.pw[comp] + bx * 8;
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    int o[8];
    idct8(&ws[r * 8], o, 18);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      lo |= (uint32_t)min(max(o[c] + 128, 0), 255) << (8 * c);
      hi |= (uint32_t)min(max(o[c + 4] + 128, 0), 255) << (8 * c);
    }
    *reinterpret_cast<uint2*>(dst + (int64_t)r * I.pw[comp]) = make_uint2(lo, hi);
  }
}

// ---- D7: chroma upsampling (triangle filter, as libjpeg's "fancy" upsampling) + YCbCr -> RGB ----
// One thread = 4 horizontally adjacent pixels (x0 = 4k): one 32-bit luma load, the chroma samples i-1 .. i+2
// of the two rows involved, one 16-byte store.
__device__ __forceinline__ uint32_t ycc_to_rgba(int yy, int cb, int cr) {
  const int R = min(max(yy + ((91881 * cr + 32768) >> 16), 0), 255);
  const int G = min(max(yy + ((-22554 * cb + 32768 - 46802 * cr) >> 16), 0), 255);
  const int Bc = min(max(yy + ((116130 * cb + 32768) >> 16), 0), 255);
  return (uint32_t)R | (uint32_t)G << 8 | (uint32_t)Bc << 16 | 0xff000000u;
}

__global__ void __launch_bounds__(256) jdec_colour_kernel(const __grid_constant__ DecBatch B) {
  const DecImage& I = B.img[blockIdx.z];
  const int x0 = (blockIdx.x * 32 + (threadIdx.x & 31)) * 4;
  const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x0 >= I.w || y >= I.h) return;
  const uint32_t y4 = *reinterpret_cast<const uint32_t*>(I.plane[0] + (int64_t)y * I.pw[0] + x0);   // planes are padded to 8
  uint32_t px[4];
  if (I.ncomp == 1) {
#pragma unroll
    for (int k = 0; k < 4; ++k) px[k] = ((y4 >> (8 * k)) & 255u) * 0x010101u | 0xff000000u;
  } else {
    const int cw = (I.w + I.hs - 1) / I.hs, chh = (I.h + I.vs - 1) / I.vs;
    const bool fancy = I.hs == 2 && cw > 2;
    int cv[2][4];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const uint8_t* P = I.plane[c + 1];
      const int stride = I.pw[c + 1];
      if (I.hs == 1) {
        const uint32_t v4 = *reinterpret_cast<const uint32_t*>(P + (int64_t)y * stride + x0);
#pragma unroll
        for (int k = 0; k < 4; ++k) cv[c][k] = (int)((v4 >> (8 * k)) & 255u);
      } else {
        const int i0 = x0 >> 1;                         // chroma samples i0, i0 + 1 cover the 4 pixels
        const int r = I.vs == 2 ? (y >> 1) : y;
        const uint8_t* r0 = P + (int64_t)r * stride;
        // horizontal neighbours i0 - 1 .. i0 + 2, clamped into the plane (the formulas below never use a clamped one)
        const int il = max(i0 - 1, 0), ir = min(i0 + 2, stride - 1);
        int s[4];                                       // column sums: 3 * near row + far row (h2v2) or the sample (h2v1)
        if (I.vs == 2 && fancy) {
          const int rn = min(max((y & 1) ? r + 1 : r - 1, 0), chh - 1);
          const uint8_t* r1 = P + (int64_t)rn * stride;
          s[0] = r0[il] * 3 + r1[il];
          s[1] = r0[i0] * 3 + r1[i0];
          s[2] = r0[i0 + 1] * 3 + r1[i0 + 1];
          s[3] = r0[ir] * 3 + r1[ir];
        } else {
          s[0] = r0[il];
          s[1] = r0[i0];
          s[2] = r0[i0 + 1];
          s[3] = r0[ir];
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {                   // chroma sample i = i0 + j -> pixels 2j, 2j + 1
          const int i = i0 + j, cur = s[1 + j], left = s[j], right = s[2 + j];
          int even, odd;
          if (!fancy) {
            even = odd = cur;
          } else if (I.vs == 2) {
            even = i == 0 ? (cur * 4 + 8) >> 4 : (cur * 3 + left + 8) >> 4;
            odd = i == cw - 1 ? (cur * 4 + 7) >> 4 : (cur * 3 + right + 7) >> 4;
          } else {
            even = i == 0 ? cur : (cur * 3 + left + 1) >> 2;
            odd = i == cw - 1 ? cur : (cur * 3 + right + 2) >> 2;
          }
          cv[c][2 * j] = even;
          cv[c][2 * j + 1] = odd;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) px[k] = ycc_to_rgba((int)((y4 >> (8 * k)) & 255u), cv[0][k] - 128, cv[1][k] - 128);
  }
  uint8_t* o = I.rgba + (int64_t)y * I.pitch + (int64_t)x0 * 4;
  if (x0 + 3 < I.w && (((uintptr_t)o) & 15) == 0) {
    *reinterpret_cast<uint4*>(o) = make_uint4(px[0], px[1], px[2], px[3]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (x0 + k < I.w) reinterpret_cast<uint32_t*>(o)[k] = px[k];
  }
}

}  // namespace jpegdec
}  // namespace stitch
