// Progressive JPEG (SOF2) decoding on the GPU: the entropy kernels.  The back end (dequantise + IDCT, chroma
// upsampling + colour) is the baseline decoder's (jpeg_decode.cuh D6, D7) over the same coefficient buffer.
//
// A progressive file is a list of scans, each coding one spectral band of one component (or the DC terms of all of
// them) at one precision (T.81 Annex G).  Which of them can run side by side follows from what they touch:
//   * FIRST scans (Ah = 0) of a band touch coefficients nothing else has written, and their symbols do not depend on
//     the coefficients: all first scans of all files of a job go through the self-synchronising decoder of the
//     baseline path in ONE set of launches -- a scan is a batch item like a baseline file is.  State per subsequence:
//     (bit position, block-in-MCU) for DC scans, (bit position, zig-zag index) for AC scans; an EOB run of n blocks is
//     one symbol that completes n blocks, so the block index of a subsequence is again a prefix sum (of blocks
//     COMPLETED).  DC differences are written in scan order and turned into values by a segmented scan per component.
//   * REFINEMENT scans (Ah > 0) read one correction bit per coefficient that is ALREADY non-zero in the block they are
//     in, so a thread that guesses its start cannot even parse: they are decoded by one thread per scan
//     (pscan_sequential, jpeg_prog_core.h).  Scans that touch disjoint (component, band) sets share a launch: the host
//     orders them in dependency levels (Y / Cb / Cr refinements of the usual script run side by side).
#pragma once
#include "jpeg_decode.cuh"
#include "jpeg_prog_core.h"

namespace stitch {
namespace jpegdec {

// A scan's bytes inside the uploaded file -> its own 4-byte aligned buffer (the word-wise readers want that).  One
// upload per file and one launch instead of a copy call per scan: the host side of a 9-file job issued ~90 of them.
struct ScanPiece {
  const uint8_t* src;
  uint8_t* dst;
  uint32_t len;
};
__global__ void __launch_bounds__(256) jprog_gather_kernel(const ScanPiece* __restrict__ pieces) {
  const ScanPiece pc = pieces[blockIdx.x];
  for (uint32_t k = blockIdx.y * 256u + threadIdx.x; k < pc.len; k += gridDim.y * 256u) pc.dst[k] = pc.src[k];
}

struct PBatch {
  const PScan* scans;         // device array: the first scans of every file of the job
  uint32_t* changed;
  int32_t round, flag_cur, flag_prev;
  uint32_t sub;
  __device__ __forceinline__ const PScan& item(int i) const { return scans[i]; }
};

// Decodes the symbols of a FIRST scan whose first bit lies in raw bytes [.., end_byte), starting from `s`
// (s.b = block in the MCU for DC scans, s.z = zig-zag index for AC scans).  Returns the state after the last of them
// and the number of blocks COMPLETED.  WRITE: blk is the scan-order index of the block in progress.
template <bool WRITE>
__device__ __forceinline__ State pdecode_range(const PScan& S, const SmemTables& T, State s, uint32_t end_byte,
                                               uint32_t* completed, int64_t blk) {
  Reader r;
  r.p = S.data;
  r.n = S.len;
  rd_start(r, s.byte, s.bit);
  uint32_t count = 0;
  bool first = true;
  int b = s.b, z = s.z;
  State out = s;
  if (s.done) { *completed = 0; return s; }
  const bool dc = S.mode == PSCAN_DC_FIRST;
  const int ny = S.ns > 1 ? S.hs * S.vs : 1;
  const int z0 = dc ? 0 : S.ss;
  const int tac = 2 + S.ta;
  for (;;) {
    rd_fill(r);
    if (first) { rd_skip(r, s.bit); first = false; }
    {
      const int real = r.cnt - r.fake;
      const uint32_t nb = (uint32_t)((real + 7) >> 3);
      if (r.pos - nb >= end_byte) {
        uint32_t c;
        int cbit;
        rd_where(r, &c, &cbit);
        if (c >= end_byte) {
          out.byte = c; out.bit = cbit; out.b = b; out.z = z; out.done = false;
          break;
        }
      }
    }
    int nblk = 0, zpos = -1, value = 0, nb2 = b, nz = z;
    if (dc) {
      const int ci = S.ns == 1 ? 0 : (b < ny ? 0 : b - ny + 1);
      const int sz = huff_decode(T, S.td[ci], r) & 15;
      if (sz) {
        const uint32_t bits = (uint32_t)(r.acc >> (64 - sz));
        rd_skip(r, sz);
        value = (bits < (1u << (sz - 1))) ? (int)bits - (1 << sz) + 1 : (int)bits;
      }
      nblk = 1;
      nb2 = (b + 1 == S.bpm) ? 0 : b + 1;
    } else {
      const int rs = huff_decode(T, tac, r);
      const int run = rs >> 4, sz = rs & 15;
      if (sz) {
        const uint32_t bits = (uint32_t)(r.acc >> (64 - sz));
        rd_skip(r, sz);
        value = (bits < (1u << (sz - 1))) ? (int)bits - (1 << sz) + 1 : (int)bits;
        zpos = z + run;
        nz = zpos + 1;
      } else if (run == 15) {
        nz = z + 16;
      } else {                                   // EOBn: this block and (1 << run) + bits - 1 more end here
        nblk = 1 << run;
        if (run) {
          nblk += (int)(uint32_t)(r.acc >> (64 - run));
          rd_skip(r, run);
        }
        nz = S.ss;
      }
      if (nz > S.se && nblk == 0) { nblk = 1; nz = S.ss; }
    }
    if (r.cnt < r.fake) {
      // the symbol ran into a marker (it was padding) or past the end: drop it
      uint32_t mpos = r.pos;
      while (mpos + 2 < r.n && r.p[mpos + 1] == 0xFFu) ++mpos;
      const uint32_t m = (mpos + 1 < r.n) ? r.p[mpos + 1] : 0xD9u;
      if (mpos + 1 < r.n && m >= 0xD0u && m <= 0xD7u) {
        b = 0;
        z = z0;
        rd_start(r, mpos + 2, 0);
        if (mpos + 2 >= end_byte) {
          out.byte = mpos + 2; out.bit = 0; out.b = 0; out.z = z0; out.done = false;
          break;
        }
        continue;
      }
      out.byte = r.n; out.bit = 0; out.b = 0; out.z = z0; out.done = true;
      break;
    }
    if (WRITE && blk >= 0 && blk < S.n_blk) {
      if (dc) S.dcdiff[blk] = value;
      else if (zpos >= 0 && zpos < 64) S.coef[pscan_block(S, (int)blk) * 64 + zpos] = (int16_t)(value * (1 << S.al));
    }
    blk += nblk;
    count += (uint32_t)nblk;
    b = nb2;
    z = nz;
  }
  *completed = count;
  return out;
}

__device__ __forceinline__ State decode_nowrite(const PScan& S, const SmemTables& T, State s, uint32_t end_byte, uint32_t* completed) {
  return pdecode_range<false>(S, T, s, end_byte, completed, 0);
}
template <bool EARLY, class BATCH>
__device__ __forceinline__ State redecode(const BATCH& B, const PScan& S, const SmemTables& T, uint32_t idx, State in, uint32_t* completed) {
  return pdecode_range<false>(S, T, in, (idx + 1) * B.sub, completed, 0);   // no checkpoints for the scans of progressive files
}
__device__ __forceinline__ uint32_t scan_start_word(const PScan& S) {
  return S.mode == PSCAN_AC_FIRST ? (uint32_t)S.ss << 16 : 0u;
}

__device__ __forceinline__ State pguess_state(const PScan& S, uint32_t i, uint32_t sub) {
  State s;
  s.byte = i * sub;
  s.bit = 0;
  s.b = 0;
  s.z = S.mode == PSCAN_AC_FIRST ? S.ss : 0;
  s.done = false;
  if (i > 0 && s.byte < S.len && S.data[s.byte - 1] == 0xFFu) {
    const uint32_t v = S.data[s.byte];
    if (v == 0 || (v >= 0xD0u && v <= 0xD7u)) s.byte += 1;
  }
  return s;
}

// PD1: first pass from guessed states (the synchronisation rounds and the prefix sums are jdec_sync_kernel<PBatch> and
// jdec_scan_kernel<PBatch>).
__global__ void __launch_bounds__(JDEC_THREADS) jprog_first_kernel(const __grid_constant__ PBatch B) {
  __shared__ SmemTables T;
  const PScan& S = B.item(blockIdx.y);
  if (blockIdx.x * JDEC_THREADS >= (unsigned)S.n_sub) return;
  build_tables(S.tab, T);
  const uint32_t i = blockIdx.x * JDEC_THREADS + threadIdx.x;
  if (i >= (uint32_t)S.n_sub) return;
  uint32_t completed = 0;
  const State e = pdecode_range<false>(S, T, pguess_state(S, i, B.sub), (i + 1) * B.sub, &completed, 0);
  S.state[i] = make_uint4(e.byte, pack_state(e), completed, 0u);
}

// PD4: write pass from the synchronised states.
__global__ void __launch_bounds__(JDEC_THREADS) jprog_write_kernel(const __grid_constant__ PBatch B) {
  __shared__ SmemTables T;
  const PScan& S = B.item(blockIdx.y);
  if (blockIdx.x * JDEC_THREADS >= (unsigned)S.n_sub) return;
  build_tables(S.tab, T);
  const uint32_t i = blockIdx.x * JDEC_THREADS + threadIdx.x;
  if (i >= (uint32_t)S.n_sub) return;
  State s;
  if (i == 0) { s = unpack_state(0u, scan_start_word(S)); }
  else { const uint4 p = S.state[i - 1]; s = unpack_state(p.x, p.y); }
  uint32_t completed = 0;
  pdecode_range<true>(S, T, s, (i + 1) * B.sub, &completed, (int64_t)S.prefix[i]);
}

// PD5: DC differences (scan order) -> values << Al in the coefficient buffer.  One CTA per (scan, component in the
// scan): segmented inclusive scan over the component's blocks in scan order, a restart interval boundary zeroes the
// predictor.  grid = (first scans, 3); scans that are not DC scans return.
__global__ void __launch_bounds__(1024) jprog_dc_kernel(const PScan* __restrict__ scans) {
  __shared__ int wsum[32];
  __shared__ int wflag[32];
  __shared__ int carry;
  const PScan& S = scans[blockIdx.x];
  if (S.mode != PSCAN_DC_FIRST || (int)blockIdx.y >= S.ns) return;
  const int ci = blockIdx.y;
  const int ny = S.ns > 1 ? S.hs * S.vs : 1;
  const int nb = ci == 0 ? ny : 1, off = ci == 0 ? 0 : ny + ci - 1;
  const int n_el = (S.n_blk / S.bpm) * nb;
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  if (t == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n_el; base += 1024) {
    const int j = base + t;
    const bool live = j < n_el;
    const int mcu = j / nb, within = j - mcu * nb;
    const int blk = mcu * S.bpm + off + within;
    const int own = live ? S.dcdiff[blk] : 0;
    int x = own, f = (live && S.ri > 0 && within == 0 && (mcu % S.ri) == 0) ? 1 : 0;
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
    int r = x;
    if (!f) {
      if (w > 0) r += wsum[w - 1];
      if (!(w > 0 && wflag[w - 1])) r += carry;
    }
    if (live) S.coef[pscan_block(S, blk) * 64] = (int16_t)(r * (1 << S.al));
    __syncthreads();
    if (t == 1023) carry = r;
    __syncthreads();
  }
}

// ---- refinement scans: one WARP per scan ----
// The bit stream has to be parsed by one thread (see the head of this file); a single GPU thread issues a dependent
// instruction every ~4 cycles, so everything that is not parsing is taken off it:
//   * bytes: the warp copies the scan's bytes into a shared-memory ring WITHOUT the stuffed zeros (32 words per step,
//     keep flags + a warp prefix sum), so lane 0 refills 32 bits with one aligned LDS and no 0xFF test;
//   * blocks: they stream through a ring of JPROG_STAGES shared-memory slots (cp.async, one coalesced 128-byte request
//     per block, issued JPROG_STAGES blocks ahead: the L2 latency of the coefficient buffer -- what made a
//     coefficient-by-coefficient loop over global memory take 2.3 s for a 12 MP file -- is hidden), and two ballots
//     give lane 0 the block's non-zero mask;
//   * the lanes list the positions of the block's zero-history coefficients (one popc each), so that a zero run is one
//     table read for the parser, pj_refine_parse_z: results as four 64-bit words.  All 32 lanes run the parser on the
//     same bits (same cost as one lane in SIMT; no divergence, no broadcast of the results);
//   * every lane applies them to its two coefficients (pj_refine_apply) and stores the ones that changed.
// DC refinement scans (one bit per block) need no parser at all: lane i of a step takes block 32 * step + i.
// Scans with restart intervals, and first scans that could not join the parallel batch, take the plain one-thread
// route (pscan_sequential / pj_refine on the word-wise Reader).
#define JPROG_STAGES 8
#define JPROG_RING 1024            // bytes of unstuffed scan data in shared memory (power of two)

__device__ __forceinline__ int br_sym(const SmemTables& T, int t, Reader& r) {
  rd_fill(r);
  return huff_decode(T, t, r);
}
__device__ __forceinline__ int br_bits(Reader& r, int nbits) {   // nbits <= 32
  if (nbits == 0) return 0;
  rd_fill(r);
  const int v = (int)(uint32_t)(r.acc >> (64 - nbits));
  rd_skip(r, nbits);
  return v;
}
// Byte-aligned restart: drop the bits in hand, continue behind the next RSTn marker (zeros when there is none).
__device__ __forceinline__ void br_restart(Reader& r, int* rst_next) {
  uint32_t o = r.pos;
  while (o + 1 < r.n && !(r.p[o] == 0xFFu && r.p[o + 1] >= 0xD0u && r.p[o + 1] <= 0xD7u)) ++o;
  rd_start(r, o + 1 < r.n ? o + 2 : r.n, 0);
  *rst_next = (*rst_next + 1) & 7;
}

// Lane 0's reader over the ring of unstuffed bytes: rd / wr count bytes since the start of the scan.
struct RingBits {
  const uint8_t* ring;
  uint32_t rd, wr;
  uint64_t acc;
  int cnt;
};
__device__ __forceinline__ void rb_fill(RingBits& b) {
  if (b.cnt > 32) return;
  uint32_t w;
  if (b.rd + 4 <= b.wr) {
    w = __byte_perm(*reinterpret_cast<const uint32_t*>(b.ring + (b.rd & (JPROG_RING - 1))), 0, 0x0123);
  } else {                                        // the last bytes of the scan, zeros behind them
    w = 0;
    for (uint32_t j = 0; j < 4; ++j)
      if (b.rd + j < b.wr) w |= (uint32_t)b.ring[(b.rd + j) & (JPROG_RING - 1)] << (24 - 8 * j);
  }
  b.rd += 4;
  b.acc |= (uint64_t)w << (32 - b.cnt);
  b.cnt += 32;
}
__device__ __forceinline__ int br_sym(const SmemTables& T, int t, RingBits& b) {
  rb_fill(b);
  const uint32_t v = (uint32_t)(b.acc >> 48);
  const uint32_t e = T.lut[t][v >> 8];
  int len = (int)(e >> 8), sym = (int)(e & 255u);
  if (!e) {
    len = 16;
    sym = 0;
    for (int l = 9; l <= 16; ++l) {
      const int code = (int)(v >> (16 - l));
      if (code <= T.maxcode[t][l] && code >= T.mincode[t][l]) {
        len = l;
        sym = T.vals[t][(T.valptr[t][l] + code - T.mincode[t][l]) & 255];
        break;
      }
    }
  }
  b.acc <<= len;
  b.cnt -= len;
  return sym;
}
__device__ __forceinline__ int br_bits(RingBits& b, int nbits) {   // nbits <= 32
  if (nbits == 0) return 0;
  rb_fill(b);
  const int v = (int)(uint32_t)(b.acc >> (64 - nbits));
  b.acc <<= nbits;
  b.cnt -= nbits;
  return v;
}

// ---- scans of one launch that depend on each other: progress words ----
// A refinement pass may touch block i as soon as the pass before it (same component, same block order) is done with
// block i: the producer publishes its block count every few blocks (fence + release store), the consumer polls it
// (acquire load, bounded: a corrupt job must not hang the GPU) before it requests a block.  All CTAs of the launch are
// resident (one warp each, at most a few dozen), and producers have the lower CTA indices.
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Warp-wide: returns once every dependency has published at least `need` blocks; `seen` caches the smallest count read.
// Every lane runs the same loop on lane 0's reading: a poll inside `if (lane == 0)` left the warp split in two behind
// it (the compiler kept separate copies of the code that follows), and everything warp-wide then ran twice per block
// -- 340 ms instead of 230 for a 3 MP file.
__device__ __forceinline__ void wait_deps(const PScan& S, uint32_t need, uint32_t& seen, int lane) {
  if (S.ndep == 0 || seen >= need) return;
  uint32_t low = 0xffffffffu;
  for (int d = 0; d < S.ndep; ++d) {
    uint32_t v = __shfl_sync(0xffffffffu, ld_acquire_u32(S.dep[d]), 0), spins = 0;
    while (v < need && ++spins < (1u << 24)) {
      __nanosleep(200);
      v = __shfl_sync(0xffffffffu, ld_acquire_u32(S.dep[d]), 0);
    }
    low = min(low, max(v, need));                 // gave up: carry on (the job's output is garbage, the GPU is free)
  }
  seen = low;
  (void)lane;
}
// Warp-wide: this scan's blocks [0, n) are final.
__device__ __forceinline__ void publish(const PScan& S, uint32_t n, int lane) {
  if (!S.progress) return;
  __syncwarp();                                   // the lanes' stores happen before lane 0's release (cumulative)
  if (lane == 0) st_release_u32(S.progress, n);
#ifdef JPROG_TRACE
  if (lane == 0 && (n <= 16 || (n & 8191) == 0)) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    printf("  publish %u at %llu us (word %p)\n", n, t / 1000 % 100000000ull, (void*)S.progress); }
#endif
}

// The plain route: one thread, coefficient blocks through the cp.async ring, word-wise Reader with restart markers.
__device__ __forceinline__ void jprog_refine_plain(const PScan& S, const SmemTables& T, int16_t (*ring)[64], int lane) {
  uint32_t seen = 0;
  wait_deps(S, (uint32_t)S.n_blk, seen, lane);
  auto issue = [&](int blk) {
    if (blk < S.n_blk) {
      const int16_t* src = S.coef + pscan_block(S, blk) * 64 + 2 * lane;
      const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&ring[blk % JPROG_STAGES][2 * lane]);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  for (int s = 0; s < JPROG_STAGES; ++s) issue(s);
  Reader r;
  r.p = S.data;
  r.n = S.len;
  rd_start(r, 0, 0);
  int rst_left = S.ri, rst_next = 0, b_in_mcu = 0;
  uint32_t eobrun = 0;
  const int p1 = 1 << S.al, tac = 2 + S.ta;
  const uint64_t band = (S.se >= 63 ? ~0ull : ((1ull << (S.se + 1)) - 1)) & ~((1ull << S.ss) - 1);
  for (int blk = 0; blk < S.n_blk; ++blk) {
    asm volatile("cp.async.wait_group %0;" ::"n"(JPROG_STAGES - 1) : "memory");
    __syncwarp();
    int16_t* c = ring[blk % JPROG_STAGES];
    const int v0 = c[lane], v1 = c[lane + 32];
    const uint32_t lo = __ballot_sync(0xffffffffu, v0 != 0), hi = __ballot_sync(0xffffffffu, v1 != 0);
    RefineOut o;
    o.corrmask = o.corrbits = o.newmask = o.negmask = 0;
    if (lane == 0) {
      if (b_in_mcu == 0 && S.ri) {
        if (rst_left == 0) { br_restart(r, &rst_next); eobrun = 0; rst_left = S.ri; }
        rst_left--;
      }
      if (S.mode == PSCAN_DC_REFINE) {
        if (br_bits(r, 1)) { o.corrmask = 1; o.corrbits = 1; }     // "correct" coefficient 0 ...
      } else {
        pj_refine_parse(r, T, tac, (uint64_t)lo | (uint64_t)hi << 32, band, &eobrun, &o);
      }
      b_in_mcu = (b_in_mcu + 1 == S.bpm) ? 0 : b_in_mcu + 1;
    }
    o.corrmask = __shfl_sync(0xffffffffu, o.corrmask, 0);
    o.corrbits = __shfl_sync(0xffffffffu, o.corrbits, 0);
    o.newmask = __shfl_sync(0xffffffffu, o.newmask, 0);
    o.negmask = __shfl_sync(0xffffffffu, o.negmask, 0);
    int16_t* dst = S.coef + pscan_block(S, blk) * 64;
    if (S.mode == PSCAN_DC_REFINE) {
      if (lane == 0 && o.corrmask) dst[0] = (int16_t)(v0 | p1);     // ... which for the DC term is a plain OR
    } else {
      const int n0 = pj_refine_apply(v0, lane, p1, o), n1 = pj_refine_apply(v1, lane + 32, p1, o);
      if (n0 != v0) dst[lane] = (int16_t)n0;
      if (n1 != v1) dst[lane + 32] = (int16_t)n1;
    }
    __syncwarp();
    issue(blk + JPROG_STAGES);
  }
  publish(S, (uint32_t)S.n_blk, lane);
}

// Scans decoded by one warp each: one CTA per scan.
__global__ void __launch_bounds__(32) jprog_seq_kernel(const PScan* __restrict__ scans) {
  __shared__ SmemTables T;
  __shared__ __align__(16) int16_t ring[JPROG_STAGES][64];
  __shared__ __align__(16) uint8_t bytes[JPROG_RING];
  __shared__ uint8_t zpos[64];             // positions of the block's zero-history coefficients inside the band
  const PScan S = scans[blockIdx.x];      // a private copy: the decoder's stores cannot alias it
  build_tables(S.tab, T);
  const int lane = threadIdx.x;
#ifdef JPROG_TRACE
  struct Trace { int cta, mode, comp, al, ndep; unsigned smid; unsigned long long t0;
    __device__ ~Trace() { unsigned long long t1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (threadIdx.x == 0) printf("cta %d mode %d comp %d al %d ndep %d sm %u: start %llu us, end %llu us\n", cta, mode, comp, al, ndep, smid, t0 / 1000 % 100000000ull, t1 / 1000 % 100000000ull); } } trace;
  trace.cta = blockIdx.x; trace.mode = S.mode; trace.comp = S.comp; trace.al = S.al; trace.ndep = S.ndep;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(trace.smid));
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(trace.t0));
#endif
  if (S.mode != PSCAN_AC_REFINE && S.mode != PSCAN_DC_REFINE) {
    // a first scan that could not join the parallel batch (a file that codes a band twice; STITCH_JPROG_SEQ)
    uint32_t seen = 0;
    wait_deps(S, (uint32_t)S.n_blk, seen, lane);
    if (lane == 0) pscan_sequential(S, T);
    __syncwarp();
    publish(S, (uint32_t)S.n_blk, lane);
    return;
  }
  if (S.ri != 0) { jprog_refine_plain(S, T, ring, lane); return; }

  // ---- the byte ring: 32 words of the scan per step, stuffed zeros dropped ----
  uint32_t raw_pos = 0, wr = 0, prev_last = 0;    // warp-uniform
  const uint32_t* data32 = reinterpret_cast<const uint32_t*>(S.data);
  auto stage = [&]() {
    const uint32_t base = raw_pos + 4 * lane;
    const uint32_t w = base < S.len ? data32[base >> 2] : 0u;
    const int valid = base < S.len ? min(4, (int)(S.len - base)) : 0;
    const uint32_t up = __shfl_up_sync(0xffffffffu, w, 1);
    uint32_t pb = lane ? up >> 24 : prev_last;
    uint32_t keep = 0;                            // bit j: byte j of the word is data
    int n = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t bj = (w >> (8 * j)) & 255u;
      if (j < valid && !(bj == 0u && pb == 0xFFu)) { keep |= 1u << j; ++n; }
      pb = bj;
    }
    int x = n;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    uint32_t at = wr + (uint32_t)(x - n);
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (keep & (1u << j)) bytes[at++ & (JPROG_RING - 1)] = (uint8_t)(w >> (8 * j));
    wr += (uint32_t)__shfl_sync(0xffffffffu, x, 31);
    prev_last = __shfl_sync(0xffffffffu, w, 31) >> 24;
    raw_pos += 128;
  };
  // at least `need` bytes ahead of the reader (or all that is left); never overwrites unread bytes (need <= 512)
  auto ensure = [&](uint32_t rd, uint32_t need) {
    while (wr - rd < need && raw_pos < S.len) stage();
    __syncwarp();
  };
  const int p1 = 1 << S.al;

  if (S.mode == PSCAN_DC_REFINE) {
    // one bit per block, in block order: a step is 32 blocks = 4 bytes of the ring
    uint32_t rd = 0, seen = 0;
    for (int g = 0; g < S.n_blk; g += 32) {
      wait_deps(S, (uint32_t)min(S.n_blk, g + 32), seen, lane);
      ensure(rd, 64);
      const int blk = g + lane;
      const uint32_t at = rd + (lane >> 3);
      const uint32_t byte = at < wr ? bytes[at & (JPROG_RING - 1)] : 0u;
      if (blk < S.n_blk && ((byte >> (7 - (lane & 7))) & 1u)) {
        int16_t* dc = S.coef + pscan_block(S, blk) * 64;
        *dc = (int16_t)(*dc | p1);
      }
      rd += 4;
      __syncwarp();
    }
    publish(S, (uint32_t)S.n_blk, lane);
    return;
  }

  uint32_t seen = 0;
  BlockWalk ahead, cur;                    // the block being requested / the block being decoded
  walk_start(S, 0, &ahead);
  walk_start(S, 0, &cur);
  auto issue = [&](int blk) {               // blocks are requested in order: `ahead` is at blk
    if (blk < S.n_blk) {
      wait_deps(S, (uint32_t)blk + 1u, seen, lane);
      // L2 loads (cp.async.cg, 16 bytes from 8 lanes): another SM may have written the block a moment ago
      if (lane < 8) {
        const int16_t* src = S.coef + walk_index(S, ahead, blk) * 64 + 8 * lane;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&ring[blk % JPROG_STAGES][8 * lane]);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
      }
      walk_next(S, &ahead);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  for (int s = 0; s < JPROG_STAGES; ++s) issue(s);
  RingBits b;
  b.ring = bytes;
  b.rd = 0;
  b.wr = 0;
  b.acc = 0;
  b.cnt = 0;
  uint32_t eobrun = 0;
  const int tac = 2 + S.ta;
  const uint64_t band = (S.se >= 63 ? ~0ull : ((1ull << (S.se + 1)) - 1)) & ~((1ull << S.ss) - 1);
  for (int blk = 0; blk < S.n_blk; ++blk) {
    // a block takes at most 63 x (16 + 1 + 1) + 30 bits = 146 bytes, the reader holds up to 8 more
    ensure(b.rd, 256);
    asm volatile("cp.async.wait_group %0;" ::"n"(JPROG_STAGES - 1) : "memory");
    __syncwarp();
    const int16_t* c = ring[blk % JPROG_STAGES];
    const int v0 = c[lane], v1 = c[lane + 32];
    const uint32_t lo = __ballot_sync(0xffffffffu, v0 != 0), hi = __ballot_sync(0xffffffffu, v1 != 0);
    // the list of zero-history positions: every lane ranks its two coefficients with one popc each
    const uint32_t zlo = ~lo & (uint32_t)band, zhi = ~hi & (uint32_t)(band >> 32);
    const int nzl = __popc(zlo);
    if ((zlo >> lane) & 1u) zpos[__popc(zlo & ((1u << lane) - 1u))] = (uint8_t)lane;
    if ((zhi >> lane) & 1u) zpos[nzl + __popc(zhi & ((1u << lane) - 1u))] = (uint8_t)(32 + lane);
    __syncwarp();
    // EVERY lane parses (the same bits, the same tables: shared-memory broadcasts): in SIMT that costs what one lane
    // costs, the warp never splits, and every lane ends up holding the results
    RefineOut o;
    b.wr = wr;
    pj_refine_parse_z(b, T, tac, (uint64_t)lo | (uint64_t)hi << 32, band, zpos, nzl + __popc(zhi), S.ss, S.se, &eobrun, &o);
    if (o.corrmask | o.newmask) {
      int16_t* dst = S.coef + walk_index(S, cur, blk) * 64;
      const int n0 = pj_refine_apply(v0, lane, p1, o), n1 = pj_refine_apply(v1, lane + 32, p1, o);
      if (n0 != v0) dst[lane] = (int16_t)n0;
      if (n1 != v1) dst[lane + 32] = (int16_t)n1;
    }
    walk_next(S, &cur);
    if ((blk & 31) == 31 || blk == S.n_blk - 1) publish(S, (uint32_t)blk + 1u, lane);
    __syncwarp();
    issue(blk + JPROG_STAGES);
  }
}

}  // namespace jpegdec
}  // namespace stitch
