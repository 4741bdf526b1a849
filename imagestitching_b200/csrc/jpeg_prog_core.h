// Progressive JPEG (SOF2, T.81 Annex G) on the GPU: what a scan looks like to the kernels, and the SEQUENTIAL decoder
// of one scan -- the code a single GPU thread runs for a refinement scan (jpeg_prog.cuh), written so that it also
// compiles as plain C++: tests/native/jprog_host.cpp drives it on the CPU against the oracle decoder.
//
// Replaces, for files the page's SUPPORTED_IMAGE_TYPES (P:4) admits, the decode inside loadImageFrom (U:27-121).
//
// Coefficients live in the baseline decoder's buffer (jpeg_decode.cuh DecImage::coef): blocks in the order of a fully
// interleaved scan (MCU by MCU: hs * vs luma blocks, then Cb, Cr), zig-zag order inside a block, so the dequantise /
// IDCT / colour kernels are shared.  A single-component scan walks the component's own blocks in raster order over
// ceil(ceil(w * hs_c / hmax) / 8) columns (T.81 A.2.3): pscan_block maps its block index into that buffer.
#pragma once
#include <cstdint>

#include "jpeg_parse.h"

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define PJ_HD __device__ __forceinline__   /* under nvcc only kernels call these; the host build is the tests' CPU drive */
#else
#define PJ_HD inline
struct uint4 { uint32_t x, y, z, w; };   // plain-C++ build (the CPU drive of the tests): layout only
struct uint2 { uint32_t x, y; };
#endif

namespace stitch {
namespace jpegdec {

enum { PSCAN_DC_FIRST = 0, PSCAN_AC_FIRST = 1, PSCAN_DC_REFINE = 2, PSCAN_AC_REFINE = 3 };

struct PScan {
  const uint8_t* data;        // entropy-coded bytes of the scan (device; 4-byte aligned, 16 readable bytes behind)
  const Tables* tab;          // Huffman tables in force (device)
  int16_t* coef;              // the image's coefficient buffer
  int32_t* dcdiff;            // DC first scans on the parallel path: differences in scan order [n_blk]
  uint4* state;               // [n_sub] self-synchronisation states (jpeg_decode.cuh)
  uint32_t* prefix;           // [n_sub] blocks completed before each subsequence
  uint2* consumed;
  uint32_t len;
  int32_t n_sub;
  int32_t n_blk;              // blocks the scan codes
  int32_t mode;               // PSCAN_*
  int32_t ns, comp;           // components in the scan; frame index of the first (the only one when ns == 1)
  int32_t td[3], ta;
  int32_t ss, se, al, ri;
  int32_t bpm;                // blocks per MCU of THIS scan
  int32_t hs, vs, mcux, img_bpm, ncomp, bw;   // image geometry; bw = block columns of a single-component scan
  // scans decoded by one warp each, all in ONE launch (jprog_seq_kernel): block i may be touched once every scan
  // listed in dep[] -- earlier scans of the same blocks in the same order -- has published progress > i
  uint32_t* progress;         // blocks of this scan that are final (device word, zeroed before the launch)
  const uint32_t* dep[4];
  int32_t ndep;
};

// Scan-order block index -> block index in the coefficient buffer.
PJ_HD int64_t pscan_block(const PScan& S, int blk) {
  if (S.ns > 1 || S.ncomp == 1) return blk;
  const int bx = blk % S.bw, by = blk / S.bw;
  int m, b;
  if (S.comp == 0) {
    m = (by / S.vs) * S.mcux + bx / S.hs;
    b = (by % S.vs) * S.hs + bx % S.hs;
  } else {
    m = by * S.mcux + bx;
    b = S.hs * S.vs + S.comp - 1;
  }
  return (int64_t)m * S.img_bpm + b;
}

// The same map for a walk over consecutive blocks: no divisions (luma sampling factors are 1 or 2).
struct BlockWalk {
  int bx, by;
};
PJ_HD void walk_start(const PScan& S, int blk, BlockWalk* w) {
  w->bx = blk % S.bw;
  w->by = blk / S.bw;
}
PJ_HD int64_t walk_index(const PScan& S, const BlockWalk& w, int blk) {
  if (S.ns > 1 || S.ncomp == 1) return blk;
  if (S.comp == 0) {
    const int sx = S.hs - 1, sy = S.vs - 1;     // shift and mask at once: factors are 1 or 2
    return (int64_t)((w.by >> sy) * S.mcux + (w.bx >> sx)) * S.img_bpm + ((w.by & sy) << sx) + (w.bx & sx);
  }
  return (int64_t)(w.by * S.mcux + w.bx) * S.img_bpm + S.hs * S.vs + S.comp - 1;
}
PJ_HD void walk_next(const PScan& S, BlockWalk* w) {
  if (++w->bx == S.bw) { w->bx = 0; ++w->by; }
}

// Geometry and parameters of scan `ps` of file `H` (pointers are left to the caller).
inline void pscan_describe(const Header& H, const ProgScan& ps, PScan* S) {
  S->mode = ps.ss == 0 ? (ps.ah == 0 ? PSCAN_DC_FIRST : PSCAN_DC_REFINE) : (ps.ah == 0 ? PSCAN_AC_FIRST : PSCAN_AC_REFINE);
  S->ns = ps.ns;
  S->comp = ps.comp;
  for (int j = 0; j < 3; ++j) S->td[j] = ps.td[j];
  S->ta = ps.ta;
  S->ss = ps.ss;
  S->se = ps.se;
  S->al = ps.al;
  S->ri = ps.ri;
  S->hs = H.hs;
  S->vs = H.vs;
  S->mcux = H.mcux;
  S->ncomp = H.ncomp;
  S->img_bpm = H.ncomp == 1 ? 1 : H.hs * H.vs + 2;
  S->len = (uint32_t)ps.len;
  if (ps.ns > 1 || H.ncomp == 1) {
    S->bpm = S->img_bpm;
    S->bw = H.mcux;
    S->n_blk = H.mcux * H.mcuy * S->img_bpm;
  } else {
    // a component's own size: ceil(w * hs_c / hmax) samples (A.1.1), ceil(that / 8) blocks
    const int cw = ps.comp == 0 ? H.w : (H.w + H.hs - 1) / H.hs, chh = ps.comp == 0 ? H.h : (H.h + H.vs - 1) / H.vs;
    S->bpm = 1;
    S->bw = (cw + 7) / 8;
    S->n_blk = S->bw * ((chh + 7) / 8);
  }
}

// One thread builds all four tables (the cooperative form is build_tables in jpeg_decode.cuh).
PJ_HD void build_tables_seq(const Tables* g, SmemTables& s) {
  for (int t = 0; t < 4; ++t) {
    for (int i = 0; i < 256; ++i) { s.vals[t][i] = g->vals[t][i]; s.lut[t][i] = 0; }
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
}

// ---- sequential bit reader: bytes one at a time, stuffed zeros dropped, zeros fed behind a marker ----
struct SeqBits {
  const uint8_t* p;
  uint32_t n, o;
  uint64_t acc;     // unread bits, top aligned
  int cnt;
  int marker;       // pending marker byte (0 = none)
};

PJ_HD void sb_fill(SeqBits& b) {
  while (b.cnt <= 56) {
    uint32_t v = 0;
    if (!b.marker && b.o < b.n) {
      v = b.p[b.o];
      if (v == 0xFFu) {
        uint32_t k = b.o + 1;
        while (k < b.n && b.p[k] == 0xFFu) ++k;          // fill bytes
        const uint32_t m = k < b.n ? b.p[k] : 0xD9u;
        if (m == 0) { b.o = k + 1; }
        else { b.marker = (int)m; b.o = k + 1; v = 0; }
      } else {
        b.o++;
      }
    }
    b.acc |= (uint64_t)v << (56 - b.cnt);
    b.cnt += 8;
  }
}
PJ_HD int sb_bits(SeqBits& b, int nbits) {       // nbits <= 16
  if (nbits == 0) return 0;
  if (b.cnt < nbits) sb_fill(b);
  const int v = (int)(b.acc >> (64 - nbits));
  b.acc <<= nbits;
  b.cnt -= nbits;
  return v;
}
PJ_HD int sb_sym(const SmemTables& T, int t, SeqBits& b) {
  if (b.cnt < 16) sb_fill(b);
  const uint32_t v = (uint32_t)(b.acc >> 48);
  const uint32_t e = T.lut[t][v >> 8];
  if (e) {
    b.acc <<= (e >> 8);
    b.cnt -= (int)(e >> 8);
    return (int)(e & 255u);
  }
  for (int l = 9; l <= 16; ++l) {
    const int code = (int)(v >> (16 - l));
    if (code <= T.maxcode[t][l] && code >= T.mincode[t][l]) {
      b.acc <<= l;
      b.cnt -= l;
      return T.vals[t][(T.valptr[t][l] + code - T.mincode[t][l]) & 255];
    }
  }
  b.acc <<= 16;
  b.cnt -= 16;
  return 0;
}
PJ_HD int pj_extend(int v, int s) { return (s && v < (1 << (s - 1))) ? v - (1 << s) + 1 : v; }

PJ_HD void sb_restart(SeqBits& b, int* rst_next) {
  b.acc = 0;
  b.cnt = 0;
  if (!b.marker) {
    while (b.o + 1 < b.n && !(b.p[b.o] == 0xFFu && b.p[b.o + 1] >= 0xD0u && b.p[b.o + 1] <= 0xD7u)) b.o++;
    if (b.o + 1 < b.n) { b.marker = b.p[b.o + 1]; b.o += 2; }
  }
  if (b.marker == 0xD0 + *rst_next) b.marker = 0;
  *rst_next = (*rst_next + 1) & 7;
}

// A correction bit for a coefficient that is already non-zero (G.1.2.3): its magnitude grows by one bit of precision.
PJ_HD void pj_correct(int16_t* cf, int p1) {
  const int v = *cf;
  if ((v & p1) == 0) *cf = (int16_t)(v >= 0 ? v + p1 : v - p1);
}

// Bit-reader interface of the refinement code below: SeqBits here, the word-wise Reader of jpeg_decode.cuh on the GPU.
PJ_HD int br_sym(const SmemTables& T, int t, SeqBits& b) { return sb_sym(T, t, b); }
PJ_HD int br_bits(SeqBits& b, int nbits) { return sb_bits(b, nbits); }
PJ_HD void br_restart(SeqBits& b, int* rst_next) { sb_restart(b, rst_next); }

#if defined(__CUDACC__)
#define PJ_CTZ64(x) (__ffsll((long long)(x)) - 1)
#define PJ_POPC64(x) __popcll(x)
#else
#define PJ_CTZ64(x) __builtin_ctzll(x)
#define PJ_POPC64(x) __builtin_popcountll(x)
#endif

// ---- AC refinement, block by block, with the block's history as a bit mask (G.1.2.3, Figure G.7) ----
// The bit stream can only be parsed by ONE thread (jpeg_prog.cuh), so that thread does nothing but parse: given the
// mask nz of coefficients that are non-zero on entry (zig-zag bit k), it walks the symbols with mask algebra -- a run
// of r zero-history coefficients drops r set bits from the zero mask, the lowest remaining one takes the new value,
// the non-zero ones passed on the way each own one correction bit -- and RECORDS what to do: the coefficients that
// get a correction bit (corrmask) with their bits in stream order (corrbits, first bit on top), the newly non-zero
// ones (newmask) and their signs (negmask).  pj_refine_apply then updates one coefficient; on the GPU every lane of
// the warp applies two.
struct RefineOut {
  uint64_t corrmask, corrbits, newmask, negmask;
};

template <class BR>
PJ_HD void pj_refine_take(BR& b, uint64_t m, RefineOut* o) {
  int n = PJ_POPC64(m);
  if (n == 0) return;
  o->corrmask |= m;
  if (n > 32) {                                   // br_bits hands out at most 32 bits
    o->corrbits = (o->corrbits << 32) | (uint32_t)br_bits(b, 32);
    n -= 32;
  }
  o->corrbits = (o->corrbits << n) | (uint32_t)br_bits(b, n);
}

// band = bits ss..se.  Same effect on the block as the coefficient-by-coefficient loop of pscan_sequential.
template <class BR>
PJ_HD void pj_refine_parse(BR& b, const SmemTables& T, int tac, uint64_t nz, uint64_t band, uint32_t* eobrun, RefineOut* o) {
  o->corrmask = o->corrbits = o->newmask = o->negmask = 0;
  uint64_t from = band;                           // positions of the band not yet passed
  if (*eobrun == 0) {
    while (from) {
      const int rs = br_sym(T, tac, b);
      const int r = rs >> 4, s = rs & 15;
      bool neg = false;
      if (s) {
        neg = br_bits(b, 1) == 0;
      } else if (r != 15) {
        *eobrun = 1u << r;
        if (r) *eobrun += (uint32_t)br_bits(b, r);
        break;
      }
      uint64_t zeros = ~nz & from;
      for (int i = 0; i < r && zeros; ++i) zeros &= zeros - 1;    // the run's r zero-history coefficients
      const uint64_t low = zeros & (0 - zeros);                   // the next one: takes the new value (or ends a ZRL)
      const uint64_t upto = zeros ? low - 1 : ~0ull;              // everything in front of it
      pj_refine_take(b, nz & from & upto, o);
      if (s && zeros) {
        o->newmask |= low;
        if (neg) o->negmask |= low;
      }
      from = zeros ? from & ~(upto | low) : 0;
    }
  }
  if (*eobrun > 0) {
    pj_refine_take(b, nz & from, o);
    --*eobrun;
  }
}

// The same parse with the positions of the band's zero-history coefficients listed (zpos[0 .. nz0), ascending -- on
// the GPU the lanes write that list, one popc each): a run then costs one table read instead of a bit loop, and the
// number of correction bits in front of the stop is (stop - k - r), no mask needed.  This is the form the warp kernel
// runs; ~65 instructions per symbol on the parsing thread instead of ~115.
template <class BR>
PJ_HD void pj_refine_parse_z(BR& b, const SmemTables& T, int tac, uint64_t nz, uint64_t band, const uint8_t* zpos, int nz0,
                             int ss, int se, uint32_t* eobrun, RefineOut* o) {
  uint64_t corrbits = 0, newmask = 0, negmask = 0;
  int k = ss, zi = 0;
  auto take = [&](int n) {
    if (n > 32) {
      corrbits = (corrbits << 32) | (uint32_t)br_bits(b, 32);
      n -= 32;
    }
    if (n > 0) corrbits = (corrbits << n) | (uint32_t)br_bits(b, n);
  };
  if (*eobrun == 0) {
    while (k <= se) {
      const int rs = br_sym(T, tac, b);
      const int r = rs >> 4, s = rs & 15;
      bool neg = false;
      if (s) {
        neg = br_bits(b, 1) == 0;
      } else if (r != 15) {
        *eobrun = 1u << r;
        if (r) *eobrun += (uint32_t)br_bits(b, r);
        break;
      }
      const int t = zi + r;                        // index of the zero-history coefficient the run stops at
      if (t < nz0) {
        const int stop = zpos[t];
        take(stop - k - r);
        if (s) {
          newmask |= 1ull << stop;
          if (neg) negmask |= 1ull << stop;
        }
        zi = t + 1;
        k = stop + 1;
      } else {                                     // fewer zeros left than the run asks for: to the end of the band
        take((se + 1 - k) - (nz0 - zi));
        zi = nz0;
        k = se + 1;
      }
    }
  }
  if (*eobrun > 0) {
    if (k <= se) take((se + 1 - k) - (nz0 - zi));
    k = se + 1;
    --*eobrun;
  }
  o->corrbits = corrbits;
  o->newmask = newmask;
  o->negmask = negmask;
  o->corrmask = nz & band & (k > 63 ? ~0ull : ((1ull << k) - 1));   // every non-zero coefficient in front of k was passed
}

// Coefficient k of the block after the scan (v = its value before).
PJ_HD int pj_refine_apply(int v, int k, int p1, const RefineOut& o) {
  const uint64_t bit = 1ull << k;
  if (o.corrmask & bit) {
    const int rank = PJ_POPC64(o.corrmask & (bit - 1)), total = PJ_POPC64(o.corrmask);
    if (((o.corrbits >> (total - 1 - rank)) & 1) && (v & p1) == 0) v = v >= 0 ? v + p1 : v - p1;
  } else if (o.newmask & bit) {
    v = (o.negmask & bit) ? -p1 : p1;
  }
  return v;
}

// The whole scan, block after block (all four kinds).  The steps are those of T.81 G.1.2 / Figures G.3 - G.7.
PJ_HD void pscan_sequential(const PScan& S, const SmemTables& T) {
  SeqBits b;
  b.p = S.data;
  b.n = S.len;
  b.o = 0;
  b.acc = 0;
  b.cnt = 0;
  b.marker = 0;
  int pred[3] = {0, 0, 0};
  int rst_left = S.ri, rst_next = 0;
  uint32_t eobrun = 0;
  const int p1 = 1 << S.al;
  const int ny = S.ns > 1 ? S.hs * S.vs : 1;
  const int tac = 2 + S.ta;
  int b_in_mcu = 0;
  for (int blk = 0; blk < S.n_blk; ++blk) {
    if (b_in_mcu == 0 && S.ri) {
      if (rst_left == 0) {
        sb_restart(b, &rst_next);
        pred[0] = pred[1] = pred[2] = 0;
        eobrun = 0;
        rst_left = S.ri;
      }
      rst_left--;
    }
    int16_t* c = S.coef + pscan_block(S, blk) * 64;
    if (S.mode == PSCAN_DC_FIRST) {
      const int ci = S.ns == 1 ? 0 : (b_in_mcu < ny ? 0 : b_in_mcu - ny + 1);
      const int s = sb_sym(T, S.td[ci], b) & 15;
      const int diff = s ? pj_extend(sb_bits(b, s), s) : 0;
      pred[ci] += diff;
      c[0] = (int16_t)(pred[ci] * p1);
    } else if (S.mode == PSCAN_DC_REFINE) {
      if (sb_bits(b, 1)) c[0] = (int16_t)(c[0] | p1);
    } else if (S.mode == PSCAN_AC_FIRST) {
      if (eobrun > 0) {
        eobrun--;
      } else {
        for (int k = S.ss; k <= S.se; ++k) {
          const int rs = sb_sym(T, tac, b);
          const int r = rs >> 4, s = rs & 15;
          if (s) {
            k += r;
            const int v = pj_extend(sb_bits(b, s), s);
            if (k <= 63) c[k] = (int16_t)(v * p1);
          } else if (r == 15) {
            k += 15;
          } else {
            eobrun = 1u << r;
            if (r) eobrun += (uint32_t)sb_bits(b, r);
            eobrun--;
            break;
          }
        }
      }
    } else {
      int k = S.ss;
      if (eobrun == 0) {
        for (; k <= S.se; ++k) {
          const int rs = sb_sym(T, tac, b);
          int r = rs >> 4, s = rs & 15;
          if (s) {
            s = sb_bits(b, 1) ? p1 : -p1;
          } else if (r != 15) {
            eobrun = 1u << r;
            if (r) eobrun += (uint32_t)sb_bits(b, r);
            break;
          }
          do {
            int16_t* cf = c + k;
            if (*cf != 0) {
              if (sb_bits(b, 1)) pj_correct(cf, p1);
            } else if (--r < 0) {
              break;
            }
            ++k;
          } while (k <= S.se);
          if (s && k <= 63) c[k] = (int16_t)s;
        }
      }
      if (eobrun > 0) {
        for (; k <= S.se; ++k) {
          int16_t* cf = c + k;
          if (*cf != 0 && sb_bits(b, 1)) pj_correct(cf, p1);
        }
        eobrun--;
      }
    }
    b_in_mcu = (b_in_mcu + 1 == S.bpm) ? 0 : b_in_mcu + 1;
  }
}

}  // namespace jpegdec
}  // namespace stitch
