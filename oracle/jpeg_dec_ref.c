/* TEST INFRASTRUCTURE (oracle) -- sequential baseline JPEG decoder, the checker of the GPU decoder.
 *
 * The reference decodes its inputs inside the closed WeChat runtime (utils/canvas.js:27-121 loadImageFrom ->
 * canvas.createImage(); pages/index/index.js:1441-1520); what reaches drawImage is the decoded RGBA bitmap.
 * The de-facto definition of that bitmap is libjpeg's default pipeline, restated here from the published
 * algorithms: ITU-T T.81 (markers, Huffman decoding F.2.2, dequantisation), the Loeffler-Ligtenberg-Moschytz
 * integer IDCT with 13-bit constants ("islow"), triangle-filter ("fancy") chroma upsampling and the 16-bit
 * fixed-point YCbCr->RGB conversion of JFIF.  PINNED: tests/test_jpeg_decode.py holds it byte-identical to
 * libjpeg(-turbo) as shipped in PIL for 4:4:4, 4:2:2, 4:2:0 and grayscale files.
 *
 * Supported: baseline / extended sequential Huffman (SOF0, SOF1) with one interleaved scan, and progressive Huffman
 * (SOF2, T.81 Annex G: spectral selection + successive approximation, interleaved DC scans, single-component scans,
 * EOB runs, tables redefined between scans); 8-bit, 1 or 3 components, luma sampling 1x1, 2x1 or 2x2 with 1x1
 * chroma, restart intervals.  A complete progressive file is decoded without libjpeg's inter-block smoothing (that only
 * acts on files whose AC scans are missing).  Anything else returns < 0.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "jpeg_tables.inc"

typedef struct {
  int w, h, ncomp;
  int hs[3], vs[3], tq[3], td[3], ta[3];
  int hmax, vmax, mcux, mcuy, ri;
  uint16_t q[4][64];            /* natural order */
  /* Huffman tables [class*2 + id]: canonical decode data (T.81 F.2.2.3) */
  int have_h[4];
  uint8_t bits[4][17], vals[4][256];
  int32_t mincode[4][17], maxcode[4][18], valptr[4][17];
  const uint8_t* scan;
  size_t scan_len;
  int progressive;              /* SOF2 */
  int comp_id[3];
  const uint8_t* file;          /* progressive: the marker walk restarts at first_sos for every decode */
  size_t file_len, first_sos;
} JDec;

static void build_huff(JDec* d, int t) {
  int code = 0, k = 0;
  for (int l = 1; l <= 16; ++l) {
    d->valptr[t][l] = k;
    d->mincode[t][l] = code;
    code += d->bits[t][l];
    k += d->bits[t][l];
    d->maxcode[t][l] = d->bits[t][l] ? code - 1 : -1;
    code <<= 1;
  }
  d->maxcode[t][17] = 0x7fffffff;
}

static int parse(JDec* d, const uint8_t* p, size_t n) {
  memset(d, 0, sizeof(*d));
  if (n < 4 || p[0] != 0xFF || p[1] != 0xD8) return -1;
  size_t o = 2;
  int have_sof = 0;
  while (o + 4 <= n) {
    if (p[o] != 0xFF) return -2;
    const int m = p[o + 1];
    if (m == 0xFF) { ++o; continue; }
    if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { o += 2; continue; }
    const size_t len = ((size_t)p[o + 2] << 8) | p[o + 3];
    if (len < 2 || o + 2 + len > n) return -3;
    const uint8_t* s = p + o + 4;
    const size_t sl = len - 2;
    if (m == 0xDB) {  /* DQT */
      size_t i = 0;
      while (i < sl) {
        const int pq = s[i] >> 4, tq = s[i] & 15;
        if (pq != 0 || tq > 3 || i + 65 > sl) return -4;
        for (int k = 0; k < 64; ++k) d->q[tq][JZIGZAG[k]] = s[i + 1 + k];
        i += 65;
      }
    } else if (m == 0xC0 || m == 0xC1 || m == 0xC2) {  /* SOF0 / SOF1 / SOF2 */
      if (sl < 6 || s[0] != 8) return -5;
      d->progressive = (m == 0xC2);
      d->h = (s[1] << 8) | s[2];
      d->w = (s[3] << 8) | s[4];
      d->ncomp = s[5];
      if ((d->ncomp != 1 && d->ncomp != 3) || sl < (size_t)(6 + 3 * d->ncomp) || d->w == 0 || d->h == 0) return -6;
      for (int c = 0; c < d->ncomp; ++c) {
        d->comp_id[c] = s[6 + 3 * c];
        d->hs[c] = s[7 + 3 * c] >> 4;
        d->vs[c] = s[7 + 3 * c] & 15;
        d->tq[c] = s[8 + 3 * c];
        if (d->tq[c] > 3) return -6;
      }
      if (d->ncomp == 1) { d->hs[0] = d->vs[0] = 1; }
      else {
        if (d->hs[1] != 1 || d->vs[1] != 1 || d->hs[2] != 1 || d->vs[2] != 1) return -7;
        if (!((d->hs[0] == 1 && d->vs[0] == 1) || (d->hs[0] == 2 && d->vs[0] == 1) || (d->hs[0] == 2 && d->vs[0] == 2)))
          return -7;
      }
      d->hmax = d->hs[0];
      d->vmax = d->vs[0];
      d->mcux = (d->w + 8 * d->hmax - 1) / (8 * d->hmax);
      d->mcuy = (d->h + 8 * d->vmax - 1) / (8 * d->vmax);
      have_sof = 1;
    } else if (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
      return -8;  /* lossless, hierarchical, arithmetic */
    } else if (m == 0xC4) {  /* DHT */
      size_t i = 0;
      while (i < sl) {
        const int tc = s[i] >> 4, th = s[i] & 15;
        if (tc > 1 || th > 1 || i + 17 > sl) return -9;
        const int t = tc * 2 + th;
        int cnt = 0;
        for (int l = 1; l <= 16; ++l) { d->bits[t][l] = s[i + l]; cnt += s[i + l]; }
        if (cnt > 256 || i + 17 + cnt > sl) return -9;
        memcpy(d->vals[t], s + i + 17, cnt);
        build_huff(d, t);
        d->have_h[t] = 1;
        i += 17 + cnt;
      }
    } else if (m == 0xDD) {
      if (sl < 2) return -10;
      d->ri = (s[0] << 8) | s[1];
    } else if (m == 0xDA && have_sof && d->progressive) {  /* first SOS of a progressive file: scans are walked later */
      d->file = p;
      d->file_len = n;
      d->first_sos = o;
      return 0;
    } else if (m == 0xDA) {  /* SOS */
      if (!have_sof || sl < 1 || s[0] != d->ncomp || sl < (size_t)(1 + 2 * d->ncomp + 3)) return -11;
      for (int c = 0; c < d->ncomp; ++c) {
        d->td[c] = s[2 + 2 * c] >> 4;
        d->ta[c] = s[2 + 2 * c] & 15;
        if (d->td[c] > 1 || d->ta[c] > 1 || !d->have_h[d->td[c]] || !d->have_h[2 + d->ta[c]]) return -11;
      }
      d->scan = p + o + 2 + len;
      d->scan_len = n - (o + 2 + len);
      return 0;
    }
    o += 2 + len;
  }
  return -12;
}

/* ---- entropy decoding (T.81 F.2.2) ---- */
typedef struct {
  const uint8_t* p;
  size_t n, o;
  uint32_t acc;
  int cnt;
  int marker;   /* pending marker byte (0 = none) */
} Bits;

static void fill(Bits* b) {
  while (b->cnt <= 24) {
    int v = 0;  /* past a marker / the end the decoder sees zero bits (libjpeg feeds zeros too) */
    if (!b->marker && b->o < b->n) {
      v = b->p[b->o];
      if (v == 0xFF) {
        int m = 0;
        size_t k = b->o + 1;
        while (k < b->n && b->p[k] == 0xFF) ++k;   /* fill bytes */
        if (k < b->n) m = b->p[k];
        if (m == 0) { b->o = k + 1; v = 0xFF; }      /* stuffed zero */
        else { b->marker = m; b->o = k + 1; v = 0; }
      } else {
        b->o++;
      }
    }
    b->acc |= (uint32_t)v << (24 - b->cnt);
    b->cnt += 8;
  }
}
static inline int getbits(Bits* b, int n) {
  if (n == 0) return 0;
  fill(b);
  const int v = (int)(b->acc >> (32 - n));
  b->acc <<= n;
  b->cnt -= n;
  return v;
}
static int decode_sym(const JDec* d, int t, Bits* b) {
  fill(b);
  int code = 0;
  for (int l = 1; l <= 16; ++l) {
    code = (code << 1) | (int)(b->acc >> 31);
    b->acc <<= 1;
    b->cnt--;
    if (d->maxcode[t][l] >= 0 && code <= d->maxcode[t][l] && code >= d->mincode[t][l])
      return d->vals[t][d->valptr[t][l] + code - d->mincode[t][l]];
  }
  return 0;  /* invalid code: libjpeg warns and yields 0 */
}
static inline int extend(int v, int s) { return (s && v < (1 << (s - 1))) ? v - (1 << s) + 1 : v; }

/* Quantised coefficients, natural order, one int16[64] per block; blocks in component planes of
 * (mcux*hs) x (mcuy*vs) blocks, row-major.  coef[c] must hold that many blocks. */
static int decode_scan(const JDec* d, int16_t* coef[3]) {
  Bits b;
  memset(&b, 0, sizeof b);
  b.p = d->scan;
  b.n = d->scan_len;
  int pred[3] = {0, 0, 0};
  const int nmcu = d->mcux * d->mcuy;
  int rst_left = d->ri, rst_next = 0;
  for (int m = 0; m < nmcu; ++m) {
    if (d->ri && rst_left == 0) {
      /* discard remaining bits, expect RSTn */
      b.acc = 0;
      b.cnt = 0;
      if (!b.marker) {  /* marker not yet seen: skip to it */
        while (b.o + 1 < b.n && !(b.p[b.o] == 0xFF && b.p[b.o + 1] >= 0xD0 && b.p[b.o + 1] <= 0xD7)) b.o++;
        if (b.o + 1 < b.n) { b.marker = b.p[b.o + 1]; b.o += 2; }
      }
      if (b.marker == 0xD0 + rst_next) b.marker = 0;
      rst_next = (rst_next + 1) & 7;
      pred[0] = pred[1] = pred[2] = 0;
      rst_left = d->ri;
    }
    const int mx = m % d->mcux, my = m / d->mcux;
    for (int c = 0; c < d->ncomp; ++c) {
      const int bw = d->mcux * d->hs[c];
      for (int v = 0; v < d->vs[c]; ++v)
        for (int h = 0; h < d->hs[c]; ++h) {
          int16_t* blk = coef[c] + ((size_t)(my * d->vs[c] + v) * bw + (mx * d->hs[c] + h)) * 64;
          memset(blk, 0, 128);
          int s = decode_sym(d, d->td[c], &b);
          int diff = s ? extend(getbits(&b, s & 15), s & 15) : 0;
          pred[c] += diff;
          blk[0] = (int16_t)pred[c];
          for (int k = 1; k < 64;) {
            const int rs = decode_sym(d, 2 + d->ta[c], &b);
            const int r = rs >> 4, sz = rs & 15;
            if (sz == 0) {
              if (r != 15) break;  /* EOB */
              k += 16;
              continue;
            }
            k += r;
            if (k > 63) break;
            blk[JZIGZAG[k]] = (int16_t)extend(getbits(&b, sz), sz);
            ++k;
          }
        }
    }
    if (d->ri) rst_left--;
  }
  return 0;
}


/* ---- progressive scans (T.81 Annex G; the same steps as libjpeg's jdphuff.c) ---- */
static inline int getbit(Bits* b) { return getbits(b, 1); }

/* End of the entropy-coded data that starts at o: the first marker that is neither a stuffed zero nor RSTn. */
static size_t scan_end(const uint8_t* p, size_t n, size_t o) {
  while (o + 1 < n) {
    if (p[o] == 0xFF && p[o + 1] != 0 && p[o + 1] != 0xFF && !(p[o + 1] >= 0xD0 && p[o + 1] <= 0xD7)) return o;
    ++o;
  }
  return n;
}

static void restart(Bits* b, int* rst_next) {
  b->acc = 0;
  b->cnt = 0;
  if (!b->marker) {
    while (b->o + 1 < b->n && !(b->p[b->o] == 0xFF && b->p[b->o + 1] >= 0xD0 && b->p[b->o + 1] <= 0xD7)) b->o++;
    if (b->o + 1 < b->n) { b->marker = b->p[b->o + 1]; b->o += 2; }
  }
  if (b->marker == 0xD0 + *rst_next) b->marker = 0;
  *rst_next = (*rst_next + 1) & 7;
}

static int16_t* block_at(const JDec* d, int16_t* coef[3], int c, int bx, int by) {
  return coef[c] + ((size_t)by * (d->mcux * d->hs[c]) + bx) * 64;
}

/* One scan: ns components cs[] (frame indices), band ss..se, successive approximation ah/al. */
static int decode_prog_scan(JDec* d, int16_t* coef[3], const uint8_t* data, size_t len, int ns, const int* cs, const int* td,
                            const int* ta, int ss, int se, int ah, int al) {
  Bits b;
  memset(&b, 0, sizeof b);
  b.p = data;
  b.n = len;
  int pred[3] = {0, 0, 0};
  int rst_left = d->ri, rst_next = 0;
  unsigned eobrun = 0;
  if (ss == 0) {                       /* DC scan: may interleave components */
    if (se != 0) return -31;
    int nmx, nmy;
    if (ns == 1) {
      const int c = cs[0];
      nmx = ((d->w * d->hs[c] + d->hmax - 1) / d->hmax + 7) / 8;
      nmy = ((d->h * d->vs[c] + d->vmax - 1) / d->vmax + 7) / 8;
    } else {
      nmx = d->mcux;
      nmy = d->mcuy;
    }
    for (int m = 0; m < nmx * nmy; ++m) {
      if (d->ri && rst_left == 0) { restart(&b, &rst_next); pred[0] = pred[1] = pred[2] = 0; rst_left = d->ri; }
      const int mx = m % nmx, my = m / nmx;
      for (int j = 0; j < ns; ++j) {
        const int c = cs[j];
        const int nh = ns == 1 ? 1 : d->hs[c], nv = ns == 1 ? 1 : d->vs[c];
        for (int v = 0; v < nv; ++v)
          for (int h = 0; h < nh; ++h) {
            int16_t* blk = block_at(d, coef, c, mx * nh + h, my * nv + v);
            if (ah == 0) {
              const int s = decode_sym(d, td[j], &b);
              const int diff = s ? extend(getbits(&b, s & 15), s & 15) : 0;
              pred[c] += diff;
              blk[0] = (int16_t)(pred[c] * (1 << al));
            } else if (getbit(&b)) {
              blk[0] |= (int16_t)(1 << al);
            }
          }
      }
      if (d->ri) rst_left--;
    }
    return 0;
  }
  /* AC scan: one component, blocks of the component's own (unpadded) size in raster order */
  if (ns != 1 || se > 63 || ss > se) return -32;
  const int c = cs[0];
  const int nbx = ((d->w * d->hs[c] + d->hmax - 1) / d->hmax + 7) / 8;
  const int nby = ((d->h * d->vs[c] + d->vmax - 1) / d->vmax + 7) / 8;
  const int p1 = 1 << al, m1 = -(1 << al);
  const int t = 2 + ta[0];
  for (int m = 0; m < nbx * nby; ++m) {
    if (d->ri && rst_left == 0) { restart(&b, &rst_next); eobrun = 0; rst_left = d->ri; }
    int16_t* blk = block_at(d, coef, c, m % nbx, m / nbx);
    if (ah == 0) {
      if (eobrun > 0) {
        eobrun--;
      } else {
        for (int k = ss; k <= se; ++k) {
          const int rs = decode_sym(d, t, &b);
          const int r = rs >> 4, s = rs & 15;
          if (s) {
            k += r;
            const int v = extend(getbits(&b, s), s);
            if (k <= 63) blk[JZIGZAG[k]] = (int16_t)(v * (1 << al));
          } else if (r == 15) {
            k += 15;
          } else {
            eobrun = 1u << r;
            if (r) eobrun += (unsigned)getbits(&b, r);
            eobrun--;
            break;
          }
        }
      }
    } else {
      int k = ss;
      if (eobrun == 0) {
        for (; k <= se; ++k) {
          const int rs = decode_sym(d, t, &b);
          int r = rs >> 4, s = rs & 15;
          if (s) {
            s = getbit(&b) ? p1 : m1;       /* size must be 1: a newly non-zero coefficient */
          } else if (r != 15) {
            eobrun = 1u << r;
            if (r) eobrun += (unsigned)getbits(&b, r);
            break;
          }
          do {
            int16_t* cf = blk + JZIGZAG[k];
            if (*cf != 0) {
              if (getbit(&b) && (*cf & p1) == 0) *cf = (int16_t)(*cf + (*cf >= 0 ? p1 : m1));
            } else if (--r < 0) {
              break;
            }
            ++k;
          } while (k <= se);
          if (s && k <= 63) blk[JZIGZAG[k]] = (int16_t)s;
        }
      }
      if (eobrun > 0) {
        for (; k <= se; ++k) {
          int16_t* cf = blk + JZIGZAG[k];
          if (*cf != 0 && getbit(&b) && (*cf & p1) == 0) *cf = (int16_t)(*cf + (*cf >= 0 ? p1 : m1));
        }
        eobrun--;
      }
    }
    if (d->ri) rst_left--;
  }
  return 0;
}

/* Walks the marker segments from the first SOS to EOI: DHT / DRI between scans take effect for the scans after them. */
static int decode_progressive(JDec* d, int16_t* coef[3]) {
  const uint8_t* p = d->file;
  const size_t n = d->file_len;
  size_t o = d->first_sos;
  int scans = 0;
  while (o + 4 <= n) {
    if (p[o] != 0xFF) return -33;
    const int m = p[o + 1];
    if (m == 0xFF) { ++o; continue; }
    if (m == 0xD9) break;
    if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { o += 2; continue; }
    const size_t len = ((size_t)p[o + 2] << 8) | p[o + 3];
    if (len < 2 || o + 2 + len > n) return -34;
    const uint8_t* s = p + o + 4;
    const size_t sl = len - 2;
    if (m == 0xC4) {
      size_t i = 0;
      while (i < sl) {
        const int tc = s[i] >> 4, th = s[i] & 15;
        if (tc > 1 || th > 1 || i + 17 > sl) return -9;
        const int t = tc * 2 + th;
        int cnt = 0;
        for (int l = 1; l <= 16; ++l) { d->bits[t][l] = s[i + l]; cnt += s[i + l]; }
        if (cnt > 256 || i + 17 + cnt > sl) return -9;
        memset(d->vals[t], 0, 256);
        memcpy(d->vals[t], s + i + 17, cnt);
        build_huff(d, t);
        d->have_h[t] = 1;
        i += 17 + cnt;
      }
    } else if (m == 0xDD) {
      if (sl < 2) return -10;
      d->ri = (s[0] << 8) | s[1];
    } else if (m == 0xDA) {
      if (sl < 1) return -11;
      const int ns = s[0];
      if (ns < 1 || ns > 3 || sl < (size_t)(1 + 2 * ns + 3)) return -11;
      if (ns != 1 && ns != d->ncomp) return -35;          /* partial interleaves: not produced by any common encoder */
      int cs[3], td[3], ta[3];
      for (int j = 0; j < ns; ++j) {
        cs[j] = -1;
        for (int c = 0; c < d->ncomp; ++c)
          if (d->comp_id[c] == s[1 + 2 * j]) cs[j] = c;
        if (cs[j] < 0 || (ns > 1 && cs[j] != j)) return -35;
        td[j] = s[2 + 2 * j] >> 4;
        ta[j] = s[2 + 2 * j] & 15;
        if (td[j] > 1 || ta[j] > 1) return -11;
      }
      const int ss = s[1 + 2 * ns], se = s[2 + 2 * ns], ah = s[3 + 2 * ns] >> 4, al = s[3 + 2 * ns] & 15;
      if (al > 13) return -36;
      for (int j = 0; j < ns; ++j)
        if ((ss == 0 && ah == 0 && !d->have_h[td[j]]) || (ss > 0 && !d->have_h[2 + ta[j]])) return -11;
      const size_t start = o + 2 + len, end = scan_end(p, n, start);
      const int rc = decode_prog_scan(d, coef, p + start, end - start, ns, cs, td, ta, ss, se, ah, al);
      if (rc) return rc;
      ++scans;
      o = end;
      continue;
    }
    o += 2 + len;
  }
  return scans ? 0 : -37;
}

/* ---- inverse DCT, "islow": LL&M with 13-bit constants, two passes ---- */
#define CB 13
#define P1 2
#define DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))
static inline uint8_t clamp8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

static void idct_1d(const int32_t* in, int stride, int32_t* out, int ostride, int shift, int first) {
  int32_t z2 = in[2 * stride], z3 = in[6 * stride];
  int32_t z1 = (z2 + z3) * 4433;
  int32_t tmp2 = z1 + z3 * (-15137), tmp3 = z1 + z2 * 6270;
  z2 = in[0];
  z3 = in[4 * stride];
  int32_t tmp0 = (z2 + z3) * (1 << CB), tmp1 = (z2 - z3) * (1 << CB);
  const int32_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  tmp0 = in[7 * stride];
  tmp1 = in[5 * stride];
  tmp2 = in[3 * stride];
  tmp3 = in[1 * stride];
  z1 = tmp0 + tmp3;
  z2 = tmp1 + tmp2;
  z3 = tmp0 + tmp2;
  int32_t z4 = tmp1 + tmp3;
  const int32_t z5 = (z3 + z4) * 9633;
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
  (void)first;
  out[0 * ostride] = DESCALE(tmp10 + tmp3, shift);
  out[7 * ostride] = DESCALE(tmp10 - tmp3, shift);
  out[1 * ostride] = DESCALE(tmp11 + tmp2, shift);
  out[6 * ostride] = DESCALE(tmp11 - tmp2, shift);
  out[2 * ostride] = DESCALE(tmp12 + tmp1, shift);
  out[5 * ostride] = DESCALE(tmp12 - tmp1, shift);
  out[3 * ostride] = DESCALE(tmp13 + tmp0, shift);
  out[4 * ostride] = DESCALE(tmp13 - tmp0, shift);
}

void jpeg_ref_idct_block(const int16_t* coef, const uint16_t* q, uint8_t* out, int out_stride) {
  int32_t in[64], ws[64], o[64];
  for (int i = 0; i < 64; ++i) in[i] = (int32_t)coef[i] * (int32_t)q[i];
  for (int c = 0; c < 8; ++c) idct_1d(in + c, 8, ws + c, 8, CB - P1, 1);        /* columns */
  for (int r = 0; r < 8; ++r) idct_1d(ws + 8 * r, 1, o + 8 * r, 1, CB + P1 + 3, 0); /* rows */
  for (int r = 0; r < 8; ++r)
    for (int c = 0; c < 8; ++c) out[r * out_stride + c] = clamp8(o[8 * r + c] + 128);
}

/* ---- public entry points ---- */
int jpeg_ref_info(const uint8_t* file, size_t n, int* w, int* h, int* ncomp, int* hs, int* vs) {
  JDec d;
  const int rc = parse(&d, file, n);
  if (rc) return rc;
  *w = d.w; *h = d.h; *ncomp = d.ncomp; *hs = d.hmax; *vs = d.vmax;
  return 0;
}

/* Decode to RGBA8 (alpha 255), rows of pitch bytes. */
int jpeg_ref_decode(const uint8_t* file, size_t n, uint8_t* rgba, int64_t pitch) {
  JDec d;
  int rc = parse(&d, file, n);
  if (rc) return rc;
  int16_t* coef[3] = {0, 0, 0};
  uint8_t* plane[3] = {0, 0, 0};
  int pw[3], ph[3];
  for (int c = 0; c < d.ncomp; ++c) {
    pw[c] = d.mcux * d.hs[c] * 8;
    ph[c] = d.mcuy * d.vs[c] * 8;
    coef[c] = (int16_t*)malloc((size_t)pw[c] * ph[c] * sizeof(int16_t));
    plane[c] = (uint8_t*)malloc((size_t)pw[c] * ph[c]);
    if (!coef[c] || !plane[c]) return -20;
  }
  if (d.progressive) {
    for (int c = 0; c < d.ncomp; ++c) memset(coef[c], 0, (size_t)pw[c] * ph[c] * sizeof(int16_t));
    rc = decode_progressive(&d, coef);
  } else {
    rc = decode_scan(&d, coef);
  }
  for (int c = 0; c < d.ncomp && !rc; ++c) {
    const int bw = pw[c] / 8, bh = ph[c] / 8;
#pragma omp parallel for schedule(static)
    for (int by = 0; by < bh; ++by)
      for (int bx = 0; bx < bw; ++bx)
        jpeg_ref_idct_block(coef[c] + ((size_t)by * bw + bx) * 64, d.q[d.tq[c]], plane[c] + (size_t)by * 8 * pw[c] + bx * 8, pw[c]);
  }
  if (!rc) {
    if (!pitch) pitch = (int64_t)d.w * 4;
    /* real (unpadded) chroma size: ceil(w * hs_c / hmax) etc. */
    const int cw = d.ncomp == 3 ? (d.w + d.hmax - 1) / d.hmax : 0, chh = d.ncomp == 3 ? (d.h + d.vmax - 1) / d.vmax : 0;
    const int fancy_h = d.ncomp == 3 && d.hmax == 2 && cw > 2;   /* libjpeg: fancy only when downsampled_width > 2 */
#pragma omp parallel for schedule(static)
    for (int y = 0; y < d.h; ++y) {
      uint8_t* o = rgba + (int64_t)y * pitch;
      const uint8_t* Y = plane[0] + (size_t)y * pw[0];
      if (d.ncomp == 1) {
        for (int x = 0; x < d.w; ++x) { o[4 * x] = o[4 * x + 1] = o[4 * x + 2] = Y[x]; o[4 * x + 3] = 255; }
        continue;
      }
      for (int x = 0; x < d.w; ++x) {
        int cbv, crv;
        for (int c = 1; c <= 2; ++c) {
          const uint8_t* P = plane[c];
          const int stride = pw[c];
          int val;
          if (d.hmax == 1) {
            val = P[(size_t)y * stride + x];
          } else if (d.vmax == 1) {          /* h2v1 */
            const uint8_t* r0 = P + (size_t)y * stride;
            const int i = x >> 1;
            if (!fancy_h) val = r0[i];
            else if ((x & 1) == 0) val = i == 0 ? r0[0] : (r0[i] * 3 + r0[i - 1] + 1) >> 2;
            else val = i == cw - 1 ? r0[i] : (r0[i] * 3 + r0[i + 1] + 2) >> 2;
          } else {                           /* h2v2 */
            const int r = y >> 1, i = x >> 1;
            if (!fancy_h) {
              val = P[(size_t)r * stride + i];
            } else {
              int rn = (y & 1) ? r + 1 : r - 1;          /* nearer neighbour row: above for even y, below for odd */
              if (rn < 0) rn = 0;
              if (rn > chh - 1) rn = chh - 1;
              const uint8_t* r0 = P + (size_t)r * stride;
              const uint8_t* r1 = P + (size_t)rn * stride;
              const int cur = r0[i] * 3 + r1[i];
              if ((x & 1) == 0) {
                if (i == 0) val = (cur * 4 + 8) >> 4;
                else val = (cur * 3 + (r0[i - 1] * 3 + r1[i - 1]) + 8) >> 4;
              } else {
                if (i == cw - 1) val = (cur * 4 + 7) >> 4;
                else val = (cur * 3 + (r0[i + 1] * 3 + r1[i + 1]) + 7) >> 4;
              }
            }
          }
          if (c == 1) cbv = val; else crv = val;
        }
        const int yy = Y[x], cb = cbv - 128, cr = crv - 128;
        o[4 * x + 0] = clamp8(yy + ((91881 * cr + 32768) >> 16));
        o[4 * x + 1] = clamp8(yy + ((-22554 * cb + 32768 - 46802 * cr) >> 16));
        o[4 * x + 2] = clamp8(yy + ((116130 * cb + 32768) >> 16));
        o[4 * x + 3] = 255;
      }
    }
  }
  for (int c = 0; c < 3; ++c) { free(coef[c]); free(plane[c]); }
  return rc;
}

/* Quantised coefficients for GPU back-end tests: component c, blocks row-major over the padded plane. */
int jpeg_ref_coefficients(const uint8_t* file, size_t n, int16_t* c0, int16_t* c1, int16_t* c2) {
  JDec d;
  int rc = parse(&d, file, n);
  if (rc) return rc;
  int16_t* coef[3] = {c0, c1, c2};
  return d.progressive ? decode_progressive(&d, coef) : decode_scan(&d, coef);   /* progressive: caller zeroes the planes */
}
