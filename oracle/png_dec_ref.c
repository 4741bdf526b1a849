/* TEST INFRASTRUCTURE (oracle): sequential CPU restatement of PNG decoding -- what loadImageFrom
 * (utils/canvas.js:27-121, called at pages/index/index.js:1441-1463) gets from the platform's libpng + zlib, which are
 * not part of /root/reference.  Restated from the published formats: PNG (ISO/IEC 15948: chunks, filters 0-4,
 * colour types, tRNS), zlib (RFC 1950), DEFLATE (RFC 1951).  One byte at a time, one bit at a time: the plainest
 * possible decoder, nothing shared with the GPU path (imagestitching_b200/csrc/png_*).
 * Pinned by tests/test_png_decode.py against Python's zlib (inflate) and PIL / libpng (bitmaps) on every colour type,
 * bit depth and filter.  16-bit samples keep their high byte (libpng's png_set_strip_16, PIL's RGB;16B unpacker).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may call this. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  const uint8_t* in;
  size_t in_len, in_pos;
  uint32_t bitbuf;
  int bitcnt;
  uint8_t* out;
  size_t out_len, out_pos;
  int err;
} Inf;

static int inf_bits(Inf* s, int need) {
  uint32_t val = s->bitbuf;
  while (s->bitcnt < need) {
    if (s->in_pos >= s->in_len) {
      s->err = 1;
      return 0;
    }
    val |= (uint32_t)s->in[s->in_pos++] << s->bitcnt;
    s->bitcnt += 8;
  }
  s->bitbuf = val >> need;
  s->bitcnt -= need;
  return (int)(val & ((1u << need) - 1u));
}

typedef struct {
  short count[16];
  short symbol[288];
} Huff;

/* returns 0 complete, > 0 incomplete, < 0 over-subscribed */
static int huff_build(Huff* h, const short* length, int n) {
  short offs[16];
  int left = 1;
  for (int len = 0; len <= 15; ++len) h->count[len] = 0;
  for (int s = 0; s < n; ++s) h->count[length[s]]++;
  if (h->count[0] == n) return 0;
  for (int len = 1; len <= 15; ++len) {
    left <<= 1;
    left -= h->count[len];
    if (left < 0) return left;
  }
  offs[1] = 0;
  for (int len = 1; len < 15; ++len) offs[len + 1] = (short)(offs[len] + h->count[len]);
  for (int s = 0; s < n; ++s)
    if (length[s]) h->symbol[offs[length[s]]++] = (short)s;
  return left;
}

static int huff_decode(Inf* s, const Huff* h) {
  int code = 0, first = 0, index = 0;
  for (int len = 1; len <= 15; ++len) {
    code |= inf_bits(s, 1);
    if (s->err) return -1;
    const int count = h->count[len];
    if (code - count < first) return h->symbol[index + (code - first)];
    index += count;
    first += count;
    first <<= 1;
    code <<= 1;
  }
  return -1;
}

static const short kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static const short kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
static const short kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
static const short kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

static int inf_codes(Inf* s, const Huff* lencode, const Huff* distcode) {
  for (;;) {
    int sym = huff_decode(s, lencode);
    if (sym < 0) return 1;
    if (sym < 256) {
      if (s->out_pos >= s->out_len) return 2;
      s->out[s->out_pos++] = (uint8_t)sym;
    } else if (sym == 256) {
      return 0;
    } else {
      sym -= 257;
      if (sym >= 29) return 3;
      const int len = kLenBase[sym] + inf_bits(s, kLenExtra[sym]);
      const int ds = huff_decode(s, distcode);
      if (ds < 0 || ds >= 30) return 4;
      const size_t dist = (size_t)kDistBase[ds] + (size_t)inf_bits(s, kDistExtra[ds]);
      if (s->err) return 5;
      if (dist > s->out_pos) return 6;
      if (s->out_pos + (size_t)len > s->out_len) return 2;
      for (int k = 0; k < len; ++k, ++s->out_pos) s->out[s->out_pos] = s->out[s->out_pos - dist];
    }
  }
}

/* zlib stream -> out[0 .. out_len); returns 0 when the stream is valid and makes exactly out_len bytes */
int png_dec_ref_inflate(const uint8_t* z, size_t z_len, uint8_t* out, size_t out_len) {
  if (z_len < 2 || (z[0] & 0x0F) != 8 || (z[0] >> 4) > 7 || ((z[0] << 8) | z[1]) % 31 != 0 || (z[1] & 0x20)) return 10;
  Inf s;
  memset(&s, 0, sizeof s);
  s.in = z + 2;
  s.in_len = z_len - 2;
  s.out = out;
  s.out_len = out_len;
  int last;
  do {
    last = inf_bits(&s, 1);
    const int type = inf_bits(&s, 2);
    if (s.err) return 11;
    if (type == 0) {
      s.bitbuf = 0;
      s.bitcnt = 0;
      if (s.in_pos + 4 > s.in_len) return 12;
      const unsigned len = s.in[s.in_pos] | ((unsigned)s.in[s.in_pos + 1] << 8);
      const unsigned nlen = s.in[s.in_pos + 2] | ((unsigned)s.in[s.in_pos + 3] << 8);
      s.in_pos += 4;
      if ((len ^ 0xFFFFu) != nlen) return 13;
      if (s.in_pos + len > s.in_len || s.out_pos + len > s.out_len) return 14;
      memcpy(s.out + s.out_pos, s.in + s.in_pos, len);
      s.in_pos += len;
      s.out_pos += len;
    } else if (type == 1) {
      short lengths[288 + 30];
      Huff lc, dc;
      int sym = 0;
      for (; sym < 144; ++sym) lengths[sym] = 8;
      for (; sym < 256; ++sym) lengths[sym] = 9;
      for (; sym < 280; ++sym) lengths[sym] = 7;
      for (; sym < 288; ++sym) lengths[sym] = 8;
      huff_build(&lc, lengths, 288);
      for (sym = 0; sym < 30; ++sym) lengths[sym] = 5;
      huff_build(&dc, lengths, 30);
      const int rc = inf_codes(&s, &lc, &dc);
      if (rc) return 20 + rc;
    } else if (type == 2) {
      static const short order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
      short lengths[320];
      Huff lc, dc;
      const int nlen = inf_bits(&s, 5) + 257, ndist = inf_bits(&s, 5) + 1, ncode = inf_bits(&s, 4) + 4;
      if (s.err || nlen > 286 || ndist > 30) return 30;
      int i = 0;
      for (; i < ncode; ++i) lengths[order[i]] = (short)inf_bits(&s, 3);
      for (; i < 19; ++i) lengths[order[i]] = 0;
      if (huff_build(&lc, lengths, 19) != 0) return 31;
      i = 0;
      while (i < nlen + ndist) {
        int sym = huff_decode(&s, &lc);
        if (sym < 0) return 32;
        if (sym < 16) {
          lengths[i++] = (short)sym;
        } else {
          int len = 0, rep;
          if (sym == 16) {
            if (i == 0) return 33;
            len = lengths[i - 1];
            rep = 3 + inf_bits(&s, 2);
          } else if (sym == 17) {
            rep = 3 + inf_bits(&s, 3);
          } else {
            rep = 11 + inf_bits(&s, 7);
          }
          if (i + rep > nlen + ndist) return 34;
          while (rep--) lengths[i++] = (short)len;
        }
      }
      if (s.err) return 35;
      if (lengths[256] == 0) return 36;
      int rc = huff_build(&lc, lengths, nlen);
      if (rc < 0 || (rc > 0 && !(nlen - lc.count[0] == 1 && lc.count[1] == 1))) return 37;   /* zlib: incomplete only as one 1-bit code */
      rc = huff_build(&dc, lengths + nlen, ndist);
      if (rc < 0 || (rc > 0 && !(ndist - dc.count[0] == 1 && dc.count[1] == 1))) return 38;
      rc = inf_codes(&s, &lc, &dc);
      if (rc) return 40 + rc;
    } else {
      return 15;
    }
  } while (!last);
  return s.out_pos == s.out_len ? 0 : 16;
}

static int paeth(int a, int b, int c) {
  const int p = a + b - c;
  const int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

/* raw = h x (filter byte + row_bytes), unfiltered in place (ISO 15948 clause 9); bpp = filter unit in bytes */
int png_dec_ref_unfilter(uint8_t* raw, int h, size_t row_bytes, int bpp) {
  for (int y = 0; y < h; ++y) {
    uint8_t* row = raw + (size_t)y * (row_bytes + 1) + 1;
    const uint8_t* up = y ? row - (row_bytes + 1) : NULL;
    const int ft = row[-1];
    if (ft > 4) return 1;
    for (size_t i = 0; i < row_bytes; ++i) {
      const int a = i >= (size_t)bpp ? row[i - bpp] : 0;
      const int b = up ? up[i] : 0;
      const int c = (up && i >= (size_t)bpp) ? up[i - bpp] : 0;
      int pred = 0;
      switch (ft) {
        case 1: pred = a; break;
        case 2: pred = b; break;
        case 3: pred = (a + b) >> 1; break;
        case 4: pred = paeth(a, b, c); break;
        default: break;
      }
      row[i] = (uint8_t)(row[i] + pred);
    }
  }
  return 0;
}

/* unfiltered rows -> RGBA8 (straight alpha).  palette: 256 x RGBA (alpha from tRNS); key: tRNS colour key */
void png_dec_ref_to_rgba(const uint8_t* raw, int w, int h, size_t row_bytes, int depth, int ctype, const uint8_t* palette, int has_key,
                         const uint16_t* key, uint8_t* rgba) {
  const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;
  for (int y = 0; y < h; ++y) {
    const uint8_t* row = raw + (size_t)y * (row_bytes + 1) + 1;
    uint8_t* o = rgba + (size_t)y * (size_t)w * 4;
    for (int x = 0; x < w; ++x, o += 4) {
      unsigned smp[4] = {0, 0, 0, 0}, full[4] = {0, 0, 0, 0};
      if (depth < 8) {
        const unsigned bit = (unsigned)x * (unsigned)depth;
        full[0] = (row[bit >> 3] >> (8 - depth - (int)(bit & 7))) & ((1u << depth) - 1u);
        smp[0] = ctype == 3 ? full[0] : full[0] * (255u / ((1u << depth) - 1u));
      } else {
        const int sb = depth / 8;
        for (int c = 0; c < channels; ++c) {
          const uint8_t* p = row + ((size_t)x * channels + c) * sb;
          smp[c] = p[0];
          full[c] = sb == 2 ? ((unsigned)p[0] << 8) | p[1] : p[0];
        }
      }
      unsigned r, g, b, a = 255;
      switch (ctype) {
        case 0:
          r = g = b = smp[0];
          if (has_key && full[0] == key[0]) a = 0;
          break;
        case 2:
          r = smp[0]; g = smp[1]; b = smp[2];
          if (has_key && full[0] == key[0] && full[1] == key[1] && full[2] == key[2]) a = 0;
          break;
        case 3:
          r = palette[4 * smp[0]]; g = palette[4 * smp[0] + 1]; b = palette[4 * smp[0] + 2]; a = palette[4 * smp[0] + 3];
          break;
        case 4:
          r = g = b = smp[0];
          a = smp[1];
          break;
        default:
          r = smp[0]; g = smp[1]; b = smp[2]; a = smp[3];
          break;
      }
      o[0] = (uint8_t)r; o[1] = (uint8_t)g; o[2] = (uint8_t)b; o[3] = (uint8_t)a;
    }
  }
}
