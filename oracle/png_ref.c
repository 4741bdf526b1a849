/* TEST INFRASTRUCTURE (oracle) -- sequential PNG writer, the checker of the GPU PNG export.
 *
 * The page exports its canvas with wx.canvasToTempFilePath({fileType: 'png', quality: 1})
 * (utils/canvas.js:205-242, called at pages/index/index.js:1577-1579); the encoder is inside the closed WeChat
 * runtime, so only the FORMAT is pinned: PNG (ISO/IEC 15948), colour type 6, 8 bit, non-interlaced, whose
 * IDAT data is a zlib stream (RFC 1950) of deflate blocks (RFC 1951).  Which filters and which deflate
 * blocks an encoder chooses is its own business; this file restates the choices of the product's GPU
 * encoder so the two files can be compared byte for byte, and tests/test_png.py pins it by decoding its
 * files with libpng/zlib (PIL): the pixels must come back exactly.
 *
 * Encoder choices: Paeth filter (type 4) on every row; the rows are cut into bands of PNG_BAND_ROWS rows,
 * each band is ONE dynamic-Huffman block followed -- except for the last band -- by an empty stored block that
 * byte-aligns the stream, and is its own IDAT chunk.  LZ77 is restricted to RUNS: the band's bytes are taken in
 * rounds of PNG_ROUND_PX pixels (with the filter-type byte in front of a row's first pixel); inside a round a
 * maximal run of equal bytes is coded as its first byte + matches of distance 1 (chunks of up to 258 bytes; a
 * chunk shorter than 3 stays literal).  That is what flat screenshot areas need -- after the Paeth filter they
 * are runs of zeros -- and it keeps every token a function of its round only.  Code lengths are Huffman
 * lengths limited to 15 bits by halving the frequencies until they fit; the code-length alphabet uses a fixed
 * 4-bit code for the lengths 0..15; the distance alphabet has the single code "distance 1".
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define PNG_BAND_ROWS 8
#define PNG_ROUND_PX 512
#define NSYM 286 /* 256 literals + end of block + 29 length codes */

static uint32_t crc_table[256];
static int crc_ready = 0;
static void crc_init(void) {
  for (uint32_t n = 0; n < 256; ++n) {
    uint32_t c = n;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
    crc_table[n] = c;
  }
  crc_ready = 1;
}
uint32_t png_ref_crc32(const uint8_t* p, size_t n) {
  if (!crc_ready) crc_init();
  uint32_t c = 0xFFFFFFFFu;
  for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ p[i]) & 255] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

static inline int paeth(int a, int b, int c) {
  const int p = a + b - c;
  const int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}
/* filtered byte i (0 .. 4w-1) of row y */
static inline uint8_t filt(const uint8_t* row, const uint8_t* prev, int i) {
  const int a = i >= 4 ? row[i - 4] : 0, b = prev ? prev[i] : 0, c = (prev && i >= 4) ? prev[i - 4] : 0;
  return (uint8_t)(row[i] - paeth(a, b, c));
}

/* Huffman code lengths (<= 15) for the symbols with freq > 0.  Tree: repeatedly merge the two nodes with the
 * smallest (frequency, index); leaves are 0..256, internal nodes 257.. in creation order.  Too deep -> halve
 * the frequencies (rounding up) and rebuild. */
void png_ref_code_lengths(const uint32_t* freq_in, uint8_t* len) {
  uint32_t freq[NSYM];
  memcpy(freq, freq_in, sizeof freq);
  for (;;) {
    uint32_t w[2 * NSYM];
    int parent[2 * NSYM], alive[2 * NSYM];
    int n = NSYM, n_alive = 0;
    for (int i = 0; i < NSYM; ++i) { w[i] = freq[i]; alive[i] = freq[i] > 0; n_alive += alive[i]; parent[i] = -1; }
    memset(len, 0, NSYM);
    if (n_alive == 1) {  /* cannot happen for a band (end-of-block + at least one literal), kept total */
      for (int i = 0; i < NSYM; ++i) if (alive[i]) len[i] = 1;
      return;
    }
    while (n_alive > 1) {
      int m1 = -1, m2 = -1;
      for (int i = 0; i < n; ++i) {
        if (!alive[i]) continue;
        if (m1 < 0 || w[i] < w[m1]) { m2 = m1; m1 = i; }
        else if (m2 < 0 || w[i] < w[m2]) m2 = i;
      }
      alive[m1] = alive[m2] = 0;
      w[n] = w[m1] + w[m2];
      parent[m1] = parent[m2] = n;
      parent[n] = -1;
      alive[n] = 1;
      ++n;
      --n_alive;
    }
    int depth[2 * NSYM], maxd = 0;
    depth[n - 1] = 0;
    for (int i = n - 2; i >= 0; --i) {
      if (parent[i] < 0) { depth[i] = 0; continue; }
      depth[i] = depth[parent[i]] + 1;
      if (i < NSYM && depth[i] > maxd) maxd = depth[i];
    }
    if (maxd <= 15) {
      for (int i = 0; i < NSYM; ++i) if (freq[i] > 0) len[i] = (uint8_t)depth[i];
      return;
    }
    for (int i = 0; i < NSYM; ++i) if (freq[i] > 0) freq[i] = (freq[i] + 1) >> 1;
  }
}

/* canonical codes (RFC 1951 3.2.2), bit-reversed for the LSB-first stream */
void png_ref_codes(const uint8_t* len, uint16_t* code) {
  int bl_count[16] = {0}, next[16];
  for (int i = 0; i < NSYM; ++i) bl_count[len[i]]++;
  bl_count[0] = 0;
  int c = 0;
  for (int b = 1; b <= 15; ++b) { c = (c + bl_count[b - 1]) << 1; next[b] = c; }
  for (int i = 0; i < NSYM; ++i) {
    if (!len[i]) { code[i] = 0; continue; }
    int v = next[len[i]]++, r = 0;
    for (int k = 0; k < len[i]; ++k) r |= ((v >> k) & 1) << (len[i] - 1 - k);
    code[i] = (uint16_t)r;
  }
}

/* length 3..258 -> literal/length symbol 257..285 and its extra bits (RFC 1951 3.2.5) */
static int len_code(int len, int* extra, int* nextra) {
  int c, e = 0, ne = 0;
  if (len < 11) c = 254 + len;
  else if (len == 258) c = 285;
  else {
    const int n = len - 3;
    int hb = 0;
    while ((n >> (hb + 1)) != 0) ++hb;          /* floor(log2(n)), n >= 8 */
    ne = hb - 2;
    c = 257 + 4 * ne + (n >> ne);
    e = n & ((1 << ne) - 1);
  }
  if (extra) *extra = e;
  if (nextra) *nextra = ne;
  return c;
}

typedef struct { uint8_t* p; size_t n, cap; uint64_t acc; int cnt; } BitW;
static void put(BitW* b, uint32_t v, int n) {
  b->acc |= (uint64_t)v << b->cnt;
  b->cnt += n;
  while (b->cnt >= 8) {
    if (b->n < b->cap) b->p[b->n] = (uint8_t)b->acc;
    b->n++;
    b->acc >>= 8;
    b->cnt -= 8;
  }
}
static void align(BitW* b) { if (b->cnt) put(b, 0, 8 - b->cnt); }
static void put_be32(uint8_t* p, uint32_t v) { p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v; }

/* Writes the PNG of an RGBA8 image.  Returns 0 and *need = file size; -1 if cap is too small (need still set). */
int png_ref_encode_rgba(const uint8_t* rgba, int w, int h, int64_t pitch, uint8_t* out, size_t cap, size_t* need) {
  if (!pitch) pitch = (int64_t)w * 4;
  const int rowb = w * 4, nb = (h + PNG_BAND_ROWS - 1) / PNG_BAND_ROWS;
  size_t o = 0;
  uint8_t hdr[33] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A, 0, 0, 0, 13, 'I', 'H', 'D', 'R'};
  put_be32(hdr + 16, (uint32_t)w);
  put_be32(hdr + 20, (uint32_t)h);
  hdr[24] = 8; hdr[25] = 6; hdr[26] = 0; hdr[27] = 0; hdr[28] = 0;
  put_be32(hdr + 29, png_ref_crc32(hdr + 12, 17));
  if (o + 33 <= cap) memcpy(out + o, hdr, 33);
  o += 33;
  uint32_t ad_a = 1, ad_b = 0;
  const size_t band_cap = (size_t)PNG_BAND_ROWS * (rowb + 1) * 2 + 1024;
  uint8_t* tmp = (uint8_t*)malloc(band_cap + 16);
  uint8_t* fb = (uint8_t*)malloc((size_t)PNG_BAND_ROWS * (rowb + 1));
  int* tok = (int*)malloc(sizeof(int) * (size_t)PNG_BAND_ROWS * (rowb + 1));
  if (!tmp || !fb || !tok) return -2;
  for (int band = 0; band < nb; ++band) {
    const int y0 = band * PNG_BAND_ROWS, y1 = y0 + PNG_BAND_ROWS < h ? y0 + PNG_BAND_ROWS : h;
    /* filtered bytes of the band, Adler, and the tokens: tok[i] = literal value (0..255), 0x1000 | length for a
     * match of distance 1 (3..258), or -1 for a byte covered by a match */
    uint32_t freq[NSYM];
    memset(freq, 0, sizeof freq);
    size_t nf = 0;
    for (int y = y0; y < y1; ++y) {
      const uint8_t* row = rgba + (int64_t)y * pitch;
      const uint8_t* prev = y > 0 ? rgba + (int64_t)(y - 1) * pitch : NULL;
      fb[nf++] = 4;
      for (int i = 0; i < rowb; ++i) fb[nf++] = filt(row, prev, i);
    }
    for (size_t i = 0; i < nf; ++i) {
      ad_a = (ad_a + fb[i]) % 65521u;
      ad_b = (ad_b + ad_a) % 65521u;
    }
    int n_match = 0;
    {
      const int npx = (y1 - y0) * w;
      size_t pos = 0;                           /* stream position of the round's first byte */
      for (int base = 0; base < npx; base += PNG_ROUND_PX) {
        const int q1 = base + PNG_ROUND_PX < npx ? base + PNG_ROUND_PX : npx;
        size_t n = 0;                           /* bytes of the round: 4 per pixel + 1 per row start */
        for (int q = base; q < q1; ++q) n += 4 + ((q % w) == 0);
        size_t i = 0;
        while (i < n) {
          size_t e = i + 1;
          while (e < n && fb[pos + e] == fb[pos + i]) ++e;
          const size_t L = e - i;               /* maximal run [i, e) */
          tok[pos + i] = fb[pos + i];
          for (size_t k = 1; k < L;) {
            const size_t clen = L - k < 258 ? L - k : 258;
            if (clen >= 3) {
              tok[pos + i + k] = 0x1000 | (int)clen;
              for (size_t m = 1; m < clen; ++m) tok[pos + i + k + m] = -1;
              ++n_match;
            } else {
              for (size_t m = 0; m < clen; ++m) tok[pos + i + k + m] = fb[pos + i];
            }
            k += clen;
          }
          i = e;
        }
        pos += n;
      }
    }
    for (size_t i = 0; i < nf; ++i) {
      if (tok[i] < 0) continue;
      if (tok[i] & 0x1000) freq[len_code(tok[i] & 0xfff, NULL, NULL)]++;
      else freq[tok[i]]++;
    }
    freq[256] = 1;
    uint8_t len[NSYM];
    uint16_t code[NSYM];
    png_ref_code_lengths(freq, len);
    png_ref_codes(len, code);
    /* chunk data = 'IDAT' + zlib bytes of this band */
    memcpy(tmp, "IDAT", 4);
    BitW b = {tmp + 4, 0, band_cap, 0, 0};
    const int last = band == nb - 1;
    if (band == 0) { put(&b, 0x78, 8); put(&b, 0x01, 8); }
    put(&b, (uint32_t)last, 1);
    put(&b, 2, 2);
    put(&b, NSYM - 257, 5);  /* HLIT: 286 codes */
    put(&b, 0, 5);           /* HDIST: 1 code */
    put(&b, 15, 4);          /* HCLEN: 19 code length codes */
    static const int order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    for (int i = 0; i < 19; ++i) put(&b, order[i] >= 16 ? 0 : 4, 3);
    for (int i = 0; i <= NSYM; ++i) {  /* 286 literal/length code lengths + the distance code: 1 bit if any match */
      const int s = i < NSYM ? len[i] : (n_match ? 1 : 0);
      const int r = ((s & 1) << 3) | ((s & 2) << 1) | ((s & 4) >> 1) | ((s & 8) >> 3);
      put(&b, (uint32_t)r, 4);
    }
    for (size_t i = 0; i < nf; ++i) {
      if (tok[i] < 0) continue;
      if (tok[i] & 0x1000) {
        int extra, nextra;
        const int c = len_code(tok[i] & 0xfff, &extra, &nextra);
        put(&b, code[c], len[c]);
        if (nextra) put(&b, (uint32_t)extra, nextra);
        put(&b, 0, 1);                          /* the single distance code (distance 1), canonical code 0 */
      } else {
        put(&b, code[tok[i]], len[tok[i]]);
      }
    }
    put(&b, code[256], len[256]);
    if (!last) {
      put(&b, 0, 3);  /* BFINAL = 0, BTYPE = 00 */
      align(&b);
      put(&b, 0x0000, 16);
      put(&b, 0xFFFF, 16);
    } else {
      align(&b);
      put(&b, (ad_b >> 8) & 255, 8); put(&b, ad_b & 255, 8); put(&b, (ad_a >> 8) & 255, 8); put(&b, ad_a & 255, 8);
    }
    const size_t dl = b.n;
    uint8_t lenb[4], crcb[4];
    put_be32(lenb, (uint32_t)dl);
    put_be32(crcb, png_ref_crc32(tmp, dl + 4));
    if (o + 4 <= cap) memcpy(out + o, lenb, 4);
    o += 4;
    if (o + dl + 4 <= cap) memcpy(out + o, tmp, dl + 4);
    o += dl + 4;
    if (o + 4 <= cap) memcpy(out + o, crcb, 4);
    o += 4;
  }
  static const uint8_t iend[12] = {0, 0, 0, 0, 'I', 'E', 'N', 'D', 0xAE, 0x42, 0x60, 0x82};
  if (o + 12 <= cap) memcpy(out + o, iend, 12);
  o += 12;
  free(tmp);
  free(fb);
  free(tok);
  *need = o;
  return o <= cap ? 0 : -1;
}
