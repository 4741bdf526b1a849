/* TEST INFRASTRUCTURE (oracle) -- not part of the product; only tests/, __graft_entry__.smoke() and
 * bench.py's CPU legs may load this.
 *
 * Sequential CPU restatement of the export step for fileType 'jpg'
 * (utils/canvas.js:205-242 -> wx.canvasToTempFilePath; the encoder itself lives in the closed WeChat
 * runtime, so the byte format is pinned to its published standard instead: ITU-T T.81 baseline
 * sequential DCT, 8-bit, YCbCr 4:2:0 (H2V2), JFIF 1.01 container, Annex K tables, IJG quality
 * scaling, restart markers every `restart_interval` MCUs).  Pinned by decoding its output with an
 * independent decoder (libjpeg inside PIL / OpenCV) in tests/test_jpeg.py.
 *
 * The forward DCT is the Loeffler-Ligtenberg-Moschytz integer algorithm (13-bit constants, two passes,
 * output scaled by 8); all arithmetic is int32, so the CUDA kernels reproduce it bit for bit.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "jpeg_tables.inc"

#define CB 13
#define P1 2
#define DESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))
#define F_0_298 2446
#define F_0_390 3196
#define F_0_541 4433
#define F_0_765 6270
#define F_0_899 7373
#define F_1_175 9633
#define F_1_501 12299
#define F_1_847 15137
#define F_1_961 16069
#define F_2_053 16819
#define F_2_562 20995
#define F_3_072 25172

/* one 8-point pass over d[0..7*stride]; first = row pass (scale up by 2^P1), else column pass */
static void fdct_1d(int32_t* d, int stride, int first) {
  int32_t t0 = d[0] + d[7 * stride], t7 = d[0] - d[7 * stride];
  int32_t t1 = d[stride] + d[6 * stride], t6 = d[stride] - d[6 * stride];
  int32_t t2 = d[2 * stride] + d[5 * stride], t5 = d[2 * stride] - d[5 * stride];
  int32_t t3 = d[3 * stride] + d[4 * stride], t4 = d[3 * stride] - d[4 * stride];
  int32_t t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
  const int sh = first ? CB - P1 : CB + P1;
  if (first) {
    d[0] = (t10 + t11) << P1;
    d[4 * stride] = (t10 - t11) << P1;
  } else {
    d[0] = DESCALE(t10 + t11, P1);
    d[4 * stride] = DESCALE(t10 - t11, P1);
  }
  int32_t z1 = (t12 + t13) * F_0_541;
  d[2 * stride] = DESCALE(z1 + t13 * F_0_765, sh);
  d[6 * stride] = DESCALE(z1 - t12 * F_1_847, sh);
  z1 = t4 + t7;
  int32_t z2 = t5 + t6, z3 = t4 + t6, z4 = t5 + t7;
  int32_t z5 = (z3 + z4) * F_1_175;
  t4 *= F_0_298;
  t5 *= F_2_053;
  t6 *= F_3_072;
  t7 *= F_1_501;
  z1 *= -F_0_899;
  z2 *= -F_2_562;
  z3 = z3 * -F_1_961 + z5;
  z4 = z4 * -F_0_390 + z5;
  d[7 * stride] = DESCALE(t4 + z1 + z3, sh);
  d[5 * stride] = DESCALE(t5 + z2 + z4, sh);
  d[3 * stride] = DESCALE(t6 + z2 + z3, sh);
  d[stride] = DESCALE(t7 + z1 + z4, sh);
}

/* IJG quality scaling of an Annex K table (natural order), 1 <= quality <= 100 */
void jpeg_ref_quant_table(int chroma, int quality, uint8_t* out64) {
  if (quality < 1) quality = 1;
  if (quality > 100) quality = 100;
  const int scale = quality < 50 ? 5000 / quality : 200 - 2 * quality;
  const unsigned char* base = chroma ? JQ_CHROMA : JQ_LUMA;
  for (int i = 0; i < 64; ++i) {
    int v = (base[i] * scale + 50) / 100;
    out64[i] = (uint8_t)(v < 1 ? 1 : v > 255 ? 255 : v);
  }
}

/* 8x8 block at (bx*8, by*8) of a plane (edge replication) -> 64 quantised coefficients, zig-zag order */
void jpeg_ref_block(const uint8_t* plane, int pw, int ph, int bx, int by, const uint8_t* q, int16_t* zz) {
  int32_t d[64];
  for (int r = 0; r < 8; ++r) {
    int yy = by * 8 + r;
    if (yy > ph - 1) yy = ph - 1;
    for (int c = 0; c < 8; ++c) {
      int xx = bx * 8 + c;
      if (xx > pw - 1) xx = pw - 1;
      d[r * 8 + c] = (int32_t)plane[(size_t)yy * pw + xx] - 128;
    }
  }
  for (int r = 0; r < 8; ++r) fdct_1d(d + r * 8, 1, 1);
  for (int c = 0; c < 8; ++c) fdct_1d(d + c, 8, 0);
  for (int k = 0; k < 64; ++k) {
    const int i = JZIGZAG[k];
    int32_t v = d[i];
    const int32_t qv = (int32_t)q[i] * 8;
    int neg = v < 0;
    if (neg) v = -v;
    v += qv >> 1;
    v = v >= qv ? v / qv : 0;
    zz[k] = (int16_t)(neg ? -v : v);
  }
}

typedef struct {
  uint16_t code[256];
  uint8_t len[256];
} HuffTab;

static void build_huff(const unsigned char* bits, const unsigned char* val, HuffTab* t) {
  memset(t, 0, sizeof *t);
  int code = 0, k = 0;
  for (int l = 1; l <= 16; ++l) {
    for (int i = 0; i < bits[l - 1]; ++i) {
      t->code[val[k]] = (uint16_t)code;
      t->len[val[k]] = (uint8_t)l;
      ++code;
      ++k;
    }
    code <<= 1;
  }
}

/* code/len tables in the layout the product uses: [4][256] = DC luma, DC chroma, AC luma, AC chroma */
void jpeg_ref_huff_tables(uint16_t* code, uint8_t* len) {
  HuffTab t[4];
  build_huff(JDC_LUMA_BITS, JDC_LUMA_VAL, &t[0]);
  build_huff(JDC_CHROMA_BITS, JDC_CHROMA_VAL, &t[1]);
  build_huff(JAC_LUMA_BITS, JAC_LUMA_VAL, &t[2]);
  build_huff(JAC_CHROMA_BITS, JAC_CHROMA_VAL, &t[3]);
  for (int i = 0; i < 4; ++i) {
    memcpy(code + 256 * i, t[i].code, sizeof t[i].code);
    memcpy(len + 256 * i, t[i].len, sizeof t[i].len);
  }
}

typedef struct {
  uint8_t* p;
  size_t n, cap;
  uint32_t acc;
  int nbits;
  int overflow;
} BitW;

static void put_byte(BitW* w, int b) {
  if (w->n < w->cap) w->p[w->n] = (uint8_t)b;
  else w->overflow = 1;
  w->n++;
}
static void put_bits(BitW* w, uint32_t v, int n) { /* n <= 16 */
  w->acc = (w->acc << n) | (v & ((1u << n) - 1));
  w->nbits += n;
  while (w->nbits >= 8) {
    const int b = (w->acc >> (w->nbits - 8)) & 0xff;
    put_byte(w, b);
    if (b == 0xff) put_byte(w, 0);
    w->nbits -= 8;
  }
}
static void flush_bits(BitW* w) { /* pad with 1-bits to a byte boundary */
  if (w->nbits > 0) put_bits(w, 0x7f, 8 - w->nbits);
  w->acc = 0;
  w->nbits = 0;
}
static int bit_size(int v) {
  int s = 0;
  if (v < 0) v = -v;
  while (v) {
    ++s;
    v >>= 1;
  }
  return s;
}
static void encode_block(BitW* w, const int16_t* zz, int* pred, const HuffTab* dc, const HuffTab* ac) {
  int diff = zz[0] - *pred;
  *pred = zz[0];
  int s = bit_size(diff);
  put_bits(w, dc->code[s], dc->len[s]);
  if (s) put_bits(w, (uint32_t)(diff < 0 ? diff - 1 : diff), s);
  int run = 0;
  for (int k = 1; k < 64; ++k) {
    const int v = zz[k];
    if (v == 0) {
      ++run;
      continue;
    }
    while (run > 15) {
      put_bits(w, ac->code[0xf0], ac->len[0xf0]);
      run -= 16;
    }
    s = bit_size(v);
    put_bits(w, ac->code[(run << 4) | s], ac->len[(run << 4) | s]);
    put_bits(w, (uint32_t)(v < 0 ? v - 1 : v), s);
    run = 0;
  }
  if (run > 0) put_bits(w, ac->code[0], ac->len[0]);
}

static void put_marker(BitW* w, int m) {
  put_byte(w, 0xff);
  put_byte(w, m);
}
static void put_u16(BitW* w, int v) {
  put_byte(w, (v >> 8) & 0xff);
  put_byte(w, v & 0xff);
}
static void put_dht(BitW* w, int tc_th, const unsigned char* bits, const unsigned char* val) {
  int n = 0;
  for (int i = 0; i < 16; ++i) n += bits[i];
  put_marker(w, 0xc4);
  put_u16(w, 2 + 1 + 16 + n);
  put_byte(w, tc_th);
  for (int i = 0; i < 16; ++i) put_byte(w, bits[i]);
  for (int i = 0; i < n; ++i) put_byte(w, val[i]);
}

/* Everything before the entropy-coded data: SOI APP0 DQT DQT SOF0 DHT x4 DRI SOS.  Returns bytes. */
size_t jpeg_ref_header(int w, int h, int quality, int restart_interval, uint8_t* out, size_t cap) {
  BitW bw = {out, 0, cap, 0, 0, 0};
  uint8_t q[2][64];
  jpeg_ref_quant_table(0, quality, q[0]);
  jpeg_ref_quant_table(1, quality, q[1]);
  put_marker(&bw, 0xd8);
  put_marker(&bw, 0xe0);
  put_u16(&bw, 16);
  put_byte(&bw, 'J'); put_byte(&bw, 'F'); put_byte(&bw, 'I'); put_byte(&bw, 'F'); put_byte(&bw, 0);
  put_byte(&bw, 1); put_byte(&bw, 1);      /* version 1.01 */
  put_byte(&bw, 0);                          /* aspect-ratio units */
  put_u16(&bw, 1); put_u16(&bw, 1);
  put_byte(&bw, 0); put_byte(&bw, 0);      /* no thumbnail */
  for (int t = 0; t < 2; ++t) {
    put_marker(&bw, 0xdb);
    put_u16(&bw, 67);
    put_byte(&bw, t);
    for (int k = 0; k < 64; ++k) put_byte(&bw, q[t][JZIGZAG[k]]);
  }
  put_marker(&bw, 0xc0);
  put_u16(&bw, 17);
  put_byte(&bw, 8);
  put_u16(&bw, h);
  put_u16(&bw, w);
  put_byte(&bw, 3);
  put_byte(&bw, 1); put_byte(&bw, 0x22); put_byte(&bw, 0);
  put_byte(&bw, 2); put_byte(&bw, 0x11); put_byte(&bw, 1);
  put_byte(&bw, 3); put_byte(&bw, 0x11); put_byte(&bw, 1);
  put_dht(&bw, 0x00, JDC_LUMA_BITS, JDC_LUMA_VAL);
  put_dht(&bw, 0x10, JAC_LUMA_BITS, JAC_LUMA_VAL);
  put_dht(&bw, 0x01, JDC_CHROMA_BITS, JDC_CHROMA_VAL);
  put_dht(&bw, 0x11, JAC_CHROMA_BITS, JAC_CHROMA_VAL);
  if (restart_interval > 0) {
    put_marker(&bw, 0xdd);
    put_u16(&bw, 4);
    put_u16(&bw, restart_interval);
  }
  put_marker(&bw, 0xda);
  put_u16(&bw, 12);
  put_byte(&bw, 3);
  put_byte(&bw, 1); put_byte(&bw, 0x00);
  put_byte(&bw, 2); put_byte(&bw, 0x11);
  put_byte(&bw, 3); put_byte(&bw, 0x11);
  put_byte(&bw, 0); put_byte(&bw, 63); put_byte(&bw, 0);
  return bw.overflow ? 0 : bw.n;
}

/* I420 planes (Y w x h; Cb, Cr ceil(w/2) x ceil(h/2)) -> JFIF file.  Returns 0, or 1 when `cap` is too small
 * (*out_len still holds the size needed). */
int jpeg_ref_encode_i420(const uint8_t* y, const uint8_t* cb, const uint8_t* cr, int w, int h, int quality,
                         int restart_interval, uint8_t* out, size_t cap, size_t* out_len) {
  BitW bw = {out, 0, cap, 0, 0, 0};
  bw.n = jpeg_ref_header(w, h, quality, restart_interval, out, cap);
  if (bw.n == 0) {
    uint8_t tmp[1024];
    bw.n = jpeg_ref_header(w, h, quality, restart_interval, tmp, sizeof tmp);
    bw.overflow = 1;
  }
  uint8_t q[2][64];
  jpeg_ref_quant_table(0, quality, q[0]);
  jpeg_ref_quant_table(1, quality, q[1]);
  HuffTab dcl, dcc, acl, acc_;
  build_huff(JDC_LUMA_BITS, JDC_LUMA_VAL, &dcl);
  build_huff(JDC_CHROMA_BITS, JDC_CHROMA_VAL, &dcc);
  build_huff(JAC_LUMA_BITS, JAC_LUMA_VAL, &acl);
  build_huff(JAC_CHROMA_BITS, JAC_CHROMA_VAL, &acc_);
  const int cw = (w + 1) / 2, chh = (h + 1) / 2;
  const int mx = (w + 15) / 16, my = (h + 15) / 16;
  int pred[3] = {0, 0, 0};
  int16_t zz[64];
  int count = 0, rst = 0;
  for (int m = 0; m < mx * my; ++m) {
    if (restart_interval > 0 && count == restart_interval) {
      flush_bits(&bw);
      put_marker(&bw, 0xd0 + (rst & 7));
      ++rst;
      count = 0;
      pred[0] = pred[1] = pred[2] = 0;
    }
    const int mcx = m % mx, mcy = m / mx;
    for (int b = 0; b < 4; ++b) {
      jpeg_ref_block(y, w, h, mcx * 2 + (b & 1), mcy * 2 + (b >> 1), q[0], zz);
      encode_block(&bw, zz, &pred[0], &dcl, &acl);
    }
    jpeg_ref_block(cb, cw, chh, mcx, mcy, q[1], zz);
    encode_block(&bw, zz, &pred[1], &dcc, &acc_);
    jpeg_ref_block(cr, cw, chh, mcx, mcy, q[1], zz);
    encode_block(&bw, zz, &pred[2], &dcc, &acc_);
    ++count;
  }
  flush_bits(&bw);
  put_marker(&bw, 0xd9);
  *out_len = bw.n;
  return bw.overflow;
}
