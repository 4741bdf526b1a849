// Constant tables of the baseline JPEG export (ITU-T T.81 Annex K: K.1/K.2 quantisation, K.3-K.6 Huffman;
// zig-zag order of Figure A.6) and the host-side header writer.  The export step it serves is
// utils/canvas.js:205-242 (wx.canvasToTempFilePath with fileType 'jpg', quality 1).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

namespace stitch {
namespace jpeg {

static const uint8_t JQ_LUMA[64] = {
    16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62,
    18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
static const uint8_t JQ_CHROMA[64] = {
    17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
/* zig-zag position k -> natural (row-major) index */
static const uint8_t JZIGZAG[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
/* Huffman tables: BITS (number of codes of length 1..16) and HUFFVAL */
static const uint8_t JDC_LUMA_BITS[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
static const uint8_t JDC_LUMA_VAL[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const uint8_t JDC_CHROMA_BITS[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
static const uint8_t JDC_CHROMA_VAL[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const uint8_t JAC_LUMA_BITS[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
static const uint8_t JAC_LUMA_VAL[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14, 0x32, 0x81,
    0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16, 0x17, 0x18,
    0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48,
    0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75,
    0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99,
    0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3,
    0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5,
    0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
static const uint8_t JAC_CHROMA_BITS[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
static const uint8_t JAC_CHROMA_VAL[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32, 0x81, 0x08,
    0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25,
    0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47,
    0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74,
    0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97,
    0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba,
    0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4,
    0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};

// IJG quality scaling of an Annex K table (natural order), 1 <= quality <= 100.
inline void quant_table(bool chroma, int quality, uint8_t* out64) {
  quality = quality < 1 ? 1 : quality > 100 ? 100 : quality;
  const int scale = quality < 50 ? 5000 / quality : 200 - 2 * quality;
  const uint8_t* base = chroma ? JQ_CHROMA : JQ_LUMA;
  for (int i = 0; i < 64; ++i) {
    const int v = (base[i] * scale + 50) / 100;
    out64[i] = (uint8_t)(v < 1 ? 1 : v > 255 ? 255 : v);
  }
}

// Canonical Huffman codes of one (BITS, HUFFVAL) table: entry = code | len << 16, 0 for unused symbols.
inline void huff_table(const uint8_t* bits, const uint8_t* val, uint32_t* out256) {
  std::memset(out256, 0, 256 * sizeof(uint32_t));
  uint32_t code = 0;
  int k = 0;
  for (int l = 1; l <= 16; ++l) {
    for (int i = 0; i < bits[l - 1]; ++i) out256[val[k++]] = code++ | ((uint32_t)l << 16);
    code <<= 1;
  }
}

// Device-side constants of one encode: quantisation divisors (q * 8, natural order) and the four code tables.
struct Tables {
  uint16_t qdiv[2][64];       // [luma, chroma][natural index]: divisor q * 8 (the DCT output is scaled by 8)
  uint32_t qmagic[2][64];     // ceil(2^32 / qdiv): floor(n / qdiv) == umulhi(n, qmagic) for n < 2^21
  uint32_t huff[4][256];      // DC luma, DC chroma, AC luma, AC chroma: code | len << 16
};

inline void build_tables(int quality, Tables* t) {
  uint8_t q[64];
  for (int c = 0; c < 2; ++c) {
    quant_table(c != 0, quality, q);
    for (int i = 0; i < 64; ++i) {
      t->qdiv[c][i] = (uint16_t)(q[i] * 8);
      t->qmagic[c][i] = (uint32_t)(((1ull << 32) + t->qdiv[c][i] - 1) / t->qdiv[c][i]);
    }
  }
  huff_table(JDC_LUMA_BITS, JDC_LUMA_VAL, t->huff[0]);
  huff_table(JDC_CHROMA_BITS, JDC_CHROMA_VAL, t->huff[1]);
  huff_table(JAC_LUMA_BITS, JAC_LUMA_VAL, t->huff[2]);
  huff_table(JAC_CHROMA_BITS, JAC_CHROMA_VAL, t->huff[3]);
}

// Everything before the entropy-coded data: SOI, APP0 (JFIF 1.01), DQT x2, SOF0 (8 bit, 3 components,
// Y 2x2 / Cb 1x1 / Cr 1x1), DHT x4, DRI, SOS.
inline std::vector<uint8_t> header(int w, int h, int quality, int restart_interval) {
  std::vector<uint8_t> o;
  auto b = [&](int v) { o.push_back((uint8_t)v); };
  auto u16 = [&](int v) { b(v >> 8); b(v & 0xff); };
  auto marker = [&](int m) { b(0xff); b(m); };
  marker(0xd8);
  marker(0xe0); u16(16);
  for (char c : {'J', 'F', 'I', 'F'}) b(c);
  b(0); b(1); b(1); b(0); u16(1); u16(1); b(0); b(0);
  uint8_t q[64];
  for (int t = 0; t < 2; ++t) {
    quant_table(t != 0, quality, q);
    marker(0xdb); u16(67); b(t);
    for (int k = 0; k < 64; ++k) b(q[JZIGZAG[k]]);
  }
  marker(0xc0); u16(17); b(8); u16(h); u16(w); b(3);
  b(1); b(0x22); b(0);
  b(2); b(0x11); b(1);
  b(3); b(0x11); b(1);
  auto dht = [&](int tc_th, const uint8_t* bits, const uint8_t* val) {
    int n = 0;
    for (int i = 0; i < 16; ++i) n += bits[i];
    marker(0xc4); u16(2 + 1 + 16 + n); b(tc_th);
    for (int i = 0; i < 16; ++i) b(bits[i]);
    for (int i = 0; i < n; ++i) b(val[i]);
  };
  dht(0x00, JDC_LUMA_BITS, JDC_LUMA_VAL);
  dht(0x10, JAC_LUMA_BITS, JAC_LUMA_VAL);
  dht(0x01, JDC_CHROMA_BITS, JDC_CHROMA_VAL);
  dht(0x11, JAC_CHROMA_BITS, JAC_CHROMA_VAL);
  if (restart_interval > 0) { marker(0xdd); u16(4); u16(restart_interval); }
  marker(0xda); u16(12); b(3);
  b(1); b(0x00); b(2); b(0x11); b(3); b(0x11);
  b(0); b(63); b(0);
  return o;
}

}  // namespace jpeg
}  // namespace stitch
