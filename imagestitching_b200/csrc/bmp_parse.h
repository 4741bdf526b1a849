// Host-side header parser of the BMP decoder (bmp_decode.cuh): BITMAPFILEHEADER + BITMAPCOREHEADER / BITMAPINFOHEADER
// and its V2-V5 extensions, palette, bit fields.  Plain C++ (tests/native/parse_fuzz.cpp runs it under AddressSanitizer
// on damaged files).  Replaces the first half of loadImageFrom (U:27-121) for 'bmp' (P:4).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>

namespace stitch {
namespace bmpdec {

constexpr int kMaxImages = 9;

struct Header {
  int32_t w = 0, h = 0, bpp = 0;
  bool top_down = false;
  uint32_t data_off = 0, row_stride = 0;
  int32_t shift[4] = {0, 0, 0, 0}, bits[4] = {0, 0, 0, 0};   // R, G, B, A fields of 16 / 32 bit pixels (bits 0 = absent)
  uint32_t palette[256];                                      // RGBA8, alpha 255
};

enum ParseStatus { kOk = 0, kNotBmp = 1, kCorrupt = 2, kUnsupported = 3 };

inline uint32_t le16(const uint8_t* p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8; }
inline uint32_t le32(const uint8_t* p) { return le16(p) | le16(p + 2) << 16; }

inline bool field_of(uint32_t mask, int32_t* shift, int32_t* bits) {
  *shift = 0;
  *bits = 0;
  if (!mask) return true;
  while (!((mask >> *shift) & 1u)) ++*shift;
  while (*shift + *bits < 32 && ((mask >> (*shift + *bits)) & 1u)) ++*bits;
  return (*shift + *bits >= 32 ? 0u : mask >> (*shift + *bits)) == 0u;   // the set bits are contiguous
}

inline ParseStatus parse(const uint8_t* f, size_t n, Header* H) {
  if (n < 26 || f[0] != 'B' || f[1] != 'M') return kNotBmp;
  const uint32_t off_bits = le32(f + 10), hs = le32(f + 14);
  const uint8_t* d = f + 14;
  uint32_t compression = 0, colours = 0;
  int pal_entry = 4;
  if (hs == 12) {
    H->w = (int32_t)le16(d + 4);
    H->h = (int32_t)le16(d + 6);
    H->bpp = (int32_t)le16(d + 10);
    pal_entry = 3;
  } else if (hs == 40 || hs == 52 || hs == 56 || hs == 64 || hs == 108 || hs == 124) {
    if (n < 14 + (size_t)hs) return kCorrupt;
    H->w = (int32_t)le32(d + 4);
    const int32_t hh = (int32_t)le32(d + 8);
    H->top_down = hh < 0;
    H->h = hh < 0 ? -hh : hh;
    H->bpp = (int32_t)le16(d + 14);
    compression = le32(d + 16);
    colours = le32(d + 32);
  } else {
    return kUnsupported;
  }
  if (H->w <= 0 || H->h <= 0 || H->w > 65535 * 16 || H->h > 65535 * 16) return kCorrupt;
  const int bpp = H->bpp;
  if (bpp != 1 && bpp != 4 && bpp != 8 && bpp != 16 && bpp != 24 && bpp != 32) return kUnsupported;
  if (compression == 1 || compression == 2) return kUnsupported;          // RLE8 / RLE4
  if (compression != 0 && compression != 3) return kUnsupported;          // JPEG / PNG payloads, alpha bit fields of CE
  uint32_t masks[4] = {0, 0, 0, 0};
  if (compression == 3) {
    if (bpp != 16 && bpp != 32) return kUnsupported;
    const uint8_t* m = d + 40;                                             // inside a V2+ header, behind a 40-byte one
    const int nm = hs >= 56 ? 4 : 3;
    if (n < (size_t)(14 + 40 + 4 * nm)) return kCorrupt;
    for (int k = 0; k < nm; ++k) masks[k] = le32(m + 4 * k);
    if (bpp == 32 && !masks[0] && !masks[1] && !masks[2] && !masks[3]) {   // no masks at all: B, G, R, A bytes
      masks[0] = 0xFF0000u; masks[1] = 0xFF00u; masks[2] = 0xFFu; masks[3] = 0xFF000000u;
    }
  } else if (bpp == 16) {
    masks[0] = 0x7C00u; masks[1] = 0x3E0u; masks[2] = 0x1Fu;               // 5-5-5
  } else if (bpp == 32) {
    masks[0] = 0xFF0000u; masks[1] = 0xFF00u; masks[2] = 0xFFu;            // B, G, R, unused
  }
  for (int k = 0; k < 4; ++k)
    if (!field_of(masks[k], &H->shift[k], &H->bits[k])) return kUnsupported;
  if (bpp <= 8) {
    if (!colours || colours > (1u << bpp)) colours = 1u << bpp;
    const size_t pal_off = 14 + (size_t)hs;
    if (pal_off + (size_t)colours * pal_entry > n) return kCorrupt;
    for (int i = 0; i < 256; ++i) H->palette[i] = 0xFF000000u;
    for (uint32_t i = 0; i < colours; ++i) {
      const uint8_t* e = f + pal_off + (size_t)i * pal_entry;
      H->palette[i] = (uint32_t)e[2] | (uint32_t)e[1] << 8 | (uint32_t)e[0] << 16 | 0xFF000000u;
    }
  }
  const uint64_t stride = (((uint64_t)H->w * bpp + 31) / 32) * 4;
  if (stride > 0x7FFFFFFFu) return kUnsupported;
  H->row_stride = (uint32_t)stride;
  H->data_off = off_bits;
  if ((uint64_t)off_bits + stride * (uint64_t)H->h > n) return kCorrupt;   // the pixel array must be there in full
  return kOk;
}

}  // namespace bmpdec
}  // namespace stitch
