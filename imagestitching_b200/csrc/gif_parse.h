// Host-side container walk of the GIF decoder (gif_decode.cuh): logical screen, colour tables, graphic control
// extension, the first image descriptor and the list of its data sub-blocks.  Plain C++ (tests/native/parse_fuzz.cpp
// runs it under AddressSanitizer on damaged files).  Replaces the first half of loadImageFrom (U:27-121) for 'gif' (P:4).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

namespace stitch {
namespace gifdec {

constexpr int kMaxImages = 9;
constexpr int kPrelude = 256;     // the root strings, in front of the index buffer

struct Piece {
  uint32_t file_off, stream_off, len;
};

struct Header {
  int32_t w = 0, h = 0;
  bool interlaced = false;
  int32_t min_code = 0;
  int32_t transparent = -1;        // palette index with alpha 0, or -1
  uint32_t palette[256];           // RGBA8
  uint32_t stream_len = 0;
  std::vector<Piece> pieces;
};

enum ParseStatus { kOk = 0, kNotGif = 1, kCorrupt = 2, kUnsupported = 3 };

inline ParseStatus parse(const uint8_t* f, size_t n, Header* H) {
  if (n < 13 || (std::memcmp(f, "GIF87a", 6) != 0 && std::memcmp(f, "GIF89a", 6) != 0)) return kNotGif;
  const int sw = f[6] | f[7] << 8, sh = f[8] | f[9] << 8;
  if (sw <= 0 || sh <= 0) return kCorrupt;
  size_t o = 13;
  const uint8_t* table = nullptr;
  int table_n = 0;
  if (f[10] & 0x80) {
    table_n = 2 << (f[10] & 7);
    if (o + 3 * (size_t)table_n > n) return kCorrupt;
    table = f + o;
    o += 3 * (size_t)table_n;
  }
  H->transparent = -1;
  while (o < n) {
    const int b = f[o++];
    if (b == 0x3B) return kCorrupt;                    // trailer before any image
    if (b == 0x21) {                                   // extension: label, sub-blocks
      if (o >= n) return kCorrupt;
      const int label = f[o++];
      bool first = true;
      for (;;) {
        if (o >= n) return kCorrupt;
        const int len = f[o++];
        if (!len) break;
        if (o + len > n) return kCorrupt;
        if (label == 0xF9 && first && len >= 4) H->transparent = (f[o] & 1) ? f[o + 3] : -1;   // graphic control
        first = false;
        o += len;
      }
      continue;
    }
    if (b != 0x2C) return kCorrupt;
    if (o + 9 > n) return kCorrupt;
    const int left = f[o] | f[o + 1] << 8, top = f[o + 2] | f[o + 3] << 8;
    const int w = f[o + 4] | f[o + 5] << 8, h = f[o + 6] | f[o + 7] << 8;
    const int packed = f[o + 8];
    o += 9;
    if (left != 0 || top != 0 || w != sw || h != sh) return kUnsupported;   // a frame smaller than the screen
    H->w = w;
    H->h = h;
    H->interlaced = (packed & 0x40) != 0;
    if (packed & 0x80) {
      table_n = 2 << (packed & 7);
      if (o + 3 * (size_t)table_n > n) return kCorrupt;
      table = f + o;
      o += 3 * (size_t)table_n;
    }
    for (int i = 0; i < 256; ++i) H->palette[i] = 0xFF000000u;             // entries the table does not have: black
    for (int i = 0; i < table_n; ++i)
      H->palette[i] = (uint32_t)table[3 * i] | (uint32_t)table[3 * i + 1] << 8 | (uint32_t)table[3 * i + 2] << 16 | 0xFF000000u;
    if (H->transparent >= 0) H->palette[H->transparent] &= 0x00FFFFFFu;
    if (o >= n) return kCorrupt;
    H->min_code = f[o++];
    if (H->min_code < 1 || H->min_code > 8) return kCorrupt;               // 1 is written as 2 by encoders, but decoders take it
    H->pieces.clear();
    uint64_t total = 0;
    for (;;) {
      if (o >= n) break;                                                   // cut short: decode what is there
      const uint32_t len = f[o++];
      if (!len) break;
      const uint32_t take = o + len > n ? (uint32_t)(n - o) : len;
      H->pieces.push_back(Piece{(uint32_t)o, (uint32_t)total, take});
      total += take;
      o += len;
      if (total >= ((uint64_t)1 << 30)) return kUnsupported;
    }
    H->stream_len = (uint32_t)total;
    return total ? kOk : kCorrupt;
  }
  return kCorrupt;
}

}  // namespace gifdec
}  // namespace stitch
