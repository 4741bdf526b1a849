// Host-side walk over a PNG FILE for the GPU decoder (png_decode.cu): IHDR, PLTE, tRNS and the list of IDAT payloads.
// What loadImageFrom (U:27-121) leaves to the platform's libpng; the rules are the PNG specification's (ISO 15948):
// chunk order, the colour-type / bit-depth table, CRC-32 on the critical chunks read here.
// IDAT CRCs and the zlib stream's Adler-32 are NOT verified (libpng's PNG_CRC_QUIET_USE / a benign error there): the
// pixels come out the same, a damaged stream is caught by the inflater's own checks.
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

namespace stitch {
namespace pngdec {

enum ParseStatus { kOk = 0, kNotPng = 1, kCorrupt = 2, kUnsupported = 3 };

struct IdatPiece {
  uint32_t file_off;   // of the payload bytes in the file
  uint32_t dst_off;    // in the concatenated zlib stream
  uint32_t len;
  uint32_t img;        // filled by the batch
};

struct Header {
  int32_t w = 0, h = 0;
  int32_t depth = 0, ctype = 0, interlace = 0;
  int32_t channels = 0;
  int32_t bpp = 0;            // bytes per complete pixel, rounded up to 1 (the filter unit)
  uint32_t row_bytes = 0;     // bytes per row without the filter-type byte
  uint64_t raw_len = 0;       // (row_bytes + 1) * h
  uint32_t z_len = 0;         // bytes of the zlib stream (all IDAT payloads)
  uint32_t palette[256];      // RGBA8 little-endian (R in the low byte), alpha from tRNS
  int32_t n_palette = 0;
  int32_t has_key = 0;        // tRNS colour key for gray / RGB
  uint16_t key[3] = {0, 0, 0};
  std::vector<IdatPiece> idat;   // payloads cut into pieces of <= kPiece bytes
};

constexpr uint32_t kPiece = 32u << 10;

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline uint32_t crc32_update(uint32_t c, const uint8_t* p, size_t n) {
  static uint32_t table[256];
  static bool init = false;
  if (!init) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t v = i;
      for (int k = 0; k < 8; ++k) v = (v & 1) ? 0xEDB88320u ^ (v >> 1) : v >> 1;
      table[i] = v;
    }
    init = true;
  }
  for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c;
}

// Adam7 (ISO 15948 clause 8.2): pass p holds the pixels (x0 + i dx, y0 + j dy).  Fills ps[] with the passes that hold
// pixels, in stream order; returns their number.
struct AdamPass {
  int32_t x0, y0, dx, dy, w, h;
  uint32_t row_bytes;
};
inline int adam7_passes(int32_t w, int32_t h, int bits, AdamPass* ps) {
  static const int8_t t[7][4] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4}, {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
  int n = 0;
  for (int p = 0; p < 7; ++p) {
    const int64_t pw = ((int64_t)w - t[p][0] + t[p][2] - 1) / t[p][2], ph = ((int64_t)h - t[p][1] + t[p][3] - 1) / t[p][3];
    if (pw <= 0 || ph <= 0) continue;
    ps[n++] = AdamPass{t[p][0], t[p][1], t[p][2], t[p][3], (int32_t)pw, (int32_t)ph, (uint32_t)((pw * bits + 7) / 8)};
  }
  return n;
}

inline ParseStatus parse(const uint8_t* f, size_t n, Header* H) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
  if (n < 8 || std::memcmp(f, sig, 8) != 0) return kNotPng;
  if (n >= ((size_t)1 << 32)) return kUnsupported;
  size_t o = 8;
  bool have_ihdr = false, have_plte = false, have_idat = false, idat_closed = false, have_iend = false;
  uint64_t z_total = 0;
  for (int i = 0; i < 256; ++i) H->palette[i] = 0xFF000000u;
  while (o + 12 <= n) {
    const uint32_t len = be32(f + o);
    if (len > 0x7FFFFFFFu || (uint64_t)o + 12 + len > n) return kCorrupt;
    const uint8_t* type = f + o + 4;
    const uint8_t* data = f + o + 8;
    const bool is_idat = std::memcmp(type, "IDAT", 4) == 0;
    if (!have_ihdr && std::memcmp(type, "IHDR", 4) != 0) return kCorrupt;
    if (!is_idat) {
      const bool critical = (type[0] & 0x20) == 0;
      const bool is_trns = std::memcmp(type, "tRNS", 4) == 0;
      if (critical || is_trns) {
        const uint32_t crc = crc32_update(0xFFFFFFFFu, type, 4 + (size_t)len) ^ 0xFFFFFFFFu;
        if (crc != be32(data + len)) {
          if (critical) return kCorrupt;
          o += 12 + (size_t)len;   // libpng drops an ancillary chunk with a bad CRC
          continue;
        }
      }
      if (have_idat) idat_closed = true;
    }
    if (std::memcmp(type, "IHDR", 4) == 0) {
      if (have_ihdr || len != 13) return kCorrupt;
      have_ihdr = true;
      const uint32_t w = be32(data), h = be32(data + 4);
      if (w == 0 || h == 0 || w > 0x7FFFFFFFu || h > 0x7FFFFFFFu) return kCorrupt;
      H->w = (int32_t)w;
      H->h = (int32_t)h;
      H->depth = data[8];
      H->ctype = data[9];
      if (data[10] != 0 || data[11] != 0 || data[12] > 1) return kCorrupt;
      H->interlace = data[12];
      const int d = H->depth;
      switch (H->ctype) {
        case 0: if (d != 1 && d != 2 && d != 4 && d != 8 && d != 16) return kCorrupt; H->channels = 1; break;
        case 2: if (d != 8 && d != 16) return kCorrupt; H->channels = 3; break;
        case 3: if (d != 1 && d != 2 && d != 4 && d != 8) return kCorrupt; H->channels = 1; break;
        case 4: if (d != 8 && d != 16) return kCorrupt; H->channels = 2; break;
        case 6: if (d != 8 && d != 16) return kCorrupt; H->channels = 4; break;
        default: return kCorrupt;
      }
      const uint64_t bits = (uint64_t)H->channels * (uint64_t)d;
      H->bpp = (int32_t)((bits + 7) / 8);
      const uint64_t rb = ((uint64_t)w * bits + 7) / 8;
      if (rb > 0x7FFFFFF0u) return kUnsupported;
      H->row_bytes = (uint32_t)rb;
      H->raw_len = (rb + 1) * (uint64_t)h;
    } else if (std::memcmp(type, "PLTE", 4) == 0) {
      if (have_plte || have_idat || len % 3 != 0 || len > 768) return kCorrupt;
      if (H->ctype == 0 || H->ctype == 4) return kCorrupt;
      have_plte = true;
      H->n_palette = (int32_t)(len / 3);
      for (uint32_t i = 0; i < len / 3; ++i)
        H->palette[i] = (H->palette[i] & 0xFF000000u) | data[3 * i] | ((uint32_t)data[3 * i + 1] << 8) | ((uint32_t)data[3 * i + 2] << 16);
    } else if (std::memcmp(type, "tRNS", 4) == 0) {
      if (have_idat) { o += 12 + (size_t)len; continue; }   // misplaced: ignored, as libpng does
      if (H->ctype == 3) {
        if (!have_plte || len > 256) { o += 12 + (size_t)len; continue; }
        for (uint32_t i = 0; i < len; ++i) H->palette[i] = (H->palette[i] & 0x00FFFFFFu) | ((uint32_t)data[i] << 24);
      } else if (H->ctype == 0 && len == 2) {
        H->has_key = 1;
        H->key[0] = (uint16_t)((data[0] << 8) | data[1]);
      } else if (H->ctype == 2 && len == 6) {
        H->has_key = 1;
        for (int c = 0; c < 3; ++c) H->key[c] = (uint16_t)((data[2 * c] << 8) | data[2 * c + 1]);
      }
    } else if (is_idat) {
      if (idat_closed) return kCorrupt;   // IDAT chunks must be consecutive
      if (H->ctype == 3 && !have_plte) return kCorrupt;
      have_idat = true;
      uint32_t off = 0;
      while (off < len) {
        const uint32_t piece = len - off < kPiece ? len - off : kPiece;
        if (z_total + piece >= ((uint64_t)1 << 29)) return kUnsupported;   // bit positions are 32-bit
        H->idat.push_back(IdatPiece{(uint32_t)(o + 8 + off), (uint32_t)z_total, piece, 0});
        z_total += piece;
        off += piece;
      }
    } else if (std::memcmp(type, "IEND", 4) == 0) {
      have_iend = true;
      break;
    } else if ((type[0] & 0x20) == 0) {
      return kCorrupt;   // unknown critical chunk
    }
    o += 12 + (size_t)len;
  }
  if (!have_ihdr || !have_idat) return kCorrupt;
  (void)have_iend;   // a file cut after its last IDAT still decodes if the stream is complete (libpng warns only)
  H->z_len = (uint32_t)z_total;
  if (H->z_len < 6) return kCorrupt;
  {   // zlib header (RFC 1950): deflate, window <= 32 KB, check bits, no preset dictionary
    uint8_t zh[2];
    int got = 0;
    for (const IdatPiece& pc : H->idat)
      for (uint32_t k = 0; k < pc.len && got < 2; ++k) zh[got++] = f[pc.file_off + k];
    if ((zh[0] & 0x0F) != 8 || (zh[0] >> 4) > 7 || ((zh[0] << 8) | zh[1]) % 31 != 0 || (zh[1] & 0x20)) return kCorrupt;
  }
  if (H->interlace) {   // Adam7: the stream holds the passes one after the other, each with its own (shorter) rows
    AdamPass ps[7];
    const int np = adam7_passes(H->w, H->h, H->channels * H->depth, ps);
    H->raw_len = 0;
    for (int k = 0; k < np; ++k) H->raw_len += ((uint64_t)ps[k].row_bytes + 1) * (uint64_t)ps[k].h;
  }
  if (H->raw_len >= ((uint64_t)1 << 31)) return kUnsupported;
  return kOk;
}

}  // namespace pngdec
}  // namespace stitch
