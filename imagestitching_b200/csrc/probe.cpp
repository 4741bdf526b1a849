// Header probe: what wx.getImageInfo hands to _prepareSingleImage (P:678-741 -- info.width, info.height,
// info.type, info.orientation), read from the first bytes of the file.  Pure CPU, no CUDA.
//
// Formats are the reference's SUPPORTED_IMAGE_TYPES (P:4): jpg/jpeg, png, gif, bmp, webp.
//   JPEG  ITU-T T.81 B.1: marker segments; SOFn carries Y (lines) and X (samples per line); APP1 "Exif\0\0"
//         wraps a TIFF structure whose IFD0 tag 0x0112 is the orientation (EXIF 2.3 table 4).
//   PNG   signature + IHDR; optional eXIf chunk (PNG ext. 1.5) holds the same TIFF structure.
//   GIF   logical screen descriptor; BMP  BITMAPCOREHEADER / BITMAPINFOHEADER; WEBP  VP8 / VP8L / VP8X.
#include <cstring>

#include "stitch.h"

namespace {

struct Reader {
  const uint8_t* p;
  size_t n;
  bool has(size_t off, size_t len) const { return off <= n && len <= n - off; }
  uint32_t be16(size_t o) const { return (uint32_t)p[o] << 8 | p[o + 1]; }
  uint32_t be32(size_t o) const { return be16(o) << 16 | be16(o + 2); }
  uint32_t le16(size_t o) const { return (uint32_t)p[o + 1] << 8 | p[o]; }
  uint32_t le24(size_t o) const { return (uint32_t)p[o + 2] << 16 | le16(o); }
  uint32_t le32(size_t o) const { return le16(o + 2) << 16 | le16(o); }
};

// Orientation from a TIFF structure at [base, base+len): IFD0 entry 0x0112, type SHORT (3), count 1.
int tiff_orientation(const Reader& r, size_t base, size_t len) {
  if (len < 8 || !r.has(base, len)) return 0;
  const bool le = r.p[base] == 'I' && r.p[base + 1] == 'I';
  const bool be = r.p[base] == 'M' && r.p[base + 1] == 'M';
  if (!le && !be) return 0;
  auto u16 = [&](size_t o) { return le ? r.le16(base + o) : r.be16(base + o); };
  auto u32 = [&](size_t o) { return le ? r.le32(base + o) : r.be32(base + o); };
  if (u16(2) != 42) return 0;
  const size_t ifd = u32(4);
  if (ifd + 2 > len) return 0;
  const uint32_t count = u16(ifd);
  for (uint32_t i = 0; i < count; ++i) {
    const size_t e = ifd + 2 + (size_t)i * 12;
    if (e + 12 > len) return 0;
    if (u16(e) != 0x0112) continue;
    const uint32_t type = u16(e + 2);
    uint32_t v = 0;
    if (type == 3) v = u16(e + 8);        // SHORT, left-justified in the value field
    else if (type == 4) v = u32(e + 8);   // LONG (seen in the wild)
    return (v >= 1 && v <= 8) ? (int)v : 0;
  }
  return 0;
}

int probe_jpeg(const Reader& r, StitchImageInfo* out) {
  size_t o = 2;
  bool have_size = false;
  while (r.has(o, 4)) {
    if (r.p[o] != 0xFF) { ++o; continue; }           // resync on garbage between segments
    const uint8_t m = r.p[o + 1];
    if (m == 0xFF) { ++o; continue; }                // fill bytes
    if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { o += 2; continue; }  // stand-alone markers
    if (m == 0xD9 || m == 0xDA) break;               // EOI / start of scan: the headers are over
    const size_t len = r.be16(o + 2);
    if (len < 2 || !r.has(o + 2, len)) break;
    if (m >= 0xC0 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
      if (len >= 7) {
        out->height = (int32_t)r.be16(o + 5);
        out->width = (int32_t)r.be16(o + 7);
        have_size = true;
      }
    } else if (m == 0xE1 && len >= 8 && std::memcmp(r.p + o + 4, "Exif\0\0", 6) == 0 && out->orientation_exif == 0) {
      out->orientation_exif = tiff_orientation(r, o + 10, len - 8);
    }
    if (have_size && out->orientation_exif) break;
    o += 2 + len;
  }
  return have_size ? STITCH_OK : STITCH_E_BAD_IMAGE;
}

int probe_png(const Reader& r, StitchImageInfo* out) {
  if (!r.has(8, 25) || std::memcmp(r.p + 12, "IHDR", 4) != 0) return STITCH_E_BAD_IMAGE;
  out->width = (int32_t)r.be32(16);
  out->height = (int32_t)r.be32(20);
  size_t o = 8;
  while (r.has(o, 12)) {  // length, type, data, crc
    const size_t len = r.be32(o);
    if (std::memcmp(r.p + o + 4, "eXIf", 4) == 0 && r.has(o + 8, len)) {
      out->orientation_exif = tiff_orientation(r, o + 8, len);
      break;
    }
    if (std::memcmp(r.p + o + 4, "IDAT", 4) == 0 || std::memcmp(r.p + o + 4, "IEND", 4) == 0) break;
    o += 12 + len;
  }
  return STITCH_OK;
}

int probe_webp(const Reader& r, StitchImageInfo* out) {
  if (!r.has(12, 8)) return STITCH_E_BAD_IMAGE;
  const uint8_t* f = r.p + 12;
  if (std::memcmp(f, "VP8 ", 4) == 0 && r.has(20, 10)) {          // lossy: frame tag 3 B, start code 9D 01 2A
    if (r.p[23] != 0x9D || r.p[24] != 0x01 || r.p[25] != 0x2A) return STITCH_E_BAD_IMAGE;
    out->width = (int32_t)(r.le16(26) & 0x3FFF);
    out->height = (int32_t)(r.le16(28) & 0x3FFF);
    return STITCH_OK;
  }
  if (std::memcmp(f, "VP8L", 4) == 0 && r.has(20, 5)) {           // lossless: 0x2F, then 14 + 14 bits (minus 1)
    if (r.p[20] != 0x2F) return STITCH_E_BAD_IMAGE;
    const uint32_t b = r.le32(21);
    out->width = (int32_t)((b & 0x3FFF) + 1);
    out->height = (int32_t)(((b >> 14) & 0x3FFF) + 1);
    return STITCH_OK;
  }
  if (std::memcmp(f, "VP8X", 4) == 0 && r.has(20, 10)) {          // extended: flags, 24-bit canvas size minus 1
    out->width = (int32_t)(r.le24(24) + 1);
    out->height = (int32_t)(r.le24(27) + 1);
    if (r.p[20] & 0x08) {                                         // EXIF chunk present: walk the chunks
      size_t o = 12;
      while (r.has(o, 8)) {
        const size_t len = r.le32(o + 4);
        if (std::memcmp(r.p + o, "EXIF", 4) == 0 && r.has(o + 8, len)) {
          size_t b = o + 8, l = len;
          if (l >= 6 && std::memcmp(r.p + b, "Exif\0\0", 6) == 0) { b += 6; l -= 6; }
          out->orientation_exif = tiff_orientation(r, b, l);
          break;
        }
        o += 8 + len + (len & 1);
      }
    }
    return STITCH_OK;
  }
  return STITCH_E_BAD_IMAGE;
}

}  // namespace

extern "C" int stitch_probe_image(const void* file, size_t bytes, StitchImageInfo* out) {
  if (!out) return STITCH_E_BAD_ARG;
  std::memset(out, 0, sizeof(*out));
  if (!file) return STITCH_E_BAD_ARG;
  const Reader r{(const uint8_t*)file, bytes};
  int st = STITCH_E_BAD_IMAGE;
  if (r.has(0, 4) && r.p[0] == 0xFF && r.p[1] == 0xD8) {
    out->type = STITCH_TYPE_JPEG;
    st = probe_jpeg(r, out);
  } else if (r.has(0, 8) && std::memcmp(r.p, "\x89PNG\r\n\x1a\n", 8) == 0) {
    out->type = STITCH_TYPE_PNG;
    st = probe_png(r, out);
  } else if (r.has(0, 10) && (std::memcmp(r.p, "GIF87a", 6) == 0 || std::memcmp(r.p, "GIF89a", 6) == 0)) {
    out->type = STITCH_TYPE_GIF;
    out->width = (int32_t)r.le16(6);
    out->height = (int32_t)r.le16(8);
    st = STITCH_OK;
  } else if (r.has(0, 26) && r.p[0] == 'B' && r.p[1] == 'M') {
    out->type = STITCH_TYPE_BMP;
    const uint32_t dib = r.le32(14);
    if (dib == 12) {  // BITMAPCOREHEADER: 16-bit sizes
      out->width = (int32_t)r.le16(18);
      out->height = (int32_t)r.le16(20);
    } else {
      const int32_t w = (int32_t)r.le32(18), h = (int32_t)r.le32(22);
      out->width = w;
      out->height = h < 0 ? -h : h;  // negative height = top-down rows
    }
    st = STITCH_OK;
  } else if (r.has(0, 12) && std::memcmp(r.p, "RIFF", 4) == 0 && std::memcmp(r.p + 8, "WEBP", 4) == 0) {
    out->type = STITCH_TYPE_WEBP;
    st = probe_webp(r, out);
  }
  if (st == STITCH_OK && (out->width <= 0 || out->height <= 0)) st = STITCH_E_BAD_IMAGE;
  if (out->orientation_exif == 0) out->orientation_exif = 1;
  return st;
}

// wx.getImageInfo's orientation strings for EXIF 1..8.
extern "C" const char* stitch_orientation_name(int32_t exif) {
  static const char* const names[9] = {"up", "up", "up-mirrored", "down", "down-mirrored",
                                       "left-mirrored", "right", "right-mirrored", "left"};
  return names[(exif >= 1 && exif <= 8) ? exif : 0];
}

extern "C" const char* stitch_type_name(int32_t type) {
  static const char* const names[6] = {"unknown", "jpeg", "png", "gif", "bmp", "webp"};
  return names[(type >= 0 && type <= 5) ? type : 0];
}

// _prepareSingleImage (P:724-739): the record the layout reads, from the probe and the file size.
extern "C" int stitch_prepare_record(const StitchImageInfo* info, int64_t file_size, StitchImageDesc* out) {
  if (!info || !out) return STITCH_E_BAD_ARG;
  std::memset(out, 0, sizeof(*out));
  out->natural_width = info->width > 1 ? info->width : 1;    // Math.max(1, info.width ...)  P:729
  out->natural_height = info->height > 1 ? info->height : 1; // P:730
  out->width = out->natural_width;                           // P:731-732
  out->height = out->natural_height;
  // P:733 stores info.orientation, which in WeChat is a STRING ('up', 'right', ...); drawWithOrientation
  // (U:160-199) switches on the numbers 2..8, so a prepared record always draws un-rotated.  The numeric
  // EXIF value stays available in StitchImageInfo for callers that want to apply it themselves.
  out->orientation = 1;
  out->file_size = file_size > 0 ? file_size : 0;            // P:734
  return STITCH_OK;
}
