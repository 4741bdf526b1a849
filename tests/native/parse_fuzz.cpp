// The host-side parsers of the file decoders (jpeg_parse.h, png_parse.h, bmp_parse.h, gif_parse.h) on damaged files,
// built with -fsanitize=address,undefined by tests/test_parse_fuzz.py: whatever the bytes, a parser returns a status
// and never reads outside the buffer it was given.  TEST CODE ONLY; nothing in the library links it.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bmp_parse.h"
#include "gif_parse.h"
#include "jpeg_parse.h"
#include "png_parse.h"

// kind: 0 JPEG, 1 PNG, 2 BMP, 3 GIF.  The file is copied into an exact-size heap block so that any overread trips ASan.
// Returns the parser's status; for accepted files also checks that what the header promises lies inside the file.
extern "C" int parse_fuzz_one(int kind, const uint8_t* data, size_t n) {
  uint8_t* f = (uint8_t*)std::malloc(n ? n : 1);
  std::memcpy(f, data, n);
  int st = -1;
  if (kind == 0) {
    stitch::jpegdec::Header H;
    st = (int)stitch::jpegdec::parse(f, n, &H);
    if (st == 0) {
      if (H.progressive) {
        for (const auto& s : H.scans)
          if (s.off + s.len > n || s.tab < 0 || s.tab >= (int)H.scan_tabs.size()) st = 100;
      } else if (H.scan_off + H.scan_len > n) {
        st = 100;
      }
      if (H.w <= 0 || H.h <= 0) st = 101;
    }
  } else if (kind == 1) {
    stitch::pngdec::Header H;
    st = (int)stitch::pngdec::parse(f, n, &H);
    if (st == 0)
      for (const auto& pc : H.idat)
        if ((size_t)pc.file_off + pc.len > n) st = 100;
  } else if (kind == 2) {
    stitch::bmpdec::Header H;
    st = (int)stitch::bmpdec::parse(f, n, &H);
    if (st == 0 && (uint64_t)H.data_off + (uint64_t)H.row_stride * (uint64_t)H.h > n) st = 100;
  } else {
    stitch::gifdec::Header H;
    st = (int)stitch::gifdec::parse(f, n, &H);
    if (st == 0)
      for (const auto& pc : H.pieces)
        if ((size_t)pc.file_off + pc.len > n) st = 100;
  }
  std::free(f);
  return st;
}

// What the parsers read from an intact file: out = {status, width, height, a, b, c, d, e}
//   JPEG: a = components, b = luma h-sampling, c = luma v-sampling, d = progressive, e = scans
//   PNG:  a = bit depth, b = colour type, c = interlace, d = row bytes
//   BMP:  a = bits per pixel, b = top-down, c = row stride, d = offset of the pixel array
//   GIF:  a = interlaced, b = LZW minimum code size, c = transparent index (-1: none), d = bytes of LZW data
extern "C" void parse_fields(int kind, const uint8_t* f, size_t n, int64_t* out) {
  for (int i = 0; i < 8; ++i) out[i] = 0;
  if (kind == 0) {
    stitch::jpegdec::Header H;
    out[0] = (int)stitch::jpegdec::parse(f, n, &H);
    out[1] = H.w; out[2] = H.h; out[3] = H.ncomp; out[4] = H.hs; out[5] = H.vs; out[6] = H.progressive; out[7] = (int64_t)H.scans.size();
  } else if (kind == 1) {
    stitch::pngdec::Header H;
    out[0] = (int)stitch::pngdec::parse(f, n, &H);
    out[1] = H.w; out[2] = H.h; out[3] = H.depth; out[4] = H.ctype; out[5] = H.interlace; out[6] = H.row_bytes;
  } else if (kind == 2) {
    stitch::bmpdec::Header H;
    out[0] = (int)stitch::bmpdec::parse(f, n, &H);
    out[1] = H.w; out[2] = H.h; out[3] = H.bpp; out[4] = H.top_down; out[5] = H.row_stride; out[6] = H.data_off;
  } else {
    stitch::gifdec::Header H;
    out[0] = (int)stitch::gifdec::parse(f, n, &H);
    out[1] = H.w; out[2] = H.h; out[3] = H.interlaced; out[4] = H.min_code; out[5] = H.transparent; out[6] = H.stream_len;
  }
}
