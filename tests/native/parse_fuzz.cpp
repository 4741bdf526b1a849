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
