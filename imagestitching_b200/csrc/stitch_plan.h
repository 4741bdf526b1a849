// Internal: per-job kernel parameters built on the host from a StitchLayout.
#pragma once
#include <cstdint>

#include "stitch.h"

// One drawWithOrientation() call (U:153-202) resolved to an axis-aligned placement in
// DEVICE space (user rect x superSample CTM, P:1426-1428).  "x axis"/"y axis" below are
// DEVICE axes; sx_len/sy_len are the bitmap extents that run along them (swapped for
// EXIF 5-8), flip_x/flip_y mirror the bitmap along that device axis.
struct ImgParam {
  const uint8_t* data;
  int64_t pitch;
  double X0, Y0, W, H;  // device rect (doubles exactly as rect*ss)
  int32_t bw, bh;       // bitmap width/height
  int32_t sx_len, sy_len;
  int32_t x_lo, x_hi, y_lo, y_hi;  // integer bounding box of covered device pixels, clipped to the canvas
  int32_t transpose, flip_x, flip_y;
  int32_t pad;
};

struct JobParam {
  ImgParam img[STITCH_MAX_IMAGES];
  int32_t n;
  int32_t cw, ch;
  int32_t pad;
  uint8_t* out_rgba;  // may be null
  int64_t out_pitch;
  uint8_t* out_y;     // may be null
  uint8_t* out_cb;
  uint8_t* out_cr;
};
