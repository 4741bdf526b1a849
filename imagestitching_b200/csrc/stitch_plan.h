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
  int32_t transparent;  // 0: opaque white canvas (P:1423-1424); 1: cleared canvas, straight-alpha RGBA out (preview, P:1598)
  uint8_t* out_rgba;  // may be null
  int64_t out_pitch;
  uint8_t* out_y;     // may be null
  uint8_t* out_cb;
  uint8_t* out_cr;
};

// ---- fused compose launch: the canvas is partitioned into regions, each region into warp units ----
enum RegionKind : int32_t {
  REGION_STREAM = 0,  // integer-aligned area downscale of one image: streaming kernel
  REGION_FILL = 1,    // white background (P:1423-1424) where no image lands
  REGION_GATHER = 2,  // anything else inside one image's bounding box: per-pixel gather
  REGION_UPSCALE = 3  // fully covered pixels of one enlarged image (bilinear on both axes), EXIF 1-4
};

struct Region {
  int32_t kind, img;
  // STREAM: source-order output index ranges [x0,x1) x [y0,y1) inside the image's device rect
  // FILL/GATHER: device pixel rect (rect of 2x2 chroma blocks when chroma is produced); UPSCALE: device pixel rect
  int32_t x0, y0, x1, y1;
  int32_t tile_w, tile_h;    // warp-unit size (output pixels); STREAM uses tile_h only (bands are 31 groups wide)
  int32_t tiles_x, tiles_y;  // warp-unit grid
  int32_t first_unit;
  int32_t g_base;            // STREAM: first 4-pixel source group of the region; unit tx owns groups g_base + 31*tx ...
};

#define STITCH_MAX_REGIONS 128
// row-table entries per warp (= source rows of the tallest unit + padding; one 16-byte entry per row PAIR).  With the
// TMA ring (24 rows x 512 B per warp) 224 rows keep two CTAs of 8 warps inside the SM's 228 KB of shared memory.
#define STREAM_TAB_MAX 224
#define STREAM_WARPS 8

struct Plan {
  Region r[STITCH_MAX_REGIONS];
  int32_t n_regions;
  int32_t n_units;
};
