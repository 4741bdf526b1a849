// BMP inputs (P:4 SUPPORTED_IMAGE_TYPES 'bmp'; the decode inside loadImageFrom, U:27-121): header and palette are
// parsed on the host (a few hundred bytes), the file is uploaded as it is and ONE kernel turns its pixel array into
// the RGBA8 bitmap -- a thread per pixel, rows bottom-up or top-down, 1 / 4 / 8 bit palettised, 16 bit (5-5-5, 5-6-5 or
// any contiguous bit fields), 24 bit, 32 bit (alpha only where a mask says so).  The rules are those of the common
// decoders (stated pin: PIL's BmpImagePlugin -- a 32-bit BI_RGB file has no alpha, n-bit fields widen by
// v * 255 / (2^n - 1)).  RLE4 / RLE8 files are refused (STITCH_E_BAD_IMAGE): run-length rows have no random access and
// such files are screen captures of the 1990s.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <cstring>

#include "bmp_parse.h"

namespace stitch {
namespace bmpdec {

struct Img {
  const uint8_t* file;          // device copy of the file
  const uint32_t* palette;      // device, 256 entries
  uint8_t* rgba;
  int64_t pitch;
  int32_t w, h, bpp, top_down;
  uint32_t data_off, row_stride;
  int32_t shift[4], bits[4];
};
struct Batch {
  Img img[kMaxImages];
  int32_t n;
};

__device__ __forceinline__ uint32_t bmp_field(uint32_t px, int shift, int bits) {
  const uint32_t v = (px >> shift) & (bits >= 32 ? 0xFFFFFFFFu : (1u << bits) - 1u);
  if (bits == 8) return v;
  if (bits > 8) return v >> (bits - 8);
  return v * 255u / ((1u << bits) - 1u);
}

__global__ void __launch_bounds__(256) bmp_rgba_kernel(const __grid_constant__ Batch B) {
  const Img& I = B.img[blockIdx.z];
  const int x = blockIdx.x * 64 + (threadIdx.x & 63);
  const int y = blockIdx.y * 4 + (threadIdx.x >> 6);
  if (x >= I.w || y >= I.h) return;
  const uint8_t* row = I.file + I.data_off + (size_t)(I.top_down ? y : I.h - 1 - y) * I.row_stride;
  uint32_t out;
  if (I.bpp <= 8) {
    const uint32_t bit = (uint32_t)x * (uint32_t)I.bpp;
    const uint32_t v = (row[bit >> 3] >> (8 - I.bpp - (int)(bit & 7))) & ((1u << I.bpp) - 1u);
    out = I.palette[v];
  } else if (I.bpp == 24) {
    const uint8_t* p = row + (size_t)x * 3;
    out = (uint32_t)p[2] | (uint32_t)p[1] << 8 | (uint32_t)p[0] << 16 | 0xFF000000u;
  } else {
    const uint8_t* p = row + (size_t)x * (I.bpp >> 3);           // the pixel array starts at any byte (54 as a rule)
    uint32_t px = (uint32_t)p[0] | (uint32_t)p[1] << 8;
    if (I.bpp == 32) px |= (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24;
    uint32_t c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) c[k] = I.bits[k] ? bmp_field(px, I.shift[k], I.bits[k]) : (k == 3 ? 255u : 0u);
    out = c[0] | c[1] << 8 | c[2] << 16 | c[3] << 24;
  }
  *reinterpret_cast<uint32_t*>(I.rgba + (size_t)y * I.pitch + (size_t)x * 4) = out;
}

}  // namespace bmpdec
}  // namespace stitch
