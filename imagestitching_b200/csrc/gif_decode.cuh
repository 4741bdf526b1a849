// GIF inputs (P:4 SUPPORTED_IMAGE_TYPES 'gif'; the decode inside loadImageFrom, U:27-121): the FIRST frame, which is
// what canvas.createImage() + drawImage paint.  The container is walked on the host (screen descriptor, colour tables,
// graphic control extension, the list of data sub-blocks: one length byte per 255 data bytes is all it reads); the file
// is uploaded as it is and three kernels do the rest:
//   G0 gather   the data sub-blocks -> one contiguous LZW stream.
//   G1 lzw      one warp per image.  LZW looks sequential twice over -- variable-width codes, and a dictionary built
//               from the output -- but every dictionary string is a SUBSTRING OF THE OUTPUT ALREADY WRITTEN: the entry
//               made while code c follows code p is out[pos(p) .. pos(p) + len(p) + 1).  So the dictionary is two
//               numbers per entry (start, length) in shared memory, a code is a copy of `length` bytes from `start`
//               that the 32 lanes share, and the root codes are a 256-byte prelude in front of the output, so they are
//               copies like the rest.  All lanes run the code loop in lock step (the same bits, the same table).  The
//               last 16 KB of output are also kept in shared memory: a dictionary of 4096 entries seldom reaches
//               further back, so a code costs shared-memory latency, not an L2 round trip (1600 x 1200 photo: 295 -> 197 ms;
//               what is left is the warp's own ~70 dependent instructions per code).
//   G2 rgba     indices -> RGBA8 through the colour table; the transparent index gets alpha 0; interlaced files store
//               rows 0, 8, 16 .. then 4, 12 .. then 2, 6 .. then 1, 3 ..
// Stated pin: PIL's GifImagePlugin, frame 0, convert("RGBA").  A frame that does not cover the logical screen is
// refused (STITCH_E_BAD_IMAGE), as is a stream with a code outside the dictionary.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

#include "gif_parse.h"

namespace stitch {
namespace gifdec {

constexpr int kWindow = 16384;     // bytes of recent output the LZW warp keeps in shared memory (power of two)

struct Img {
  const uint8_t* file;
  uint8_t* stream;              // gathered LZW bytes, 4-byte aligned, 8 zero bytes behind
  uint8_t* idx;                 // kPrelude root bytes + w * h palette indices (rows in stream order)
  const uint32_t* palette;
  uint8_t* rgba;
  int64_t pitch;
  uint32_t stream_len;
  int32_t w, h, interlaced, min_code;
};
struct Batch {
  Img img[kMaxImages];
  int32_t n;
  const Piece* pieces;          // all images' pieces; piece_img[] says whose
  const uint8_t* piece_img;
  uint32_t* status;             // != 0: a damaged stream
};

// G0
__global__ void __launch_bounds__(64) gif_gather_kernel(const __grid_constant__ Batch B) {
  const Piece pc = B.pieces[blockIdx.x];
  const Img& I = B.img[B.piece_img[blockIdx.x]];
  for (uint32_t k = threadIdx.x; k < pc.len; k += 64) I.stream[pc.stream_off + k] = I.file[pc.file_off + k];
}

// G1: grid = images, one warp each; every lane executes the same code loop.
__global__ void __launch_bounds__(32) gif_lzw_kernel(const __grid_constant__ Batch B) {
  __shared__ uint32_t start[4096];     // where the entry's string begins in idx[]
  __shared__ uint16_t len[4096];
  __shared__ uint8_t ring[kWindow];    // the last kWindow bytes of output, slot = position mod kWindow
  const Img& I = B.img[blockIdx.x];
  const int lane = threadIdx.x;
  const int clear = 1 << I.min_code, eoi = clear + 1;
  uint8_t* out = I.idx;
  for (int i = lane; i < kPrelude; i += 32) out[i] = (uint8_t)i;
  for (int i = lane; i < clear; i += 32) { start[i] = (uint32_t)i; len[i] = 1; }
  __syncwarp();
  const uint32_t total = (uint32_t)kPrelude + (uint32_t)I.w * (uint32_t)I.h;
  const uint32_t* words = reinterpret_cast<const uint32_t*>(I.stream);
  const uint32_t n_bits = I.stream_len * 8u;
  uint64_t acc = 0;                    // unread bits, LSB first
  int cnt = 0;
  uint32_t rd = 0;                     // next word to fetch
  uint32_t bitpos = 0;                 // bits consumed
  int size = I.min_code + 1, next = clear + 2;
  int prev = -1;
  uint32_t pos = kPrelude, prev_pos = 0;
  bool bad = false;
  while (pos < total) {
    if (cnt < size) {
      acc |= (uint64_t)words[rd++] << cnt;     // the stream buffer is zero-padded behind its end
      cnt += 32;
    }
    if (bitpos + (uint32_t)size > n_bits) break;              // the data ran out: the rest of the picture stays index 0
    const int code = (int)(acc & ((1u << size) - 1u));
    acc >>= size;
    cnt -= size;
    bitpos += (uint32_t)size;
    if (code == clear) {
      size = I.min_code + 1;
      next = clear + 2;
      prev = -1;
      continue;
    }
    if (code == eoi) break;
    uint32_t src, L;
    if (code < next && (code < clear || code > eoi)) {
      src = start[code];
      L = len[code];
    } else if (code == next && prev >= 0) {                   // the string being defined: prev's string + its own first byte
      src = prev_pos;
      L = (uint32_t)len[prev] + 1u;
    } else {
      bad = true;
      break;
    }
    if (L > total - pos) L = total - pos;
    // the copy: bytes [src, src + L) were written by earlier codes, except -- in the second case -- the last one, which
    // is this copy's own first byte
    // A source byte is one of three things: a root (the prelude holds i at i: no load at all), a byte of the last
    // kWindow - L of output (the shared-memory ring every stored byte also goes to: ~30 cycles instead of an L2 round
    // trip per code), or older output (L2).  The ring slots this copy writes cannot be ones it still reads: a read
    // counts as inside the window only when its slot is out of this copy's reach.
    for (uint32_t j = lane; j < L; j += 32) {
      const uint32_t s = src + j, from = s < pos ? s : src;
      uint8_t v;
      if (from < (uint32_t)kPrelude) v = (uint8_t)from;
      else if (from + (uint32_t)kWindow >= pos + L) v = ring[from & (kWindow - 1)];
      else v = __ldcg(out + from);
      out[pos + j] = v;
      ring[(pos + j) & (kWindow - 1)] = v;
    }
    __syncwarp();
    if (prev >= 0 && next < 4096) {
      // every lane stores the same two values and later reads what it stored itself: no exchange between lanes, no barrier
      start[next] = prev_pos;
      len[next] = (uint16_t)(len[prev] + 1);
      ++next;
      if (next == (1 << size) && size < 12) ++size;
    }
    prev = code;
    prev_pos = pos;
    pos += L;
  }
  // indices the stream did not deliver: 0, as decoders leave them
  for (uint32_t j = pos + lane; j < total; j += 32) out[j] = 0;
  if (bad && lane == 0) atomicOr(B.status, 1u);
}

// G2
__global__ void __launch_bounds__(256) gif_rgba_kernel(const __grid_constant__ Batch B) {
  const Img& I = B.img[blockIdx.z];
  const int x = blockIdx.x * 64 + (threadIdx.x & 63);
  const int r = blockIdx.y * 4 + (threadIdx.x >> 6);          // row in stream order
  if (x >= I.w || r >= I.h) return;
  int y = r;
  if (I.interlaced) {
    const int n0 = (I.h + 7) / 8, n1 = (I.h + 3) / 8, n2 = (I.h + 1) / 4;
    if (r < n0) y = r * 8;
    else if (r < n0 + n1) y = (r - n0) * 8 + 4;
    else if (r < n0 + n1 + n2) y = (r - n0 - n1) * 4 + 2;
    else y = (r - n0 - n1 - n2) * 2 + 1;
  }
  const uint32_t v = I.idx[kPrelude + (size_t)r * I.w + x];
  *reinterpret_cast<uint32_t*>(I.rgba + (size_t)y * I.pitch + (size_t)x * 4) = I.palette[v];
}

}  // namespace gifdec
}  // namespace stitch
