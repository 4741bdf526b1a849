// Internal interface of the PNG-input decoder (png_decode.cu) towards the job runtime (stitch_api.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

#include "png_parse.h"

namespace stitch {
namespace pngdec {

constexpr int kMaxImages = 9;

// One PNG file to decode into a device bitmap (RGBA8, straight alpha).
struct DecodeItem {
  const uint8_t* file;   // host memory
  size_t bytes;
  Header hdr;
  uint8_t* rgba;         // device destination
  int64_t pitch;
};

struct Arena {
  void* p = nullptr;
  size_t cap = 0;
};

// Error bits the kernels raise in the batch's status word (copied to *status_host, pinned, behind the launches).
enum : uint32_t {
  kErrStream = 1,        // the deflate stream is damaged (bad block, distance before the start, wrong length)
  kErrFilter = 2,        // a row's filter-type byte is not 0..4
  kErrCapacity = 4,      // more deflate blocks than the chain table holds
  kErrResolve = 8,       // the resolve rounds did not converge (cannot happen for a valid stream)
};

// Enqueues the decode of `n` files on `stream`; nothing here waits for the GPU unless the arena has to grow.
// *status_host (pinned) receives the OR of the error bits when the stream reaches the end of the batch.
// Returns cudaSuccess or the first CUDA error; *launches counts kernels.
cudaError_t decode_batch(Arena& arena, const DecodeItem* items, int n, cudaStream_t stream, uint32_t* status_host, int sm_count, int* launches);
void free_arena(Arena& arena);

}  // namespace pngdec
}  // namespace stitch
