// Tensor maps of the streamed bitmaps: what the compose kernel's TMA loads (cp.async.bulk.tensor.2d) address.
//
// One pair per image of a job: `box` fetches STREAM_TMA_ROWS rows x 128 pixels (512 B per row: one warp's band of
// 32 four-pixel groups) in ONE instruction, `row` a single row of the same width for the last, partial stage of a
// unit, so no source row is ever fetched beyond the rows a unit consumes.  The tensor is the bitmap itself --
// u32 elements, {bmp_width, bmp_height}, row stride = pitch -- so columns past the bitmap edge are zero-filled
// by the hardware instead of being read from the row padding (or from beyond a tightly allocated buffer).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <unordered_map>
#include <vector>

#include "stitch.h"

#ifndef STREAM_TMA_ROWS
#define STREAM_TMA_ROWS 8     // source rows per stage = per TMA box
#endif
#ifndef STREAM_TMA_STAGES
#define STREAM_TMA_STAGES 3   // stages per warp ring
#endif
#define STREAM_TMA_BOX_PX 128  // 32 lanes x 4 pixels
#ifndef STREAM_TMA_EVICT_FIRST
#define STREAM_TMA_EVICT_FIRST 0
#endif

namespace stitch {

struct TmaMaps {
  CUtensorMap box[STITCH_MAX_IMAGES];
  CUtensorMap row[STITCH_MAX_IMAGES];
};

typedef CUresult (*TmaEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point table: libstitch.so does not link libcuda.
inline TmaEncodeFn tma_encode_fn() {
  static TmaEncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (TmaEncodeFn)p;
  }();
  return fn;
}

// Maps are a pure function of (address, pitch, size): a job that reuses its bitmap buffers (a lane's staging, the
// resident job slots of the batch sweep: 32 slots x 9 bitmaps) finds them here instead of encoding 18 descriptors per
// submit (~60 us of host time per job when every lookup missed, against 80 us of GPU time).
struct TmaCache {
  struct Key {
    const void* data; int64_t pitch; int32_t w, h;
    bool operator==(const Key& o) const { return data == o.data && pitch == o.pitch && w == o.w && h == o.h; }
  };
  struct KeyHash {
    size_t operator()(const Key& k) const {
      uint64_t x = (uint64_t)(uintptr_t)k.data ^ ((uint64_t)k.pitch * 0x9E3779B97F4A7C15ull) ^ ((uint64_t)(uint32_t)k.w << 32) ^ (uint32_t)k.h;
      x ^= x >> 29; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 32;
      return (size_t)x;
    }
  };
  struct Maps { CUtensorMap box, row; };
  std::unordered_map<Key, Maps, KeyHash> e;
  static constexpr size_t kCap = 8192;   // 2 MB of descriptors at most; cleared when full

  // cudaSuccess and the two maps, or an error (no driver entry point / encode refused)
  cudaError_t get(const void* data, int64_t pitch, int32_t w, int32_t h, CUtensorMap* box, CUtensorMap* row) {
    const Key key{data, pitch, w, h};
    auto it = e.find(key);
    if (it != e.end()) { *box = it->second.box; *row = it->second.row; return cudaSuccess; }
    const TmaEncodeFn enc = tma_encode_fn();
    if (!enc) return cudaErrorNotSupported;
    Maps n;
    const cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h};
    const cuuint64_t strides[1] = {(cuuint64_t)pitch};
    const cuuint32_t es[2] = {1, 1};
    const cuuint32_t b8[2] = {STREAM_TMA_BOX_PX, STREAM_TMA_ROWS}, b1[2] = {STREAM_TMA_BOX_PX, 1};
    if (enc(&n.box, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(data), dims, strides, b8, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS ||
        enc(&n.row, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(data), dims, strides, b1, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return cudaErrorInvalidValue;
    if (e.size() >= kCap) e.clear();
    e.emplace(key, n);
    *box = n.box; *row = n.row;
    return cudaSuccess;
  }
};

}  // namespace stitch
