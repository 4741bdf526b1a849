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
#include <vector>

#include "stitch.h"

#ifndef STITCH_STREAM_TMA
#define STITCH_STREAM_TMA 1   // 0: per-lane cp.async (LDGSTS) staging, the round-1 form (kept for A/B runs)
#endif
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

#if STITCH_STREAM_TMA
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

// Maps are a pure function of (address, pitch, size): a job that reuses its bitmap buffers (a lane's staging, a
// resident job slot of the batch sweep) finds them here instead of encoding 18 descriptors per submit.
struct TmaCache {
  struct Entry { const void* data; int64_t pitch; int32_t w, h; CUtensorMap box, row; };
  std::vector<Entry> e;
  size_t next = 0;
  static constexpr size_t kCap = 64;

  // cudaSuccess and the two maps, or an error (no driver entry point / encode refused)
  cudaError_t get(const void* data, int64_t pitch, int32_t w, int32_t h, CUtensorMap* box, CUtensorMap* row) {
    for (const Entry& x : e)
      if (x.data == data && x.pitch == pitch && x.w == w && x.h == h) { *box = x.box; *row = x.row; return cudaSuccess; }
    const TmaEncodeFn enc = tma_encode_fn();
    if (!enc) return cudaErrorNotSupported;
    Entry n;
    n.data = data; n.pitch = pitch; n.w = w; n.h = h;
    const cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h};
    const cuuint64_t strides[1] = {(cuuint64_t)pitch};
    const cuuint32_t es[2] = {1, 1};
    const cuuint32_t b8[2] = {STREAM_TMA_BOX_PX, STREAM_TMA_ROWS}, b1[2] = {STREAM_TMA_BOX_PX, 1};
    if (enc(&n.box, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(data), dims, strides, b8, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS ||
        enc(&n.row, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(data), dims, strides, b1, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return cudaErrorInvalidValue;
    if (e.size() < kCap) e.push_back(n);
    else { e[next] = n; next = (next + 1) % kCap; }
    *box = n.box; *row = n.row;
    return cudaSuccess;
  }
};
#else
struct TmaCache {};
#endif

}  // namespace stitch
