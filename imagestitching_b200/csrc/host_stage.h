// Pageable host buffers at pinned speed (stitch_run only).
//
// The page hands the draw loop decoded bitmaps and takes the exported canvas in ordinary memory (a numpy array, a
// Node Buffer / Uint8Array): pageable.  cudaMemcpyAsync on pageable memory is staged by the driver through ONE thread
// and a small bounce buffer -- 9 x 12 MP up and the 89 MB canvas down took ~3x the pinned time (profiles/README.md).
// Here the library does the staging itself: a pool of copy threads moves row chunks between the caller's memory and
// pinned memory of the lane (uploads: a ring of chunks that are sent as they fill; downloads: a pinned copy of the
// canvas whose finished rectangles are copied out by the pool behind a CUDA event), so several cores share the
// memcpy (and the page faults of a freshly allocated output) while the DMA engines run at the pinned rate.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>
#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>
#endif

namespace stitch {

class CopyPool {
 public:
  static CopyPool& get() {
    static CopyPool* p = new CopyPool();   // lives until exit: worker threads must never outlive their queue
    return *p;
  }
  void submit(std::function<void()> fn) {
    {
      std::lock_guard<std::mutex> lk(mu_);
      q_.push_back(std::move(fn));
    }
    cv_.notify_one();
  }
  int threads() const { return (int)workers_.size(); }

 private:
  CopyPool() {
    int n = 0;
    if (const char* e = std::getenv("STITCH_COPY_THREADS")) n = std::atoi(e);
    if (n <= 0) {
      const int hw = (int)std::thread::hardware_concurrency();
      n = std::min(8, std::max(2, hw / 2));
    }
    n = std::min(n, 32);
    for (int i = 0; i < n; ++i)
      workers_.emplace_back([this] {
        for (;;) {
          std::function<void()> fn;
          {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [this] { return !q_.empty(); });
            fn = std::move(q_.front());
            q_.pop_front();
          }
          fn();
        }
      });
    for (std::thread& t : workers_) t.detach();
  }
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<std::function<void()>> q_;
  std::vector<std::thread> workers_;
};

// true when `p` is ordinary (pageable) host memory: not registered / allocated through CUDA
inline bool is_pageable(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();   // older runtimes report unregistered memory as an error
    return true;
  }
  return a.type == cudaMemoryTypeUnregistered;
}

// memcpy with non-temporal stores for large blocks: the destination (a pinned ring chunk the DMA engine reads next, or a
// caller's canvas nobody reads soon) should not be pulled into the cache first (read-for-ownership = a third of the
// memory traffic of a plain copy).  Falls back to memcpy without AVX2 or for small blocks.
#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target("avx2"))) inline void copy_stream_avx2(uint8_t* d, const uint8_t* s, size_t n) {
  const size_t head = (32 - ((uintptr_t)d & 31)) & 31;
  if (head) { std::memcpy(d, s, head); d += head; s += head; n -= head; }
  size_t i = 0;
  for (; i + 128 <= n; i += 128) {
    const __m256i a = _mm256_loadu_si256((const __m256i*)(s + i)), b = _mm256_loadu_si256((const __m256i*)(s + i + 32));
    const __m256i c = _mm256_loadu_si256((const __m256i*)(s + i + 64)), e = _mm256_loadu_si256((const __m256i*)(s + i + 96));
    _mm256_stream_si256((__m256i*)(d + i), a);
    _mm256_stream_si256((__m256i*)(d + i + 32), b);
    _mm256_stream_si256((__m256i*)(d + i + 64), c);
    _mm256_stream_si256((__m256i*)(d + i + 96), e);
  }
  _mm_sfence();
  if (i < n) std::memcpy(d + i, s + i, n - i);
}
inline void copy_big(void* d, const void* s, size_t n) {
  static const bool nt = [] {
    const char* e = std::getenv("STITCH_COPY_NT");
    return !(e && e[0] == '0') && __builtin_cpu_supports("avx2");
  }();
  if (nt && n >= (64u << 10)) copy_stream_avx2((uint8_t*)d, (const uint8_t*)s, n);
  else std::memcpy(d, s, n);
}
#else
inline void copy_big(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
#endif

// rows [r0, r1) of a pitched host image -> packed rows of `dpitch` bytes (pad bytes opaque)
inline void copy_rows_in(uint8_t* dst, int64_t dpitch, const uint8_t* src, int64_t spitch, size_t row_bytes, int r0, int r1) {
  if (spitch == dpitch && (int64_t)row_bytes == dpitch) {
    copy_big(dst, src + (int64_t)r0 * spitch, (size_t)(r1 - r0) * (size_t)dpitch);
    return;
  }
  for (int r = r0; r < r1; ++r) {
    uint8_t* d = dst + (int64_t)(r - r0) * dpitch;
    std::memcpy(d, src + (int64_t)r * spitch, row_bytes);
    if ((int64_t)row_bytes < dpitch) std::memset(d + row_bytes, 0xff, (size_t)(dpitch - (int64_t)row_bytes));
  }
}

// Upload ring of one lane: kSlots chunks of kChunk bytes of pinned memory.
struct UploadRing {
  static constexpr size_t kChunk = 4u << 20;
  static constexpr int kSlots = 12;
  uint8_t* base = nullptr;
  cudaEvent_t sent[kSlots] = {};      // the slot's H2D copy has been consumed by the DMA engine
  bool used[kSlots] = {};
  std::atomic<int> ready[kSlots];     // 1: the pool has filled the slot
  uint64_t next = 0;                  // running chunk counter: slot = next % kSlots

  cudaError_t ensure() {
    if (base) return cudaSuccess;
    cudaError_t e = cudaMallocHost((void**)&base, kChunk * kSlots);
    if (e != cudaSuccess) return e;
    for (int i = 0; i < kSlots; ++i) {
      e = cudaEventCreateWithFlags(&sent[i], cudaEventDisableTiming);
      if (e != cudaSuccess) return e;
      ready[i].store(0);
    }
    return cudaSuccess;
  }
  void release() {
    if (base) cudaFreeHost(base);
    base = nullptr;
    for (int i = 0; i < kSlots; ++i)
      if (sent[i]) { cudaEventDestroy(sent[i]); sent[i] = nullptr; }
  }
};

}  // namespace stitch
