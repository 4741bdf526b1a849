// stitch_device_digest_noalloc: the device digest (synth.cuh::digest_kernel, stitch_device_digest) for callers that
// must not allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// No counterpart in the reference.
//
// stitch_device_digest keeps its accumulator in a lane's cudaMalloc'ed scratch block and returns the value through a
// cudaMallocHost'ed word.  This variant allocates NOTHING: its scratch is the file-scope __device__ array below
// (32 KB, held by every GPU that loads the library until the process exits) and the result goes to a device word the
// caller passes.  The sum is the same order-independent 64-bit sum, so both entry points return identical values.
//
// Two launches on the caller's stream: D1 writes one partial sum per block into the scratch (a block owns a fixed chunk
// of kChunkBytes, so the number of partials -- and with it the largest buffer -- is bounded by the array), D2 adds the
// partials up in a fixed order and stores the digest.  Nothing in the scratch has to start at zero: D1 overwrites
// every slot D2 reads; slots beyond this call's block count keep an earlier call's values and are not read.
// EVERY call on a GPU shares the array: two calls that overlap on that GPU (different streams, or threads) race.
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "stitch.h"

namespace {

constexpr int kThreads = 256;
constexpr size_t kChunkBytes = 256u << 10;                                            // bytes one block digests
constexpr int kMaxPartials = (int)(STITCH_NOALLOC_DIGEST_MAX_BYTES / kChunkBytes);   // 4096 slots = 32 KB
static_assert((size_t)kMaxPartials * kChunkBytes == STITCH_NOALLOC_DIGEST_MAX_BYTES, "limit is a whole number of chunks");
constexpr int kMaxDevices = 64;

// The scratch.  In HOST code this name is only a stand-in for the symbol, not a device address: the address each
// device gave its copy comes from cudaGetSymbolAddress (scratch_for below).  Device code may name it directly.
static __device__ __align__(256) unsigned long long g_digest_partials[kMaxPartials];

__device__ __forceinline__ unsigned long long mix(unsigned long long x) {   // == synth.cuh::digest_mix
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// D1: block b digests words [b * chunk_words, +chunk_words) (+ the buffer's tail bytes in the last block) -> partials[b]
__global__ void __launch_bounds__(kThreads) digest_partials_kernel(const uint8_t* __restrict__ p, size_t bytes,
                                                                    unsigned long long* __restrict__ partials) {
  __shared__ unsigned long long warp_sum[kThreads / 32];
  const size_t nw = bytes >> 2;
  const size_t chunk_words = kChunkBytes >> 2;
  const size_t w0 = (size_t)blockIdx.x * chunk_words;
  const size_t w1 = w0 + chunk_words < nw ? w0 + chunk_words : nw;
  const uint32_t* w = reinterpret_cast<const uint32_t*>(p);
  const bool aligned = (reinterpret_cast<uintptr_t>(p) & 3) == 0;
  unsigned long long acc = 0;
  for (size_t i = w0 + threadIdx.x; i < w1; i += kThreads) {
    uint32_t v;
    if (aligned) v = w[i];
    else v = p[4 * i] | (p[4 * i + 1] << 8) | (p[4 * i + 2] << 16) | ((uint32_t)p[4 * i + 3] << 24);
    acc += mix((unsigned long long)v ^ ((unsigned long long)i * 0xD6E8FEB86659FD93ull));
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {   // the 0..3 bytes after the last whole word, and the length
    uint32_t v = 0;
    for (size_t k = nw * 4; k < bytes; ++k) v |= (uint32_t)p[k] << (8 * (k & 3));
    acc += mix((unsigned long long)v ^ ((unsigned long long)nw * 0xD6E8FEB86659FD93ull) ^ ((unsigned long long)bytes << 32));
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, d);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < kThreads / 32; ++k) s += warp_sum[k];
    partials[blockIdx.x] = s;
  }
}

// D2: one block adds partials[0 .. n) and stores the digest
__global__ void __launch_bounds__(kThreads) digest_finish_kernel(const unsigned long long* __restrict__ partials, int n,
                                                                  unsigned long long* __restrict__ out) {
  __shared__ unsigned long long warp_sum[kThreads / 32];
  unsigned long long acc = 0;
  for (int i = threadIdx.x; i < n; i += kThreads) acc += partials[i];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, d);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < kThreads / 32; ++k) s += warp_sum[k];
    *out = s;
  }
}

// Device address of `device`'s copy of the scratch: asked for once per device, with that device current.
cudaError_t scratch_for(int device, unsigned long long** out) {
  static std::mutex mu;
  static unsigned long long* addr[kMaxDevices] = {};
  std::lock_guard<std::mutex> lk(mu);
  if (!addr[device]) {
    unsigned long long* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_digest_partials);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    addr[device] = p;
  }
  *out = addr[device];
  return cudaSuccess;
}

int fail(char* err, size_t cap, int code, const char* fmt, unsigned long long a = 0, unsigned long long b = 0) {
  if (err && cap) std::snprintf(err, cap, fmt, a, b);
  return code;
}

}  // namespace

extern "C" int stitch_device_digest_noalloc(int32_t device, void* stream, const void* dev_ptr, size_t bytes, uint64_t* dev_out,
                                            char* err, size_t err_capacity) {
  if (err && err_capacity) err[0] = 0;
  if (!dev_ptr || !dev_out) return fail(err, err_capacity, STITCH_E_BAD_ARG, "stitch_device_digest_noalloc: null buffer");
  if (device < 0 || device >= kMaxDevices) return fail(err, err_capacity, STITCH_E_BAD_ARG, "stitch_device_digest_noalloc: bad device index");
  if (bytes > (size_t)STITCH_NOALLOC_DIGEST_MAX_BYTES)
    return fail(err, err_capacity, STITCH_E_BAD_ARG,
                "stitch_device_digest_noalloc: %llu bytes exceed the static scratch limit of %llu bytes "
                "(STITCH_NOALLOC_DIGEST_MAX_BYTES); use stitch_device_digest",
                (unsigned long long)bytes, (unsigned long long)STITCH_NOALLOC_DIGEST_MAX_BYTES);
  cudaError_t e = cudaSetDevice(device);
  unsigned long long* scratch = nullptr;
  if (e == cudaSuccess) e = scratch_for(device, &scratch);
  if (e != cudaSuccess) {
    if (err && err_capacity) std::snprintf(err, err_capacity, "stitch_device_digest_noalloc: %s", cudaGetErrorString(e));
    return STITCH_E_CUDA;
  }
  const size_t nw = bytes >> 2;
  int blocks = (int)((nw * 4 + kChunkBytes - 1) / kChunkBytes);
  if (blocks < 1) blocks = 1;   // a buffer of 0..3 bytes still has its tail term
  cudaStream_t s = (cudaStream_t)stream;
  digest_partials_kernel<<<blocks, kThreads, 0, s>>>((const uint8_t*)dev_ptr, bytes, scratch);
  digest_finish_kernel<<<1, kThreads, 0, s>>>(scratch, blocks, (unsigned long long*)dev_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    if (err && err_capacity) std::snprintf(err, err_capacity, "stitch_device_digest_noalloc: %s", cudaGetErrorString(e));
    return STITCH_E_CUDA;
  }
  return STITCH_OK;
}
