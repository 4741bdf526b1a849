// K0: on-device synthetic bitmaps (test / bench helper).  Integer-only so the
// tests can reproduce it exactly on the CPU (tests/util.py synth_image).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace stitch {

// ---- K0: on-device synthetic bitmaps (integer-only so tests/oracle can reproduce it exactly) ----
__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// pattern 0: smooth triangle waves + +-32 noise, opaque.  1: uniform noise, opaque.
// 2: uniform noise with random alpha.  3: 1-px checkerboard.  4: horizontal ramp.
__host__ __device__ inline uint32_t synth_pixel(int x, int y, int width, uint64_t seed, int pattern) {
  const uint64_t h = splitmix64(seed ^ ((uint64_t)y * (uint64_t)width + (uint64_t)x));
  uint32_t c[4];
  if (pattern == 1 || pattern == 2) {
    for (int k = 0; k < 4; ++k) c[k] = (uint32_t)(h >> (8 * k)) & 0xff;
    if (pattern == 1) c[3] = 255;
  } else if (pattern == 3) {
    const uint32_t v = ((x ^ y) & 1) ? 255u : 0u;
    c[0] = c[1] = c[2] = v;
    c[3] = 255;
  } else if (pattern == 4) {
    const uint32_t v = (uint32_t)((x * 255LL) / (width > 1 ? width - 1 : 1));
    c[0] = v;
    c[1] = 255 - v;
    c[2] = (uint32_t)(y & 255);
    c[3] = 255;
  } else {
    for (int k = 0; k < 3; ++k) {
      const int t = (x * 3 + y * 5 + 40 * k) & 511;
      const int tri = t < 256 ? t : 511 - t;
      const int noise = (int)((h >> (8 * k)) & 63) - 32;
      int v = 128 + ((tri - 128) * 3) / 4 + noise;
      v = v < 0 ? 0 : (v > 255 ? 255 : v);
      c[k] = (uint32_t)v;
    }
    c[3] = 255;
  }
  return c[0] | (c[1] << 8) | (c[2] << 16) | (c[3] << 24);
}

__global__ void __launch_bounds__(256) generate_kernel(uint8_t* dst, int width, int height, int64_t pitch, uint64_t seed,
                                                       int pattern) {
  const int x = blockIdx.x * 256 + threadIdx.x;
  const int y = blockIdx.y;
  if (x >= width || y >= height) return;
  reinterpret_cast<uint32_t*>(dst + (int64_t)y * pitch)[x] = synth_pixel(x, y, width, seed, pattern);
}

inline cudaError_t launch_generate(uint8_t* dst, int width, int height, int64_t pitch, uint64_t seed, int pattern,
                                   cudaStream_t stream) {
  dim3 grid((width + 255) / 256, height);
  generate_kernel<<<grid, 256, 0, stream>>>(dst, width, height, pitch, seed, pattern);
  return cudaGetLastError();
}

// 64-bit digest of a device buffer: sum over its bytes' 32-bit words of splitmix64(word ^ position * K) -- order
// independent, so any grid computes it; tests and the batch sweep compare digests of canvases produced on different
// GPUs instead of copying 89 MB per job back.  CPU twin: tests/util.py::digest.
__device__ __forceinline__ uint64_t digest_mix(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__global__ void __launch_bounds__(256) digest_kernel(const uint8_t* __restrict__ p, size_t bytes, unsigned long long* out) {
  const size_t nw = bytes >> 2;
  uint64_t acc = 0;
  const uint32_t* w = reinterpret_cast<const uint32_t*>(p);
  const bool aligned = (reinterpret_cast<uintptr_t>(p) & 3) == 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nw; i += (size_t)gridDim.x * blockDim.x) {
    uint32_t v;
    if (aligned) v = w[i];
    else v = p[4 * i] | (p[4 * i + 1] << 8) | (p[4 * i + 2] << 16) | ((uint32_t)p[4 * i + 3] << 24);
    acc += digest_mix((uint64_t)v ^ ((uint64_t)i * 0xD6E8FEB86659FD93ull));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    uint32_t v = 0;
    for (size_t k = nw * 4; k < bytes; ++k) v |= (uint32_t)p[k] << (8 * (k & 3));
    acc += digest_mix((uint64_t)v ^ ((uint64_t)nw * 0xD6E8FEB86659FD93ull) ^ ((uint64_t)bytes << 32));
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, d);
  if ((threadIdx.x & 31) == 0) atomicAdd(out, (unsigned long long)acc);
}

}  // namespace stitch
