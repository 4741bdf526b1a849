// scratch micro-benchmark: HBM ceilings for the compose kernel's read/write mix and access patterns
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("%s: %s\n",#x,cudaGetErrorString(e)); return 1;}}while(0)
// linear: every thread grid-strides over uint4; reads R bytes, writes W bytes (1 store per `ratio` loads)
__global__ void lin(const uint4* __restrict__ in, uint4* __restrict__ out, size_t nin, int ratio) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  uint4 acc = make_uint4(0,0,0,0); int c = 0; size_t o = i;
  for (; i < nin; i += stride) {
    uint4 v = __ldcs(in + i); acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
    if (++c == ratio) { out[o] = acc; o += stride; c = 0; }
  }
  if (acc.x == 0x12345) out[0] = acc;
}
// banded: warp w owns a 512-byte column band of an image (pitch bytes per row) and streams down `rows` rows
// with UNR independent loads in flight; writes 16B per lane every `ratio` rows... (approximates our kernel)
template <int UNR>
__global__ void band(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int pitch, int rows_total, int rows_per_unit,
                     int bands, int nimg, size_t img_bytes, int ratio_num, int ratio_den, unsigned* ctr, int nunits) {
  int lane = threadIdx.x & 31;
  for (;;) {
    unsigned u = 0; if (lane == 0) u = atomicAdd(ctr, 1u); u = __shfl_sync(~0u, u, 0);
    if (u >= (unsigned)nunits) break;
    int b = u % bands; unsigned t = u / bands; int ny = (rows_total + rows_per_unit - 1) / rows_per_unit;
    int ty = t % ny, img = t / ny;
    int r0 = ty * rows_per_unit, r1 = min(r0 + rows_per_unit, rows_total);
    const uint8_t* p = in + img * img_bytes + (size_t)r0 * pitch + (size_t)b * 496 + lane * 16;
    uint4 acc = make_uint4(0,0,0,0); int e = 0;
    size_t orow = ((size_t)img * rows_total + r0) * ratio_den / ratio_num;
    for (int r = r0; r < r1; r += UNR) {
      uint4 v[UNR];
#pragma unroll
      for (int k = 0; k < UNR; ++k) v[k] = __ldcs((const uint4*)(p + (size_t)min(k, r1 - 1 - r) * pitch));
      p += (size_t)UNR * pitch;
#pragma unroll
      for (int k = 0; k < UNR; ++k) { acc.x ^= v[k].x; acc.y ^= v[k].y; acc.z ^= v[k].z; acc.w ^= v[k].w;
        e += ratio_den; if (e >= ratio_num) { e -= ratio_num;
          if (lane < 14) *(uint4*)(out + orow * 7280 + (size_t)b * 224 + lane * 16) = acc; ++orow; } }
    }
  }
}

// CTA-granular units: the 8 warps of a CTA take 8 consecutive band units and stay in lockstep (barrier every SYNC rows)
template <int UNR, int SYNC>
__global__ void band_cta(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int pitch, int rows_total, int rows_per_unit,
                     int bands, int nimg, size_t img_bytes, int ratio_num, int ratio_den, unsigned* ctr, int nunits) {
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ unsigned su;
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) su = atomicAdd(ctr, 8u);
    __syncthreads();
    unsigned u0 = su;
    if (u0 >= (unsigned)nunits) break;
    unsigned u = u0 + warp; bool live = u < (unsigned)nunits; if (!live) u = nunits - 1;
    int b = u % bands; unsigned t = u / bands; int ny = (rows_total + rows_per_unit - 1) / rows_per_unit;
    int ty = t % ny, img = t / ny;
    int r0 = ty * rows_per_unit, r1 = min(r0 + rows_per_unit, rows_total);
    const uint8_t* p = in + img * img_bytes + (size_t)r0 * pitch + (size_t)b * 496 + lane * 16;
    uint4 acc = make_uint4(0,0,0,0); int e = 0;
    size_t orow = ((size_t)img * rows_total + r0) * ratio_den / ratio_num;
    for (int r = r0; r < r0 + rows_per_unit; r += UNR) {
      if (SYNC && ((r - r0) % SYNC) == 0) __syncthreads();
      if (r < r1 && live) {
      uint4 v[UNR];
#pragma unroll
      for (int k = 0; k < UNR; ++k) v[k] = __ldcs((const uint4*)(p + (size_t)min(k, r1 - 1 - r) * pitch));
      p += (size_t)UNR * pitch;
#pragma unroll
      for (int k = 0; k < UNR; ++k) { acc.x ^= v[k].x; acc.y ^= v[k].y; acc.z ^= v[k].z; acc.w ^= v[k].w;
        e += ratio_den; if (e >= ratio_num) { e -= ratio_num;
          if (lane < 14) *(uint4*)(out + orow * 7280 + (size_t)b * 224 + lane * 16) = acc; ++orow; } }
      }
    }
  }
}
// wide bands: each lane reads NG x 16 B per row (warp covers NG*512 contiguous bytes per row)
template <int UNR, int NG>
__global__ void band_wide(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int pitch, int rows_total, int rows_per_unit,
                     int bands, int nimg, size_t img_bytes, int ratio_num, int ratio_den, unsigned* ctr, int nunits) {
  int lane = threadIdx.x & 31;
  for (;;) {
    unsigned u = 0; if (lane == 0) u = atomicAdd(ctr, 1u); u = __shfl_sync(~0u, u, 0);
    if (u >= (unsigned)nunits) break;
    int b = u % bands; unsigned t = u / bands; int ny = (rows_total + rows_per_unit - 1) / rows_per_unit;
    int ty = t % ny, img = t / ny;
    int r0 = ty * rows_per_unit, r1 = min(r0 + rows_per_unit, rows_total);
    size_t colb = (size_t)b * (NG * 512 - 16); if (colb + NG * 512 > (size_t)pitch) colb = pitch - NG * 512;
    const uint8_t* p = in + img * img_bytes + (size_t)r0 * pitch + colb + lane * 16;
    uint4 acc = make_uint4(0,0,0,0); int e = 0;
    size_t orow = ((size_t)img * rows_total + r0) * ratio_den / ratio_num;
    for (int r = r0; r < r1; r += UNR) {
      uint4 v[UNR][NG];
#pragma unroll
      for (int k = 0; k < UNR; ++k)
#pragma unroll
        for (int g = 0; g < NG; ++g) v[k][g] = __ldcs((const uint4*)(p + (size_t)min(k, r1 - 1 - r) * pitch + g * 512));
      p += (size_t)UNR * pitch;
#pragma unroll
      for (int k = 0; k < UNR; ++k) {
#pragma unroll
        for (int g = 0; g < NG; ++g) { acc.x ^= v[k][g].x; acc.y ^= v[k][g].y; acc.z ^= v[k][g].z; acc.w ^= v[k][g].w; }
        e += ratio_den; if (e >= ratio_num) { e -= ratio_num;
          if (lane < 14 * NG) *(uint4*)(out + orow * 7280 + (colb * 1820 / 4032 / 16) * 16 + lane * 16) = acc; ++orow; } }
    }
  }
}
int main() {
  const int W = 4032, H = 3024, N = 9; const size_t img = (size_t)W * H * 4, nin = img * N, nout = (size_t)1820 * 12288 * 4;
  uint8_t *in[3], *out; unsigned* ctr;
  for (int i = 0; i < 3; ++i) { CK(cudaMalloc(&in[i], nin)); CK(cudaMemset(in[i], i + 1, nin)); }
  CK(cudaMalloc(&out, nout + (1 << 20))); CK(cudaMalloc(&ctr, 4));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto report = [&](const char* name, float ms, int it, double bytes) { printf("%-44s %8.2f us  %7.1f GB/s\n", name, ms * 1000 / it, bytes * it / ms / 1e6); };
  const int IT = 30;
  for (int blocks : {148 * 4, 148 * 8, 148 * 16}) {
    for (int ratio : {5, 1000000}) {
      for (int w = 0; w < 3; ++w) lin<<<blocks, 256>>>((uint4*)in[w], (uint4*)out, nin / 16, ratio);
      cudaEventRecord(e0);
      for (int it = 0; it < IT; ++it) lin<<<blocks, 256>>>((uint4*)in[it % 3], (uint4*)out, nin / 16, ratio);
      cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); float ms; cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "linear %d blocks, %s", blocks, ratio == 5 ? "read 439MB + write 88MB" : "read only");
      report(nm, ms, IT, (double)nin + (ratio == 5 ? nin / 5.0 : 0));
    }
  }
  const int bands = 33;
  for (int rpu : {140, 280, 70}) {
    for (int occ : {16, 24, 32}) {
      int ny = (H + rpu - 1) / rpu, nunits = bands * ny * N;
      auto go = [&](int it) { cudaMemsetAsync(ctr, 0, 4); band<8><<<148 * occ / 8, 256>>>(in[it % 3], out, W * 4, H, rpu, bands, N, img, 3024, 1365, ctr, nunits); };
      for (int w = 0; w < 3; ++w) go(w);
      cudaEventRecord(e0); for (int it = 0; it < IT; ++it) go(it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "banded 512B x %d rows/unit, %d warps/SM", rpu, occ);
      report(nm, ms, IT, (double)nin + nout);
    }
  }

  for (int rpu : {140, 72}) {
    for (int occ : {16, 24}) {
      int ny = (H + rpu - 1) / rpu, nunits = bands * ny * N;
      { auto go = [&](int it) { cudaMemsetAsync(ctr, 0, 4); band_cta<8, 8><<<148 * occ / 8, 256>>>(in[it % 3], out, W * 4, H, rpu, bands, N, img, 3024, 1365, ctr, nunits); };
      for (int w = 0; w < 3; ++w) go(w);
      cudaEventRecord(e0); for (int it = 0; it < IT; ++it) go(it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "banded CTA-lockstep(8 rows) %d rows/unit, %d w/SM", rpu, occ);
      report(nm, ms, IT, (double)nin + nout); }
      { auto go = [&](int it) { cudaMemsetAsync(ctr, 0, 4); band_cta<8, 0><<<148 * occ / 8, 256>>>(in[it % 3], out, W * 4, H, rpu, bands, N, img, 3024, 1365, ctr, nunits); };
      for (int w = 0; w < 3; ++w) go(w);
      cudaEventRecord(e0); for (int it = 0; it < IT; ++it) go(it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "banded CTA-units no barrier %d rows/unit, %d w/SM", rpu, occ);
      report(nm, ms, IT, (double)nin + nout); }
    }
  }
  for (int rpu : {140, 72}) {
    for (int occ : {8, 16}) {
      { const int NG = 2; int bw = (W * 4 + NG * 512 - 17) / (NG * 512 - 16); int ny = (H + rpu - 1) / rpu, nunits = bw * ny * N;
      auto go = [&](int it) { cudaMemsetAsync(ctr, 0, 4); band_wide<4, NG><<<148 * occ / 8, 256>>>(in[it % 3], out, W * 4, H, rpu, bw, N, img, 3024, 1365, ctr, nunits); };
      for (int w = 0; w < 3; ++w) go(w);
      cudaEventRecord(e0); for (int it = 0; it < IT; ++it) go(it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "wide 1KB x %d rows/unit, %d w/SM", rpu, occ);
      report(nm, ms, IT, (double)nin + nout); }
      { const int NG = 4; int bw = (W * 4 + NG * 512 - 17) / (NG * 512 - 16); int ny = (H + rpu - 1) / rpu, nunits = bw * ny * N;
      auto go = [&](int it) { cudaMemsetAsync(ctr, 0, 4); band_wide<2, NG><<<148 * occ / 8, 256>>>(in[it % 3], out, W * 4, H, rpu, bw, N, img, 3024, 1365, ctr, nunits); };
      for (int w = 0; w < 3; ++w) go(w);
      cudaEventRecord(e0); for (int it = 0; it < IT; ++it) go(it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "wide 2KB x %d rows/unit, %d w/SM", rpu, occ);
      report(nm, ms, IT, (double)nin + nout); }
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
