// scratch micro-benchmark: the compose kernel's banded access pattern (512 B per warp and row, 9 x 4032x3024 RGBA8
// in, 1820x12288 out) staged with 2-D TMA boxes into a per-warp mbarrier ring, against the register / LDGSTS forms.
//   selfissue<R,NS>: every consumer warp owns a ring of NS stages x R rows x 512 B; its lane 0 re-arms the stage it has
//                    just finished and issues ONE cp.async.bulk.tensor.2d box (128 px x R rows) for it.
//   cta<R,NS>      : 8 consumer warps + 1 producer warp per CTA; CTA-wide stages of R rows x 4 KB (four 256-px boxes),
//                    full / empty mbarriers, CTA-level units of 8 adjacent bands.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmabench tmabench.cu
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}"
      ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int x, int y, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}

struct Maps { CUtensorMap m[9]; CUtensorMap one[9]; CUtensorMap wide[9]; };

// ---- per-warp ring, self-issued boxes ----
template <int R, int NS>
__global__ void __launch_bounds__(256, 2) selfissue(const __grid_constant__ Maps maps, uint8_t* __restrict__ out, int rows_total,
                                                    int rows_per_unit, int bands, int ratio_num, int ratio_den, unsigned* ctr, int nunits) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t* ring = smem + (size_t)warp * (NS * R * 512 + 128);
  const uint32_t ring_s = smem_u32(ring), bar_s = ring_s + NS * R * 512;
  if (lane == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(bar_s + 8 * s, 1);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();
  uint32_t phases = 0;
  for (;;) {
    unsigned u = 0; if (lane == 0) u = atomicAdd(ctr, 1u); u = __shfl_sync(~0u, u, 0);
    if (u >= (unsigned)nunits) break;
    const int b = u % bands; const unsigned t = u / bands; const int ny = (rows_total + rows_per_unit - 1) / rows_per_unit;
    const int ty = t % ny, img = t / ny;
    const int r0 = ty * rows_per_unit, r1 = min(r0 + rows_per_unit, rows_total);
    const int nrows = r1 - r0, nst = (nrows + R - 1) / R;
    const int x = b * 124;  // 31 groups of 4 px
    auto issue = [&](int st) {   // lane 0 only
      const int slot = st % NS, y = r0 + st * R, k = min(R, nrows - st * R);
      const uint32_t bar = bar_s + 8 * slot, dst = ring_s + slot * (R * 512);
      if (k == R) { mbar_expect_tx(bar, R * 512); tma_load_2d(dst, &maps.m[img], x, y, bar); }
      else { mbar_expect_tx(bar, k * 512); for (int j = 0; j < k; ++j) tma_load_2d(dst + j * 512, &maps.one[img], x, y + j, bar); }
    };
    if (lane == 0) for (int s = 0; s < NS && s < nst; ++s) issue(s);
    uint4 acc = make_uint4(0, 0, 0, 0); int e = 0;
    size_t orow = ((size_t)img * rows_total + r0) * ratio_den / ratio_num;
    for (int st = 0; st < nst; ++st) {
      const int slot = st % NS;
      mbar_wait(bar_s + 8 * slot, (phases >> slot) & 1u);
      phases ^= 1u << slot;
      const uint32_t base = ring_s + slot * (R * 512) + lane * 16;
      const int k = min(R, nrows - st * R);
#pragma unroll
      for (int j = 0; j < R; ++j) {
        if (j < k) {
          const uint4 v = lds128(base + j * 512);
          acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
          e += ratio_den;
          if (e >= ratio_num) { e -= ratio_num; if (lane < 14) *(uint4*)(out + orow * 7280 + (size_t)b * 224 + lane * 16) = acc; ++orow; }
        }
      }
      __syncwarp();
      if (lane == 0 && st + NS < nst) issue(st + NS);
    }
  }
}

// ---- CTA-wide stages, dedicated producer warp ----
template <int R, int NS>
__global__ void __launch_bounds__(288, 2) ctawide(const __grid_constant__ Maps maps, uint8_t* __restrict__ out, int rows_total,
                                                  int rows_per_unit, int cbands, int ratio_num, int ratio_den, unsigned* ctr, int nunits) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ unsigned s_unit[2];
  __shared__ __align__(8) uint64_t bars[2 * NS + 2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t ring_s = smem_u32(smem), full_s = smem_u32(bars), empty_s = full_s + 8 * NS;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(full_s + 8 * s, 1); mbar_init(empty_s + 8 * s, 8); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  // one ticket per CTA unit; the producer draws it and publishes it through smem + a named barrier per unit
  uint32_t phases_full = 0, phases_empty = 0;
  int stage_ctr = 0;  // running stage counter (same sequence in producer and consumers)
  for (;;) {
    if (threadIdx.x == 256) s_unit[0] = atomicAdd(ctr, 1u);
    __syncthreads();
    const unsigned u = s_unit[0];
    __syncthreads();
    if (u >= (unsigned)nunits) break;
    const int cb = u % cbands; const unsigned t = u / cbands; const int ny = (rows_total + rows_per_unit - 1) / rows_per_unit;
    const int ty = t % ny, img = t / ny;
    const int r0 = ty * rows_per_unit, r1 = min(r0 + rows_per_unit, rows_total);
    const int nrows = r1 - r0, nst = (nrows + R - 1) / R;
    const int x0 = cb * 8 * 124;
    if (warp == 8) {
      if (lane == 0) {
        for (int st = 0; st < nst; ++st) {
          const int slot = (stage_ctr + st) % NS;
          if (stage_ctr + st >= NS) { mbar_wait(empty_s + 8 * slot, (phases_empty >> slot) & 1u); phases_empty ^= 1u << slot; }
          const uint32_t bar = full_s + 8 * slot, dst = ring_s + slot * (R * 4096);
          mbar_expect_tx(bar, R * 4096);
#pragma unroll
          for (int q = 0; q < 4; ++q) tma_load_2d(dst + q * (R * 1024), &maps.wide[img], x0 + q * 256, r0 + st * R, bar);
        }
      }
      stage_ctr += nst;
      phases_empty = __shfl_sync(~0u, phases_empty, 0);
    } else {
      const int b = cb * 8 + warp;
      const int gi = warp * 31 + lane;          // group inside the CTA band
      const uint32_t lbase = ring_s + (gi >> 6) * (R * 1024) + (gi & 63) * 16;
      uint4 acc = make_uint4(0, 0, 0, 0); int e = 0;
      size_t orow = ((size_t)img * rows_total + r0) * ratio_den / ratio_num;
      for (int st = 0; st < nst; ++st) {
        const int slot = (stage_ctr + st) % NS;
        mbar_wait(full_s + 8 * slot, (phases_full >> slot) & 1u);
        phases_full ^= 1u << slot;
        const uint32_t base = lbase + slot * (R * 4096);
        const int k = min(R, nrows - st * R);
#pragma unroll
        for (int j = 0; j < R; ++j) {
          if (j < k) {
            const uint4 v = lds128(base + j * 1024);
            acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
            e += ratio_den;
            if (e >= ratio_num) { e -= ratio_num; if (lane < 14 && b < 33) *(uint4*)(out + orow * 7280 + (size_t)b * 224 + lane * 16) = acc; ++orow; }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_s + 8 * slot);
      }
      stage_ctr += nst;
    }
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int W = 4032, H = 3024, N = 9;
  const size_t img = (size_t)W * H * 4, nin = img * N, nout = (size_t)1820 * 12288 * 4;
  uint8_t* in[3]; uint8_t* out; unsigned* ctr;
  for (int i = 0; i < 3; ++i) { CK(cudaMalloc(&in[i], nin)); CK(cudaMemset(in[i], i + 1, nin)); }
  CK(cudaMalloc(&out, nout + (1 << 20))); CK(cudaMalloc(&ctr, 4));
  EncodeFn encode = nullptr; cudaDriverEntryPointQueryResult qres;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qres));
  if (!encode) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
  Maps maps[3][3];  // [input set][box rows variant: 0 -> R=8, 1 -> R=4, 2 -> R=16]
  const int RR[3] = {8, 4, 16};
  for (int s = 0; s < 3; ++s)
    for (int v = 0; v < 3; ++v)
      for (int i = 0; i < N; ++i) {
        cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H}, strides[1] = {(cuuint64_t)W * 4};
        cuuint32_t es[2] = {1, 1};
        void* base = in[s] + (size_t)i * img;
        cuuint32_t box[2] = {128, (cuuint32_t)RR[v]};
        CUresult r = encode(&maps[s][v].m[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        cuuint32_t box1[2] = {128, 1};
        CUresult r1 = encode(&maps[s][v].one[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, base, dims, strides, box1, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        cuuint32_t boxw[2] = {256, (cuuint32_t)RR[v]};
        CUresult r2 = encode(&maps[s][v].wide[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, base, dims, strides, boxw, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r || r1 || r2) { printf("encode failed %d %d %d\n", (int)r, (int)r1, (int)r2); return 1; }
      }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int IT = 30, bands = 33;
  auto timeit = [&](const char* name, auto go) -> int {
    for (int w = 0; w < 3; ++w) go(w);
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0); for (int it = 0; it < IT; ++it) go(it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-56s %8.2f us  %7.1f GB/s\n", name, ms * 1000 / IT, ((double)nin + nout) * IT / ms / 1e6);
    return 0;
  };
#define SELF(R_, NS_, V_)                                                                                                   \
  for (int rpu : {140, 72}) {                                                                                              \
    const int ny = (H + rpu - 1) / rpu, nunits = bands * ny * N;                                                           \
    const int smem = 8 * (NS_ * R_ * 512 + 128);                                                                           \
    CK(cudaFuncSetAttribute(selfissue<R_, NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));                        \
    int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, selfissue<R_, NS_>, 256, smem);                        \
    char nm[128]; snprintf(nm, 128, "self-issue TMA R=%d NS=%d, %d rows/unit, %d CTA/SM", R_, NS_, rpu, occ);              \
    if (timeit(nm, [&](int it) { cudaMemsetAsync(ctr, 0, 4);                                                               \
          selfissue<R_, NS_><<<148 * occ, 256, smem>>>(maps[it % 3][V_], out, H, rpu, bands, 3024, 1365, ctr, nunits); })) return 1; \
  }
  SELF(8, 3, 0) SELF(8, 2, 0) SELF(4, 4, 1) SELF(4, 6, 1) SELF(16, 2, 2)
#define CTAW(R_, NS_, V_)                                                                                                   \
  for (int rpu : {140, 72}) {                                                                                              \
    const int cb = (bands + 7) / 8, ny = (H + rpu - 1) / rpu, nunits = cb * ny * N;                                         \
    const int smem = NS_ * R_ * 4096;                                                                                      \
    CK(cudaFuncSetAttribute(ctawide<R_, NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));                          \
    int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ctawide<R_, NS_>, 288, smem);                          \
    char nm[128]; snprintf(nm, 128, "CTA-wide TMA R=%d NS=%d, %d rows/unit, %d CTA/SM (5th band col 1/8 full)", R_, NS_, rpu, occ); \
    if (timeit(nm, [&](int it) { cudaMemsetAsync(ctr, 0, 4);                                                               \
          ctawide<R_, NS_><<<148 * occ, 288, smem>>>(maps[it % 3][V_], out, H, rpu, cb, 3024, 1365, ctr, nunits); })) return 1; \
  }
  CTAW(8, 3, 0) CTAW(4, 4, 1) CTAW(4, 6, 1) CTAW(8, 4, 0)
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
