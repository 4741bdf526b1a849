// micro-benchmark: issue/pipe rates of FFMA, FFMA2, PRMT and mixes on sm_100a (scratch, not product)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITER 2048
__device__ __forceinline__ void ffma2(float& dx, float& dy, float ax, float ay, float bx, float by) {
  asm volatile("{.reg .b64 ra, rb, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%5};\n mov.b64 rd, {%0,%1};\n fma.rn.f32x2 rd, ra, rb, rd;\n mov.b64 {%0,%1}, rd;}"
      : "+f"(dx), "+f"(dy) : "f"(ax), "f"(ay), "f"(bx), "f"(by));
}
template <int MODE>
__global__ void k(float* out, const uint32_t* in, long long* cyc) {
  float a[16];
  uint32_t p[8];
  float w0 = in[threadIdx.x & 31] * 1e-9f + 0.5f, w1 = w0 + 0.1f;
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = i;
#pragma unroll
  for (int i = 0; i < 8; ++i) p[i] = in[i];
  long long t0 = clock64();
  for (int it = 0; it < ITER; ++it) {
    if (MODE == 0) {  // 16 FFMA
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], w0, w1);
    } else if (MODE == 1) {  // 8 FFMA2 (per-half weights)
#pragma unroll
      for (int i = 0; i < 16; i += 2) ffma2(a[i], a[i + 1], w0, w1, a[(i + 2) & 15], a[(i + 3) & 15]);
    } else if (MODE == 2) {  // 16 PRMT
#pragma unroll
      for (int i = 0; i < 8; ++i) { p[i] = __byte_perm(p[i], 0x3f800000u, 0x7604); p[i] = __byte_perm(p[i], p[(i+1)&7], 0x7614); }
    } else if (MODE == 3) {  // 8 PRMT + 8 FFMA
#pragma unroll
      for (int i = 0; i < 8; ++i) { p[i] = __byte_perm(p[i], p[(i+1)&7], 0x7604); a[i] = fmaf(a[i], w0, w1); }
    } else if (MODE == 4) {  // 8 PRMT + 4 FFMA2
#pragma unroll
      for (int i = 0; i < 8; i += 2) { p[i] = __byte_perm(p[i], p[(i+1)&7], 0x7604); p[i+1] = __byte_perm(p[i+1], p[(i+2)&7], 0x7604);
        ffma2(a[i], a[i + 1], w0, w1, a[i + 8], a[i + 9]); }
    } else if (MODE == 5) {  // 8 FFMA2 broadcast weight
#pragma unroll
      for (int i = 0; i < 16; i += 2) ffma2(a[i], a[i + 1], w0, w0, a[(i + 2) & 15], a[(i + 3) & 15]);
    } else if (MODE == 6) {  // 12 PRMT + 6 FFMA2 + 2 LOP3 : the row-step math mix
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        uint32_t x = __byte_perm(p[i & 3], 0x3f800000u, 0x7604 + ((i & 3) << 4)), y = __byte_perm(p[(i + 1) & 3], 0x3f800000u, 0x7604 + ((i & 3) << 4));
        ffma2(a[2 * i], a[2 * i + 1], w0, w0, __uint_as_float(x), __uint_as_float(y));
      }
      p[4] &= p[0] & p[1]; p[4] &= p[2] & p[3];
      p[0] += 0x01010101u; p[1] ^= p[4]; p[2] += p[1]; p[3] ^= p[0];
    } else if (MODE == 7) {  // 16 IDP2A (dp2a.lo / dp2a.hi alternating)
#pragma unroll
      for (int i = 0; i < 8; ++i) { p[i] = __dp2a_lo(p[(i + 1) & 7], p[(i + 2) & 7], p[i]); p[i] = __dp2a_hi(p[(i + 3) & 7], p[(i + 2) & 7], p[i]); }
    } else if (MODE == 8) {  // 16 IDP4A
#pragma unroll
      for (int i = 0; i < 8; ++i) { p[i] = __dp4a(p[(i + 1) & 7], p[(i + 2) & 7], p[i]); p[i] = __dp4a(p[(i + 3) & 7], p[(i + 2) & 7], p[i]); }
    } else if (MODE == 9) {  // row-PAIR mix of the integer vertical pass for 4 pixels: 8 PRMT + 12 IDP2A + 4 LOP3
      uint32_t acc[12];
#pragma unroll
      for (int i = 0; i < 12; ++i) acc[i] = __float_as_uint(a[i]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t rg = __byte_perm(p[i], p[i + 4], 0x5140), ba = __byte_perm(p[i], p[i + 4], 0x7362);
        acc[3 * i] = __dp2a_lo(p[(i + 1) & 3] | 0x10001u, rg, acc[3 * i]);
        acc[3 * i + 1] = __dp2a_hi(p[(i + 1) & 3] | 0x10001u, rg, acc[3 * i + 1]);
        acc[3 * i + 2] = __dp2a_lo(p[(i + 1) & 3] | 0x10001u, ba, acc[3 * i + 2]);
      }
      p[4] &= p[0] & p[1]; p[5] &= p[2] & p[3]; p[6] ^= p[4]; p[7] ^= p[5];
#pragma unroll
      for (int i = 0; i < 12; ++i) a[i] = __uint_as_float(acc[i]);
    } else if (MODE == 10) {  // 8 IMAD (32-bit) for reference
#pragma unroll
      for (int i = 0; i < 8; ++i) { p[i] = p[(i + 1) & 7] * p[(i + 2) & 7] + p[i]; }
    }
  }
  long long t1 = clock64();
  float s = 0; uint32_t q = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) q ^= p[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + q;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE> void run(const char* name, int ninstr) {
  float* out; uint32_t* in; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&in, 4096); cudaMemset(in, 1, 4096); cudaMalloc(&cyc, 8);
  for (int warps = 4; warps <= 32; warps *= 2) {
    k<MODE><<<148, warps * 32>>>(out, in, cyc);
    k<MODE><<<148, warps * 32>>>(out, in, cyc);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    double per_smsp = (double)ninstr * ITER * (warps / 4.0);
    printf("%-28s warps/SM=%2d  cycles=%lld  issue-cycles/instr/SMSP=%.3f\n", name, warps, c, c / per_smsp);
  }
}
int main() {
  run<0>("16 FFMA", 16); run<1>("8 FFMA2", 8); run<5>("8 FFMA2 bcast", 8); run<2>("16 PRMT", 16);
  run<3>("8 PRMT + 8 FFMA", 16); run<4>("8 PRMT + 4 FFMA2", 12); run<6>("rowstep mix (12P+6F2+~8alu)", 26);
  run<7>("16 IDP2A", 16); run<8>("16 IDP4A", 16); run<9>("int row-PAIR mix (8P+12IDP2A+4L)", 24); run<10>("8 IMAD", 8);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
