// scratch micro-benchmark: host-to-device copy rate of the e2e leg's upload (9 x 48.8 MB) from pinned memory,
// default vs write-combined allocation, alone and with a concurrent 89 MB device-to-host copy.
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("%s: %s\n",#x,cudaGetErrorString(e)); return 1;}}while(0)
int main() {
  const size_t img = (size_t)4032 * 3024 * 4, nin = img * 9, nout = (size_t)1820 * 12288 * 4;
  uint8_t *d_in, *d_out, *h_out; CK(cudaMalloc(&d_in, nin)); CK(cudaMalloc(&d_out, nout)); CK(cudaHostAlloc(&h_out, nout, cudaHostAllocDefault));
  cudaStream_t s1, s2; cudaStreamCreate(&s1); cudaStreamCreate(&s2);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int wc = 0; wc < 2; ++wc) {
    uint8_t* h_in; CK(cudaHostAlloc(&h_in, nin, wc ? cudaHostAllocWriteCombined : cudaHostAllocDefault));
    memset(h_in, 7, nin);
    for (int both = 0; both < 2; ++both) {
      for (int chunk = 0; chunk < 2; ++chunk) {
        const int IT = 10;
        for (int it = -2; it < IT; ++it) {
          if (it == 0) cudaEventRecord(e0, s1);
          if (chunk) for (int i = 0; i < 9; ++i) cudaMemcpyAsync(d_in + i * img, h_in + i * img, img, cudaMemcpyHostToDevice, s1);
          else cudaMemcpyAsync(d_in, h_in, nin, cudaMemcpyHostToDevice, s1);
          if (both) cudaMemcpyAsync(h_out, d_out, nout, cudaMemcpyDeviceToHost, s2);
        }
        cudaEventRecord(e1, s1); CK(cudaStreamSynchronize(s1)); CK(cudaStreamSynchronize(s2));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("%s, %s, %s: %.3f ms per 439 MB = %.1f GB/s\n", wc ? "write-combined" : "default pinned", chunk ? "9 copies" : "1 copy",
               both ? "with concurrent D2H" : "H2D alone", ms / IT, nin * IT / ms / 1e6);
      }
    }
    cudaFreeHost(h_in);
  }
  return 0;
}
