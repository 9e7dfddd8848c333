// H2D rate of one tick's worth of keypoints from page-locked host memory: cudaHostAlloc default / write-combined /
// cudaHostRegister of malloc'ed pages, one big copy vs the same bytes in 4 pieces. nvcc -O2 -o /tmp/h2d_probe profiles/h2d_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
static double rate(void* dst, const void* src, size_t n, int pieces, int reps) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int w = 0; w < 3; ++w) cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, 0);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  for (int r = 0; r < reps; ++r)
    for (int p = 0; p < pieces; ++p)
      cudaMemcpyAsync((char*)dst + p * (n / pieces), (const char*)src + p * (n / pieces), n / pieces, cudaMemcpyHostToDevice, 0);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return (double)n * reps / (ms / 1e3) / 1e9;
}
int main() {
  const size_t n = 36836480;
  void *d, *h0, *h1, *h2;
  cudaMalloc(&d, n);
  cudaHostAlloc(&h0, n, cudaHostAllocDefault);
  cudaHostAlloc(&h1, n, cudaHostAllocWriteCombined);
  h2 = aligned_alloc(4096, n); memset(h2, 1, n); cudaHostRegister(h2, n, cudaHostRegisterDefault);
  memset(h0, 1, n); memset(h1, 1, n);
  for (int round = 0; round < 2; ++round) {
    printf("default pinned : 1 copy %.1f GB/s, 4 pieces %.1f GB/s\n", rate(d, h0, n, 1, 50), rate(d, h0, n, 4, 50));
    printf("write-combined : 1 copy %.1f GB/s, 4 pieces %.1f GB/s\n", rate(d, h1, n, 1, 50), rate(d, h1, n, 4, 50));
    printf("host-registered: 1 copy %.1f GB/s, 4 pieces %.1f GB/s\n", rate(d, h2, n, 1, 50), rate(d, h2, n, 4, 50));
  }
  return 0;
}
