// Microbenchmark: HBM bandwidth of this B200 for write-only and 1-read : 3-write traffic (the mix of the fused
// QKV projection: 165 MB in, 495 MB out) against a 1 : 1 copy, with plain 128-bit coalesced accesses.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__global__ void k_copy(const uint4* __restrict__ a, uint4* __restrict__ b, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
}
__global__ void k_write(uint4* __restrict__ b, size_t n) {
  const uint4 v = make_uint4(1, 2, 3, threadIdx.x);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = v;
}
__global__ void k_r1w3(const uint4* __restrict__ a, uint4* __restrict__ b, size_t n) {  // b has 3n elements
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = a[i];
    b[i] = v;
    b[n + i] = v;
    b[2 * n + i] = v;
  }
}
__global__ void k_read(const uint4* __restrict__ a, uint4* __restrict__ b, size_t n) {
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = a[i];
    acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
  }
  if (acc.x == 0x12345678u) b[0] = acc;
}
template <typename F>
float timeit(F f) {
  f();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) f();
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms / 10;
}
int main() {
  const size_t n = (size_t)32 << 20;  // uint4 elements: 512 MB
  uint4 *a, *b;
  cudaMalloc(&a, n * 16);
  cudaMalloc(&b, 3 * n * 16);
  cudaMemset(a, 1, n * 16);
  const int grid = 148 * 8, blk = 512;
  float ms = timeit([&] { k_copy<<<grid, blk>>>(a, b, n); });
  printf("copy 1:1         %7.0f GB/s\n", 2.0 * n * 16 / ms / 1e6);
  ms = timeit([&] { k_write<<<grid, blk>>>(b, 3 * n); });
  printf("write only       %7.0f GB/s\n", 3.0 * n * 16 / ms / 1e6);
  ms = timeit([&] { k_r1w3<<<grid, blk>>>(a, b, n); });
  printf("read 1 : write 3 %7.0f GB/s\n", 4.0 * n * 16 / ms / 1e6);
  ms = timeit([&] { k_read<<<grid, blk>>>(a, b, n); });
  printf("read only        %7.0f GB/s\n", 1.0 * n * 16 / ms / 1e6);
  return 0;
}
