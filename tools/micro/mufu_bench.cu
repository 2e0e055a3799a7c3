// Microbenchmark: per-SM throughput of ex2.approx in f32, f16x2 and bf16x2 form (B200).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
template <int MODE>
__global__ void k(uint32_t* out, int iters) {
  uint32_t a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = 0x3c003c00u + threadIdx.x + i;  // ~1.0 in f16x2
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(a[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(a[i]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(a[i]));
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name) {
  uint32_t* d;
  cudaMalloc(&d, 148 * 1024 * 4);
  const int iters = 4096;
  k<MODE><<<148, 1024>>>(d, 16);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<148, 1024>>>(d, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double ops = 148.0 * 1024 * iters * 8;  // lane-instructions
  printf("%s: %.3f ms, %.1f lane-instr/ns total, %.2f lane-instr/clk/SM at 1.9 GHz\n", name, ms, ops / ms / 1e6,
         ops / ms / 1e6 / 148 / 1.9);
  cudaFree(d);
}
int main() {
  run<0>("ex2.f32   ");
  run<1>("ex2.f16x2 ");
  run<2>("ex2.bf16x2");
  return 0;
}
