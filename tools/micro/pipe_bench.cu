// Microbenchmark: per-SM issue rate of the non-MUFU instructions of the softmax inner loop (B200):
// cvt.rn.bf16x2.f32 (F2FP), prmt, 3-input fmax, packed fp32 fma / add, and F2FP interleaved with MUFU.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
template <int MODE>
__global__ void k(uint32_t* out, int iters) {
  uint32_t a[8];
  float f[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    a[i] = 0x3f800000u + threadIdx.x * 8 + i;
    f[i] = 1.0f + 0.001f * (threadIdx.x + i);
  }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(a[i]) : "f"(f[i]), "f"(__uint_as_float(a[i])));
      if (MODE == 1) asm volatile("prmt.b32 %0, %0, %1, 0x7632;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]));
      if (MODE == 2) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(f[(i + 1) & 7]), "f"(f[(i + 2) & 7]));
      if (MODE == 3) {
        uint64_t x, y;
        asm volatile("mov.b64 %0, {%1,%2};" : "=l"(x) : "f"(f[i]), "f"(f[(i + 1) & 7]));
        asm volatile("fma.rn.f32x2 %0, %1, %1, %1;" : "=l"(y) : "l"(x));
        asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(f[i]), "=f"(f[(i + 1) & 7]) : "l"(y));
      }
      if (MODE == 4) {  // F2FP + MUFU together: do they share a pipe?
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(a[i]) : "f"(f[i]), "f"(__uint_as_float(a[i])));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
      }
      if (MODE == 5) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]));
      if (MODE == 6) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= a[i] ^ __float_as_uint(f[i]);
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
  double ops = 148.0 * 1024 * iters * 8;
  printf("%s: %.3f ms, %.2f lane-iterations/clk/SM at 1.9 GHz\n", name, ms, ops / ms / 1e6 / 148 / 1.9);
  cudaFree(d);
}
int main() {
  run<0>("cvt.rn.bf16x2.f32 (F2FP)   ");
  run<1>("prmt                       ");
  run<2>("max.f32 3-input            ");
  run<3>("fma.rn.f32x2               ");
  run<4>("F2FP + ex2 per iteration   ");
  run<5>("add.u32                    ");
  run<6>("ex2.approx.ftz.f32         ");
  return 0;
}
