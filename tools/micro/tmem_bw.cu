// Microbenchmark (B200): tcgen05.ld / tcgen05.st throughput from 16 warps (4 per SM sub-partition), the softmax
// warp layout of attn_spatial_tc*_kernel: bytes per clock per SM for x32 / x16 shapes, with the wait after every
// load or after a batch of loads.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "../../timelapse_video_interpolation_b200/csrc/tvi_common.cuh"
using namespace tvi;
namespace tvi {
void set_error(const char*, ...) {}
int fail(int c, const char*, ...) { return c; }
}  // namespace tvi

// MODE 0: ld.x32 + wait each; 1: 2 x ld.x32 then wait; 2: ld.x16 + wait; 3: st.x16 + wait; 4: ld.x32 + st.x16 + waits
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(uint32_t* out, long long* cyc, int iters) {
  __shared__ uint32_t s_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    tmem_alloc(&s_base, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t base = s_base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;
  uint32_t r[32], r2[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) r[i] = threadIdx.x + i, r2[i] = i;
  tmem_st_x32(base, r);
  tmem_st_x32(base + 32, r);
  tmem_st_x32(base + 64, r);
  tmem_st_x32(base + 96, r);
  tmem_st_wait();
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const uint32_t col = (it & 1) * 64;
    if (MODE == 0) {
      tmem_ld_x32(base + col, r);
      tmem_ld_wait();
      acc += r[0] + r[31];
    } else if (MODE == 1) {
      tmem_ld_x32(base + col, r);
      tmem_ld_x32(base + col + 32, r2);
      tmem_ld_wait();
      acc += r[0] + r[31] + r2[0] + r2[31];
    } else if (MODE == 2) {
      tmem_ld_x16(base + col, r);
      tmem_ld_wait();
      acc += r[0] + r[15];
    } else if (MODE == 3) {
      r[0] = it;
      tmem_st_x16(base + col, r);
      tmem_st_wait();
    } else {
      tmem_ld_x32(base + col, r);
      tmem_ld_wait();
      acc += r[0] + r[31];
      r2[0] = acc;
      tmem_st_x16(base + col, r2);
      tmem_st_wait();
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(s_base, 512);
  }
}
template <int MODE>
void run(const char* name, double bytes_per_iter_per_warp) {
  uint32_t* d;
  long long* c;
  cudaMalloc(&d, 148 * 512 * 4);
  cudaMalloc(&c, 148 * 8);
  const int iters = 20000;
  k<MODE><<<148, 512>>>(d, c, 100);
  k<MODE><<<148, 512>>>(d, c, iters);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(e)); return; }
  long long h[148];
  cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += (double)h[i];
  avg /= 148;
  printf("%-34s %8.1f clk per iteration per warp; %7.1f B/clk/SM (16 warps)\n", name, avg / iters, 16.0 * bytes_per_iter_per_warp * iters / avg);
  cudaFree(d);
  cudaFree(c);
}
int main() {
  run<0>("ld.x32 + wait", 4096);
  run<1>("2 x ld.x32 + wait", 8192);
  run<2>("ld.x16 + wait", 2048);
  run<3>("st.x16 + wait", 2048);
  run<4>("ld.x32 + wait + st.x16 + wait", 4096 + 2048);
  return 0;
}
