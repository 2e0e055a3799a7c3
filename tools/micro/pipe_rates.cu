// Microbenchmark (B200, sm_100a): issue rate per SM sub-partition of the instructions the softmax role of
// attn_spatial_tc_kernel is built from, alone and in the mixes the loop uses.  Every kernel runs 148 CTAs x
// 16 warps (the kernel's softmax warp count) and times itself with clock64, so the result is in
// cycles per warp-instruction per sub-partition, independent of the SM clock.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu && ./pipe_rates
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define OPS_PER_ITER 16

enum Mode {
  kFfma = 0, kFfma2, kFadd2, kFfmaImm, kImadShl, kFmnmx, kF2fp, kMufu, kLop3, kFfmaSat, kIadd,
  kMufu_Ffma2_1to1, kMufu_Ffma2_1to2, kMufu_Ffma2_1to4, kMufu_F2fp, kFfma2_F2fp, kFfma2_Fmnmx, kFfma2_Imad,
  kMufu_Ffma2_F2fp, kNumModes
};
static const char* kNames[kNumModes] = {
  "FFMA (3 reg)", "FFMA2", "FADD2", "FFMA (imm)", "IMAD (x*2^23+y)", "FMNMX", "F2FP.BF16 pack", "MUFU.EX2", "LOP3", "FFMA.SAT", "IADD3",
  "MUFU + FFMA2 (1:1)", "MUFU + 2 FFMA2", "MUFU + 4 FFMA2", "MUFU + F2FP (1:1)", "FFMA2 + F2FP (1:1)", "FFMA2 + FMNMX (1:1)", "FFMA2 + IMAD (1:1)",
  "MUFU + 2 FFMA2 + F2FP"
};
// warp-instructions per inner unit, for the report
static const int kInstrPerUnit[kNumModes] = {1,1,1,1,1,1,1,1,1,1,1, 2,3,5,2,2,2,2, 4};

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(uint32_t* out, long long* cyc, int iters) {
  float f[8];
  uint32_t a[8];
  uint64_t p[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    f[i] = 0.5f + 0.001f * (threadIdx.x + i);
    a[i] = 0x3f000000u + threadIdx.x * 8 + i;
    asm volatile("mov.b64 %0, {%1,%2};" : "=l"(p[i]) : "f"(f[i]), "f"(f[i] * 0.5f));
  }
  const float c0 = 0.999f, c1 = 1e-4f;
  uint64_t pc0, pc1;
  asm volatile("mov.b64 %0, {%1,%1};" : "=l"(pc0) : "f"(c0));
  asm volatile("mov.b64 %0, {%1,%1};" : "=l"(pc1) : "f"(c1));
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < OPS_PER_ITER; ++u) {
      const int i = u & 7;
      if (MODE == kFfma) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(c0), "f"(c1));
      if (MODE == kFfma2) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pc0), "l"(pc1));
      if (MODE == kFadd2) asm volatile("add.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(pc1));
      if (MODE == kFfmaImm) asm volatile("fma.rn.f32 %0, %0, 0f3F7FBE77, 0f38D1B717;" : "+f"(f[i]));
      if (MODE == kImadShl) asm volatile("mad.lo.s32 %0, %0, 8388608, %1;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]));
      if (MODE == kFmnmx) asm volatile("max.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(c1));
      if (MODE == kF2fp) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(a[i]) : "f"(f[i]), "f"(__uint_as_float(a[i])));
      if (MODE == kMufu) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
      if (MODE == kLop3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]), "r"(a[(i + 2) & 7]));
      if (MODE == kFfmaSat) asm volatile("fma.rn.sat.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(c0), "f"(c1));
      if (MODE == kIadd) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]));
      if (MODE == kMufu_Ffma2_1to1 || MODE == kMufu_Ffma2_1to2 || MODE == kMufu_Ffma2_1to4 || MODE == kMufu_Ffma2_F2fp) {
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pc0), "l"(pc1));
        if (MODE != kMufu_Ffma2_1to1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[(i + 4) & 7]) : "l"(pc0), "l"(pc1));
        if (MODE == kMufu_Ffma2_1to4) {
          asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[(i + 2) & 7]) : "l"(pc0), "l"(pc1));
          asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[(i + 6) & 7]) : "l"(pc0), "l"(pc1));
        }
        if (MODE == kMufu_Ffma2_F2fp) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(a[i]) : "f"(c0), "f"(__uint_as_float(a[i])));
      }
      if (MODE == kMufu_F2fp) {
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(a[i]) : "f"(c0), "f"(__uint_as_float(a[i])));
      }
      if (MODE == kFfma2_F2fp) {
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pc0), "l"(pc1));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(a[i]) : "f"(c0), "f"(__uint_as_float(a[i])));
      }
      if (MODE == kFfma2_Fmnmx) {
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pc0), "l"(pc1));
        asm volatile("max.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(c1));
      }
      if (MODE == kFfma2_Imad) {
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pc0), "l"(pc1));
        asm volatile("mad.lo.s32 %0, %0, 8388608, %1;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]));
      }
    }
  }
  const long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float lo, hi;
    asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i]));
    s ^= a[i] ^ __float_as_uint(f[i]) ^ __float_as_uint(lo) ^ __float_as_uint(hi);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(int threads = 512) {
  uint32_t* d;
  long long* c;
  cudaMalloc(&d, 148 * 512 * 4);
  cudaMalloc(&c, 148 * 8);
  const int iters = 2048;
  k<MODE><<<148, threads>>>(d, c, 16);
  k<MODE><<<148, threads>>>(d, c, iters);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += (double)h[i];
  avg /= 148;
  // 16 warps per SM = 4 per sub-partition; units per warp = iters * OPS_PER_ITER
  const double units_per_smsp = (threads / 128.0) * iters * OPS_PER_ITER;
  printf("[%d warps/SMSP] %-26s %7.3f clk per unit per SMSP  (%d instr/unit -> %.3f clk per warp-instruction)\n",
         threads / 128, kNames[MODE], avg / units_per_smsp,
         kInstrPerUnit[MODE], avg / units_per_smsp / kInstrPerUnit[MODE]);
  cudaFree(d);
  cudaFree(c);
}

template <int M>
struct Runner {
  static void go() {
    run<M>();
    Runner<M + 1>::go();
  }
};
template <>
struct Runner<kNumModes> {
  static void go() {}
};

int main() {
  Runner<0>::go();
  for (int th = 128; th <= 512; th *= 2) {
    run<kMufu>(th);
    run<kMufu_Ffma2_F2fp>(th);
    run<kFfma2>(th);
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("CUDA error: %s\n", cudaGetErrorString(e));
    return 1;
  }
  return 0;
}
