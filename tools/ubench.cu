// Micro-benchmarks that decide the attention softmax design on B200 (sm_100a):
// per-SM throughput of MUFU ex2 (f32 / f16x2 / bf16x2), packed fp32 FMA, a packed poly-3 exp2.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdint.h>

#define ITERS 4096
template <int MODE>
__global__ void k(float* out, float seed) {
  float a[8];
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 1e-3f + i;
  uint32_t h[8];
  for (int i = 0; i < 8; ++i) h[i] = 0x38003800u + threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 3) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i]));
    }
    if (MODE == 4) {
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        asm volatile("{.reg .b64 t; mov.b64 t, {%0, %1}; fma.rn.f32x2 t, t, t, t; mov.b64 {%0, %1}, t;}" : "+f"(a[i]), "+f"(a[i + 1]));
      }
    }
    if (MODE == 5) {
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f16x2 %0, %0, %0, %0;" : "+r"(h[i]));
    }
    if (MODE == 6) {  // cvt pack f32x2 -> f16x2
#pragma unroll
      for (int i = 0; i < 8; i += 2) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(a[i] ), "f"(a[i+1]+__uint_as_float(h[i])));
    }
    if (MODE == 7) {  // 3-input max
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(a[(i + 1) & 7]), "f"(a[(i + 2) & 7]));
    }
  }
  float s = 0;
  for (int i = 0; i < 8; ++i) s += a[i] + __uint_as_float(h[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, double elems_per_instr, int instr_per_iter) {
  float* out;
  cudaMalloc(&out, 148 * 8 * 1024 * sizeof(float));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int blocks = 148 * 2, threads = 1024;
  k<MODE><<<blocks, threads>>>(out, 0.5f);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(out, 0.5f);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double instr = (double)blocks * threads * ITERS * instr_per_iter;
  double per_sm_clk = instr / (ms * 1e-3) / 148 / 1.965e9;
  printf("%-28s %8.3f ms  %7.2f lane-instr/clk/SM  %7.2f elems/clk/SM  err=%s\n", name, ms, per_sm_clk,
         per_sm_clk * elems_per_instr, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}

int main() {
  run<0>("ex2.approx.ftz.f32", 1, 8);
  run<1>("ex2.approx.f16x2", 2, 8);
  run<2>("ex2.approx.ftz.bf16x2", 2, 8);
  run<3>("fma.rn.f32", 1, 8);
  run<4>("fma.rn.f32x2", 2, 4);
  run<5>("fma.rn.f16x2", 2, 8);
  run<6>("cvt.rn.f16x2.f32", 2, 4);
  run<7>("max.f32 3-input", 2, 8);
  return 0;
}
