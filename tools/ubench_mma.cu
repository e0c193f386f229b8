// How fast can ONE thread issue tcgen05.mma (M=128, K=16) for small N?  Decides the attention tiling.
#include "../stable_diffusion_pytorch_b200/csrc/common.cuh"
void sdb_set_error(const char*, ...) {}
void sdb_count_launch() {}
using namespace sdb;

__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
               ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

template <int N, int TS, int CHUNK>
__global__ void __launch_bounds__(128, 1) k(long long* out, int iters) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t ad = umma_smem_desc(base, 64), bd = umma_smem_desc(base + 16384, 64);
    long long t0 = clock64();
    uint32_t ph = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int c = 0; c < CHUNK; ++c) {
        if (TS) umma_ts(tm + 256, tm + (c & 7) * 8, bd + 2 * (c & 1), idesc, 1u);
        else umma_f16(tm + 256, ad + 2 * (c & 1), bd + 2 * (c & 1), idesc, 1u);
      }
      umma_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), ph); ph ^= 1;
    }
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

template <int N, int TS, int CHUNK>
void run(const char* name) {
  long long* out; cudaMalloc(&out, 8);
  cudaFuncSetAttribute(k<N, TS, CHUNK>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const int iters = 2000;
  k<N, TS, CHUNK><<<148, 128, 100 * 1024>>>(out, iters);
  cudaDeviceSynchronize();
  k<N, TS, CHUNK><<<148, 128, 100 * 1024>>>(out, iters);
  long long h = 0; cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
  printf("%-34s %8.1f clk per commit-group of %d MMAs -> %7.1f clk/MMA  (%s)\n", name, (double)h / iters, CHUNK,
         (double)h / iters / CHUNK, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}

// G groups of (4 x TS-MMA N=32 + commit) issued back to back, one wait at the end; optionally 12 other
// warps hammer TMEM with tcgen05.ld / tcgen05.st meanwhile (as the softmax warps do).
template <int LOAD>
__global__ void __launch_bounds__(512, 1) k2(long long* out, int iters) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[8];
  __shared__ uint32_t slot;
  __shared__ volatile int stop;
  const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bar[i]), 1); fence_barrier_init(); stop = 0; }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 15) {
    const uint32_t idesc = umma_idesc_bf16(128, 32);
    const uint64_t bd = umma_smem_desc(base + 16384, 64);
    long long t0 = clock64();
    uint32_t ph = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        if (elect_one()) {
#pragma unroll
          for (int c = 0; c < 4; ++c) umma_ts(tm + 448, tm + 256 + c * 8, bd + 2 * (c & 1), idesc, 1u);
          umma_commit(smem_u32(&bar[g]));
        }
        __syncwarp();
      }
      mbar_wait(smem_u32(&bar[7]), ph); ph ^= 1;
    }
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 480) out[0] = t1 - t0;
    stop = 1;
  } else if (LOAD && warp < 12) {
    uint32_t v[32];
    const uint32_t ta = tm + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
    uint32_t acc = 0;
    while (!stop) {
      tmem_ld_32x32(ta, v); tmem_ld_wait();
      for (int i = 0; i < 32; ++i) acc += v[i];
      if (LOAD > 1) {
        uint32_t w16[16];
        for (int i = 0; i < 16; ++i) w16[i] = v[i] + acc;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
          ::"r"(ta + 32), "r"(w16[0]),"r"(w16[1]),"r"(w16[2]),"r"(w16[3]),"r"(w16[4]),"r"(w16[5]),"r"(w16[6]),"r"(w16[7]),
          "r"(w16[8]),"r"(w16[9]),"r"(w16[10]),"r"(w16[11]),"r"(w16[12]),"r"(w16[13]),"r"(w16[14]),"r"(w16[15]) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      }
    }
    if (acc == 0x12345) out[1] = acc;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}
template <int LOAD>
void run2(const char* name) {
  long long* out; cudaMalloc(&out, 16);
  cudaFuncSetAttribute(k2<LOAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const int iters = 500;
  k2<LOAD><<<148, 512, 100 * 1024>>>(out, iters);
  cudaDeviceSynchronize();
  k2<LOAD><<<148, 512, 100 * 1024>>>(out, iters);
  long long h = 0; cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
  printf("%-44s %8.1f clk per 8 x (4 MMA + commit) -> %6.1f clk/MMA (%s)\n", name, (double)h / iters, (double)h / iters / 32,
         cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}

int main() {
  run2<0>("TS N=32 8x(4 MMA+commit), TMEM idle");
  run2<1>("TS N=32 8x(4 MMA+commit), 12 warps LDTM");
  run2<2>("TS N=32 8x(4 MMA+commit), 12 warps LDTM+STTM");
  run<32, 0, 1>("SS N=32  1 MMA then wait");
  run<32, 0, 4>("SS N=32  4 MMAs then wait");
  run<32, 0, 32>("SS N=32 32 MMAs then wait");
  run<32, 1, 4>("TS N=32  4 MMAs then wait");
  run<32, 1, 32>("TS N=32 32 MMAs then wait");
  run<64, 0, 2>("SS N=64  2 MMAs then wait");
  run<64, 0, 32>("SS N=64 32 MMAs then wait");
  run<128, 0, 2>("SS N=128 2 MMAs then wait");
  run<128, 0, 32>("SS N=128 32 MMAs then wait");
  run<256, 0, 32>("SS N=256 32 MMAs then wait");
  run<128, 1, 32>("TS N=128 32 MMAs then wait");
  return 0;
}
