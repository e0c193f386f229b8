// Can a tcgen05.mma A operand start at an arbitrary 128-byte ROW of a SWIZZLE_128B tile?  (The conv kernel's "halo"
// loader wants the nine shifted views of one (rows+2) x (cols+2) pixel block instead of nine TMA loads.)
// Shared memory holds R rows x 64 bf16 written the way TMA writes a SWIZZLE_128B box into a 1024-byte aligned
// buffer (16-byte chunk index ^= row & 7); A(r, k) = code(r, k).  B = the 64 x 64 identity, so D(i, n) = A(start + i, n).
// For every start row the kernel prints how many of the 128 x 64 outputs match, with the descriptor's base-offset
// field (bits 49..51) left at 0 and set to (start_address >> 7) & 7.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/ubench_halo tools/ubench_halo.cu && /tmp/ubench_halo
#include "../stable_diffusion_pytorch_b200/csrc/common.cuh"
void sdb_set_error(const char*, ...) {}
void sdb_count_launch() {}
void sdb_register_timeout_flag(SdbFlagReader) {}
using namespace sdb;

constexpr int kRows = 400;

__device__ __forceinline__ float code(int r, int k) { return (float)((r * 7 + k * 3) % 251) - 125.0f; }

__global__ void __launch_bounds__(128, 1) k(int start_row, int use_base_offset, int* result) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  uint8_t* gen = smem + (base - smem_u32(smem));
  __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(gen);
  __nv_bfloat16* Bm = reinterpret_cast<__nv_bfloat16*>(gen + kRows * 128);
  for (int i = threadIdx.x; i < kRows * 64; i += 128) {
    const int r = i >> 6, kk = i & 63;
    const int chunk = (kk >> 3) ^ (r & 7);
    A[r * 64 + chunk * 8 + (kk & 7)] = __float2bfloat16(code(r, kk));
  }
  for (int i = threadIdx.x; i < 64 * 64; i += 128) {
    const int n = i >> 6, kk = i & 63;
    const int chunk = (kk >> 3) ^ (n & 7);
    Bm[n * 64 + chunk * 8 + (kk & 7)] = __float2bfloat16(n == kk ? 1.0f : 0.0f);
  }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 64); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, 64);
    const uint32_t a_addr = base + (uint32_t)start_row * 128u;
    uint64_t ad = umma_smem_desc(a_addr, 128);
    if (use_base_offset) ad |= (uint64_t)((a_addr >> 7) & 7u) << 49;
    const uint64_t bd = umma_smem_desc(base + kRows * 128, 128);
    for (int ks = 0; ks < 4; ++ks) umma_f16(tm, ad + 2 * ks, bd + 2 * ks, idesc, ks ? 1u : 0u);
    umma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int good = 0;
  for (int c = 0; c < 64; c += 32) {
    uint32_t v[32];
    tmem_ld_32x32(tm + ((uint32_t)(warp * 32) << 16) + c, v);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j)
      good += (__uint_as_float(v[j]) == code(start_row + warp * 32 + lane, c + j)) ? 1 : 0;
  }
  atomicAdd(result, good);
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 64); }
}

int main() {
  int* res; cudaMalloc(&res, 4);
  const int smem = kRows * 128 + 64 * 128 + 2048;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int starts[] = {0, 1, 2, 3, 7, 8, 9, 129, 130, 131, 261, 262};
  for (int s : starts)
    for (int bo = 0; bo < 2; ++bo) {
      cudaMemset(res, 0, 4);
      k<<<1, 128, smem>>>(s, bo, res);
      cudaError_t e = cudaDeviceSynchronize();
      int h = 0; cudaMemcpy(&h, res, 4, cudaMemcpyDeviceToHost);
      printf("start row %3d  base_offset %s : %5d / 8192 match  (%s)\n", s, bo ? "set" : "0  ", h, cudaGetErrorString(e));
    }
  return 0;
}
