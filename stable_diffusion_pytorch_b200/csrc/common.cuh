// Shared device/host helpers for libsdb200 (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/sdb200.h"

// ---------------------------------------------------------------------------------------------
// host-side error plumbing: entry points never throw / exit; they return SDB_E* and leave a
// message for sdb_last_error().
// ---------------------------------------------------------------------------------------------
void sdb_set_error(const char* fmt, ...);

#define SDB_CHECK_ARG(cond, ...)                 \
  do {                                           \
    if (!(cond)) {                               \
      sdb_set_error(__VA_ARGS__);                \
      return SDB_EINVAL;                         \
    }                                            \
  } while (0)

#define SDB_CUDA_OK(expr)                                                              \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      sdb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return SDB_ECUDA;                                                                \
    }                                                                                  \
  } while (0)

// every kernel launch goes through this: counts it (bench.py reports gpu_launches) and checks it
#define SDB_LAUNCH_OK()                       \
  do {                                        \
    sdb_count_launch();                       \
    SDB_CUDA_OK(cudaGetLastError());          \
  } while (0)
void sdb_count_launch();

static inline cudaStream_t sdb_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

int sdb_num_sms();

// host functions that cross translation units
namespace sdb {
int sdb_attention_f32_impl(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                           void* out, int64_t ldo, int B, int heads, int n, int m, int d, float scale,
                           cudaStream_t st);                                   // precise_f32.cu
int sdb_attention_ctx_dispatch(const void* q, long long ldq, const void* k, long long ldk, const void* v,
                               long long ldv, void* out, long long ldo, int B, int heads, int n, int m, int d,
                               float scale, cudaStream_t st);
// tcgen05 dense attention for head dim 128 (attention_d128_umma.cu); SDB_EUNSUPPORTED = operands not TMA-addressable
int sdb_attention_d128_dispatch(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                void* out, int64_t ldo, int B, int heads, int n, int m, float scale, cudaStream_t stream);                  // attention_ctx.cu
}  // namespace sdb

// Encodes a tiled tensor map (rank <= 5) for a bf16/fp32 tensor.  dims/strides innermost first;
// strides in BYTES for dims 1..rank-1 (dim 0 is contiguous).  swizzle_bytes in {0, 32, 64, 128}.
int sdb_make_tensor_map(CUtensorMap* out, const void* base, int elem_bytes, int rank,
                        const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                        int swizzle_bytes);

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__

// set by any kernel whose bounded mbarrier wait expired (bad descriptor / pipeline bug); host
// reads it through sdb_debug_timeout_flag().  Never hit in a correct run.
// The library is a handful of translation units compiled in parallel without relocatable device code,
// so every unit carries its own copy of the flag and registers a reader with runtime.cu
// (sdb_register_timeout_flag) at load time.
static __device__ int g_sdb_timeout_flag = 0;
typedef int (*SdbFlagReader)(void);   // returns the unit's flag and clears it; -1 on a CUDA error
void sdb_register_timeout_flag(SdbFlagReader fn);
static int sdb_tu_read_timeout_flag() {
  int v = 0, zero = 0;
  if (cudaMemcpyFromSymbol(&v, g_sdb_timeout_flag, sizeof(int)) != cudaSuccess) return -1;
  if (v) cudaMemcpyToSymbol(g_sdb_timeout_flag, &zero, sizeof(int));
  return v;
}
namespace {
struct SdbFlagRegistrar {
  SdbFlagRegistrar() { sdb_register_timeout_flag(&sdb_tu_read_timeout_flag); }
};
static SdbFlagRegistrar g_sdb_flag_registrar;
}  // namespace

namespace sdb {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "elect.sync _|P1, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}

// ---- mbarrier -------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// non-blocking probe (try_wait may suspend the thread for a hardware time slice; a scheduler that
// polls several barriers round-robin must not)
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// try_wait with a suspend-time hint: the hardware parks the thread until the phase completes (or the hint runs out)
// instead of returning after its short default slice.  A polling warp then issues a handful of instructions per wait
// instead of thousands -- ncu on the attention kernel showed a third of all issued instructions were TMA / MMA warps
// spinning on the un-hinted form, competing with the softmax warps of the same SM sub-partitions.
#ifndef SDB_MBAR_SUSPEND_NS
#define SDB_MBAR_SUSPEND_NS 0x989680u   // 10 ms upper bound per try; completion wakes the thread at once
#endif
__device__ __forceinline__ bool mbar_try_wait_hint(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity), "r"(SDB_MBAR_SUSPEND_NS)
      : "memory");
  return ok != 0;
}
// Bounded wait: gives up after ~2 s, raises the global flag and carries on so that a pipeline
// bug can never hang the GPU box.  Returns false on timeout.
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return true;
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait_hint(bar, parity)) {
    if ((++spins & 0xf) == 0) {
      uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > 2000000000ull) {
        atomicExch(&g_sdb_timeout_flag, 1);
        return false;
      }
    }
  }
  return true;
}

// mbar_wait without the suspend hint: the un-hinted try_wait returns after its short default slice, so a waiter on a
// critical path polls at a higher rate
__device__ __forceinline__ bool mbar_wait_spin(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return true;
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3ff) == 0) {
      uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > 2000000000ull) {
        atomicExch(&g_sdb_timeout_flag, 1);
        return false;
      }
    }
  }
  return true;
}
// mbar_wait for warps with slack (producers that run several stages ahead): polls with a sleep in between so the
// polling does not take issue slots from the compute warps on the same SM sub-partition
__device__ __forceinline__ bool mbar_wait_sleep(uint32_t bar, uint32_t parity, uint32_t ns) {
  if (mbar_try_wait(bar, parity)) return true;
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    __nanosleep(ns);
    if ((++spins & 0x3ff) == 0) {
      uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > 2000000000ull) {
        atomicExch(&g_sdb_timeout_flag, 1);
        return false;
      }
    }
  }
  return true;
}

// ---- TMA ------------------------------------------------------------------------------------
__device__ __forceinline__ void prefetch_tmap(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, uint32_t bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* tmap, uint32_t bar, int c0,
                                            int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1), "r"(c2),
      "r"(c3)
      : "memory");
}

// ---- tcgen05 / TMEM -------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16 and fp16 inputs.
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   bar)
               : "memory");
}
// ---- CTA pair (cta_group::2): two CTAs of a cluster on the SMs of one TPC issue ONE MMA of M = 256 ------------------
// Each CTA stages its own 128 rows of A and HALF of the B tile; the leader (cluster rank 0) issues the MMA, which reads
// both CTAs' shared memory at the same offsets and writes each CTA's 128 accumulator rows into that CTA's TMEM.
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_2cta() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_f16_2cta(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// commit of the pair's MMAs, arriving on the barrier at offset `bar` in BOTH CTAs
__device__ __forceinline__ void umma_commit_2cta(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ uint32_t sdb_cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void sdb_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_addr` (a shared::cta address of this CTA's layout) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t sdb_mapa(uint32_t local_addr, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  return remote;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
// TMA tile loads of a CTA pair: the bytes land in THIS CTA's shared memory, completion is signalled on `cluster_bar`,
// a shared::cluster barrier address (the leader's)
__device__ __forceinline__ void tma_load_2d_2cta(uint32_t dst, const void* tmap, uint32_t cluster_bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(cluster_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d_2cta(uint32_t dst, const void* tmap, uint32_t cluster_bar, int c0, int c1,
                                                 int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(cluster_bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_2cta(uint32_t dst, const void* tmap, uint32_t cluster_bar, int c0, int c1,
                                                 int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(cluster_bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread t = lane base + t)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32"
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15,"
      " %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]),
        "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]),
        "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32"
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major operand tile in shared memory, rows of `swizzle_bytes` bytes (one swizzle atom wide),
// 8-row groups packed densely: SBO = 8 * swizzle_bytes.  (cute/arch/mma_sm100_desc.hpp layout.)
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t saddr, uint32_t swizzle_bytes) {
  const uint64_t layout = swizzle_bytes == 128 ? 2ull : (swizzle_bytes == 64 ? 4ull : 6ull);
  const uint64_t sbo = (8ull * swizzle_bytes) >> 4;
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3ffff) >> 4);  // start address, bits [0,14)
  d |= 1ull << 16;                                     // LBO (unused for swizzled K-major)
  d |= sbo << 32;                                      // SBO, bits [32,46)
  d |= 1ull << 46;                                     // descriptor version (Blackwell)
  d |= layout << 61;                                   // swizzle mode
  return d;
}

// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, both operands K-major, M x N tile.
__host__ __device__ __forceinline__ uint32_t umma_idesc_bf16(uint32_t M, uint32_t N) {
  uint32_t d = 0;
  d |= 1u << 4;          // D format = F32
  d |= 1u << 7;          // A format = BF16
  d |= 1u << 10;         // B format = BF16
  d |= (N >> 3) << 17;   // N >> 3
  d |= (M >> 4) << 24;   // M >> 4
  return d;
}

// ---- misc math ------------------------------------------------------------------------------
// x * sigmoid(x); ex2.approx + fast divide (a few ulp, far inside the bf16 rounding of every consumer)
__device__ __forceinline__ float silu_f(float x) { return __fdividef(x, 1.0f + __expf(-x)); }
// Channel sums handed from a conv / GEMM epilogue to the GroupNorm that follows (SdbConvArgs.chan_sums): 64-bit fixed
// point so that the atomic accumulation is exact and order-independent.  Sum: signed, 2^-30 steps (|sum| < 8.6e9);
// sum of squares: unsigned, 2^-24 steps (< 1.1e12 per image and channel: an rms of 2000 over a 512 x 512 image).
// A partial sum is rounded once to that grid: <= 6e-8 / 64 elements, far inside GroupNorm's eps.
constexpr float kChanSumScale = 1073741824.0f;   // 2^30
constexpr float kChanSqScale = 16777216.0f;      // 2^24

__device__ __forceinline__ float gelu_erf_f(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

}  // namespace sdb
#endif  // __CUDACC__
