// K4a, second shape of the same pipeline: ONE softmax warpgroup per CTA, 64-key tiles, 3 or 4 CTAs resident per SM.
// The two-warpgroup kernel (attention_umma.cu) leaves an SMSP with two row warps; ncu shows neither the MUFU pipe (52 %)
// nor the issue slots (54 %) full -- each warp alternates MUFU bursts, polynomial FMA bursts and TMEM round trips and
// one partner is not enough to fill the gaps.  Here a CTA is 128 query rows (S 64 + P 32 + O 32 = 128 TMEM columns), so
// three or four CTAs share an SM and every SMSP holds three or four row warps in different phases.  K / V tiles are
// fetched per CTA (twice the L2 -> smem traffic of the paired kernel, still < 25 % of L2 throughput).
#include "common.cuh"

namespace sdb {

constexpr int kF1Threads = 256;   // warpgroup 0: warp 0 TMA, warp 1 MMA issue, warp 2 TMEM alloc; warpgroup 1: rows
constexpr int kF1BM = 128;
constexpr int kF1BN = 64;
constexpr int kF1QBytes = kF1BM * 32 * 2;       // 8 KB
constexpr int kF1KVBytes = kF1BN * 32 * 2;      // 4 KB each for K and V
constexpr int kF1StageBytes = 2 * kF1KVBytes;
__host__ __device__ constexpr int f1_stages(int ctas) { return ctas >= 4 ? 5 : 6; }
__host__ __device__ constexpr int f1_smem(int ctas) { return kF1QBytes + f1_stages(ctas) * kF1StageBytes + 1024 + 256; }

template <int OFF>
__device__ __forceinline__ void tmem_ld_32x32_at64(uint32_t taddr, uint32_t (&v)[64]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32"
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15,"
      " %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[OFF + 0]), "=r"(v[OFF + 1]), "=r"(v[OFF + 2]), "=r"(v[OFF + 3]), "=r"(v[OFF + 4]), "=r"(v[OFF + 5]),
        "=r"(v[OFF + 6]), "=r"(v[OFF + 7]), "=r"(v[OFF + 8]), "=r"(v[OFF + 9]), "=r"(v[OFF + 10]), "=r"(v[OFF + 11]),
        "=r"(v[OFF + 12]), "=r"(v[OFF + 13]), "=r"(v[OFF + 14]), "=r"(v[OFF + 15]), "=r"(v[OFF + 16]), "=r"(v[OFF + 17]),
        "=r"(v[OFF + 18]), "=r"(v[OFF + 19]), "=r"(v[OFF + 20]), "=r"(v[OFF + 21]), "=r"(v[OFF + 22]), "=r"(v[OFF + 23]),
        "=r"(v[OFF + 24]), "=r"(v[OFF + 25]), "=r"(v[OFF + 26]), "=r"(v[OFF + 27]), "=r"(v[OFF + 28]), "=r"(v[OFF + 29]),
        "=r"(v[OFF + 30]), "=r"(v[OFF + 31])
      : "r"(taddr)
      : "memory");
}

// kPoly = element pairs out of every 8 exponentiated on the FMA pipe; kCtas = CTAs resident per SM (3 or 4);
// kSpread = the polynomial pairs are spread over the group of 8 instead of leading it (1.105 -> 1.085 ms at kPoly = 2)
// kDbg (measurement aid, only 0 is instantiated in the library; tools/experiments/README.md has the numbers):
// bit 0 = skip the scale/shift FFMA2, bit 1 = skip the row maximum, bit 2 = release S only after the maximum,
// bits 3 / 4 = sleeping polls in the TMA / MMA warps.
template <int kPoly, int kCtas, bool kSpread, int kDbg = 0>
__global__ void __launch_bounds__(kF1Threads, kCtas) attention_umma_wg1_kernel(const __grid_constant__ FaParams p) {
  constexpr int kStages = f1_stages(kCtas);
  constexpr int kRowRegs = kCtas >= 4 ? 104 : 136;
  extern __shared__ __align__(1024) uint8_t f1_smem_raw[];
  const uint32_t base = (smem_u32(f1_smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base;
  const uint32_t sKV = base + kF1QBytes;                      // stages x (K 4 KB | V 4 KB)
  const uint32_t bars = sKV + kStages * kF1StageBytes;
  const uint32_t bar_q = bars;
  auto bar_kv_full = [&](int s) { return bars + 8u * (1 + s); };
  auto bar_kv_empty = [&](int s) { return bars + 8u * (1 + kStages + s); };
  const uint32_t bar_s_full = bars + 8u * (1 + 2 * kStages);
  const uint32_t bar_s_free = bar_s_full + 8u;
  const uint32_t bar_p_full = bar_s_full + 16u;
  const uint32_t bar_o_full = bar_s_full + 24u;
  const uint32_t tmem_slot = bar_s_full + 32u;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(f1_smem_raw + (tmem_slot - smem_u32(f1_smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kF1BM;
  const int T = (p.m + kF1BN - 1) / kF1BN;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmQ);
    prefetch_tmap(&p.tmK);
    prefetch_tmap(&p.tmV);
    mbar_init(bar_q, 1);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_kv_full(s), 1);
      mbar_init(bar_kv_empty(s), 1);
    }
    mbar_init(bar_s_full, 1);
    mbar_init(bar_s_free, 4);
    mbar_init(bar_p_full, 4);
    mbar_init(bar_o_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 128);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // TMEM columns: S [0,64), P [64,96), O [96,128)

  if (warp < 4) {
    setmaxnreg_dec<24>();
    if (warp == 0 && lane == 0) {
      // ===================== TMA producer =====================
      mbar_arrive_expect_tx(bar_q, kF1QBytes);
      tma_load_3d(sQ, &p.tmQ, bar_q, h * 32, q0, b);
      bool ok = true;
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kStages;
        ok = (kDbg & 8) ? mbar_wait_sleep(bar_kv_empty(s), (((uint32_t)(j / kStages)) & 1u) ^ 1u, 256)
                        : mbar_wait(bar_kv_empty(s), (((uint32_t)(j / kStages)) & 1u) ^ 1u);
        if (!ok) break;
        const uint32_t sk = sKV + (uint32_t)s * kF1StageBytes;
        mbar_arrive_expect_tx(bar_kv_full(s), kF1StageBytes);
        tma_load_3d(sk, &p.tmK, bar_kv_full(s), h * 32, j * kF1BN, b);
        tma_load_3d(sk + kF1KVBytes, &p.tmV, bar_kv_full(s), h * 32, j * kF1BN, b);
      }
    } else if (warp == 1) {
      // ===================== MMA issuer (warp-uniform loop, one elected lane issues) =====================
      const uint32_t idesc_s = umma_idesc_bf16(kF1BM, kF1BN);
      const uint32_t idesc_o = umma_idesc_bf16(kF1BM, 32) | (1u << 16);   // B (= V) is MN-major
      const uint64_t qdesc = umma_smem_desc(sQ, 64);
      const uint32_t tcol = tmem_base;
      auto issue_s = [&](int stage) {
        const uint64_t bdesc = umma_smem_desc(sKV + (uint32_t)stage * kF1StageBytes, 64);
        if (elect_one()) {
          umma_f16(tcol, qdesc, bdesc, idesc_s, 0u);
          umma_f16(tcol, qdesc + 2, bdesc + 2, idesc_s, 1u);
          umma_commit(bar_s_full);
        }
        __syncwarp();
      };
      bool ok = mbar_wait(bar_q, 0) && mbar_wait(bar_kv_full(0), 0);
      if (ok) {
        tc_fence_after();
        issue_s(0);
      }
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kStages, sn = (j + 1) % kStages;
        const uint32_t par = (uint32_t)j & 1u;
        if (j + 1 < T) {
          ok = mbar_wait(bar_kv_full(sn), ((uint32_t)((j + 1) / kStages)) & 1u) && mbar_wait(bar_s_free, par);
          if (!ok) break;
          tc_fence_after();
          issue_s(sn);
        }
        ok = (kDbg & 16) ? mbar_wait_sleep(bar_p_full, par, 32) : mbar_wait(bar_p_full, par);
        if (!ok) break;
        tc_fence_after();
        const uint64_t vdesc = umma_smem_desc_mn64(sKV + (uint32_t)s * kF1StageBytes + kF1KVBytes);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kF1BN / 16; ++kk)
            umma_f16_ts(tcol + 96u, tcol + 64u + (uint32_t)kk * 8u, vdesc + (uint64_t)(kk * 64), idesc_o,
                        (j | kk) ? 1u : 0u);
          umma_commit(bar_o_full);
          umma_commit(bar_kv_empty(s));
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== softmax warpgroup: one query row per thread =====================
    setmaxnreg_inc<kRowRegs>();
    const int quarter = warp & 3;
    const uint32_t tw = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const float c = p.scale_log2;
    constexpr float kLazyLog2 = 8.0f;   // lagging reference maximum, see attention_umma.cu
    float m_run = -INFINITY;
    bool ok = true;
    for (int j = 0; j < T; ++j) {
      const uint32_t par = (uint32_t)j & 1u;
      if (ok) ok = mbar_wait(bar_s_full, par);
      tc_fence_after();
      uint32_t su[64];
      tmem_ld_32x32_at64<0>(tw + 0, su);
      tmem_ld_32x32_at64<32>(tw + 32, su);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (!(kDbg & 4) && lane == 0) mbar_arrive(bar_s_free);
      if (j == T - 1 && (p.m & (kF1BN - 1)) != 0) {     // ragged key tail
        const int valid = p.m - j * kF1BN;
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= valid) su[i] = 0xff800000u;
      }
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (kDbg & 2) {
        mx[0] = __uint_as_float(su[0]);
      } else {
#pragma unroll
        for (int i = 0; i < 64; i += 8) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mx[k] = max3(mx[k], __uint_as_float(su[i + 2 * k]), __uint_as_float(su[i + 2 * k + 1]));
        }
      }
      const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      if (kDbg & 4) {
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_s_free);
      }
      if (j > 0) {
        if (ok) ok = mbar_wait(bar_o_full, par ^ 1u);
        tc_fence_after();
      }
      const bool grow = (m_tile - m_run) * c > kLazyLog2;
      if (j == 0) {
        m_run = m_tile;
      } else if (__any_sync(0xffffffffu, grow)) {
        const float m_new = grow ? m_tile : m_run;
        const float alpha = ex2_mufu((m_run - m_new) * c);
        m_run = m_new;
        uint32_t oa[16], ob[16];
        tmem_ld_32x16(tw + 96, oa);
        tmem_ld_32x16(tw + 112, ob);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          oa[i] = __float_as_uint(__uint_as_float(oa[i]) * alpha);
          ob[i] = __float_as_uint(__uint_as_float(ob[i]) * alpha);
        }
        tmem_st_32x16(tw + 96, oa);
        tmem_st_32x16(tw + 112, ob);
      }
      const float negm = -m_run * c;
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t pk[16];
#pragma unroll
        for (int g = 0; g < 2; ++g) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int i = ch * 32 + g * 16 + 2 * k;
            float x0, x1;
            if (kDbg & 1) {
              x0 = __uint_as_float(su[i]) ; x1 = __uint_as_float(su[i + 1]);
            } else {
              ffma2(x0, x1, __uint_as_float(su[i]), __uint_as_float(su[i + 1]), c, c, negm, negm);
            }
            if (kSpread ? ((k * kPoly) % 8 < kPoly) : (k < kPoly)) {
              ex2_poly2(x0, x1);
            } else {
              x0 = ex2_mufu(x0);
              x1 = ex2_mufu(x1);
            }
            pk[g * 8 + k] = __byte_perm(__float_as_uint(x0), __float_as_uint(x1), 0x7632);
          }
        }
        tmem_st_32x16(tw + 64 + ch * 16, pk);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full);
    }
    if (ok) ok = mbar_wait(bar_o_full, ((uint32_t)(T - 1)) & 1u);
    tc_fence_after();
    float acc[32];
    {
      uint32_t o[32];
      tmem_ld_32x32(tw + 96, o);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) acc[i] = __uint_as_float(o[i]);
    }
    const int row = q0 + quarter * 32 + lane;
    if (row < p.n && ok) {
      float den = acc[0];
#pragma unroll
      for (int i = 1; i < 24; ++i)
        if (i == p.d) den = acc[i];
      const float inv = 1.0f / den;
      if (p.lse) p.lse[((long long)b * p.heads + h) * p.n + row] = fmaf(m_run, c, log2f(den));
      __nv_bfloat16* og = p.out + ((long long)b * p.n + row) * p.ldo + h * p.d;
#pragma unroll
      for (int i = 0; i < 24; i += 4) {
        if (i < p.d) {
          uint2 v = make_uint2(pack_bf16x2(acc[i] * inv, acc[i + 1] * inv), pack_bf16x2(acc[i + 2] * inv, acc[i + 3] * inv));
          *reinterpret_cast<uint2*>(og + i) = v;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

template <int kPoly, int kCtas, bool kSpread = false, int kDbg = 0>
static int launch_attention_wg1(const FaParams& p, dim3 grid, cudaStream_t stream) {
  static bool attr_set[64] = {false};
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    SDB_CUDA_OK(cudaFuncSetAttribute(attention_umma_wg1_kernel<kPoly, kCtas, kSpread, kDbg>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     f1_smem(kCtas)));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  attention_umma_wg1_kernel<kPoly, kCtas, kSpread, kDbg><<<grid, kF1Threads, f1_smem(kCtas), stream>>>(p);
  return SDB_OK;
}

}  // namespace sdb
