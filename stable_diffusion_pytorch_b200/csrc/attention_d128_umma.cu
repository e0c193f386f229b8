// K4e: dense attention with head dim 128 on tcgen05 / TMEM / TMA -- the AutoencoderKL bottleneck (one head of
// d = mid channels over (H/8)^2 tokens; models/autoencoder.py:181-183 -> modules/transformer.py:85-114), the one attention
// shape of this code base that is bound by the tensor pipe and not by the exponentials: 4 n^2 d flops against n^2 exps.
//
// The pipeline of attention_umma_wg1.cu with wider operands: per CTA 128 query rows of one (batch, head), 64-key tiles.
//   S = Q K^T : Q as two [128 x 64] SWIZZLE_128B halves, K as two [64 x 64] halves, eight K = 16 MMAs into TMEM [0, 64)
//   P         : one query row per softmax thread, running maximum with the lazy (2^8) rescale, P rounded to bf16 into
//               TMEM [64, 96); the row sum is kept in registers (no spare V column to carry it through the MMA)
//   O += P V  : V exactly as its TMA box lands ([64 keys] x [64 dims] halves, MN-major B operand), two N = 64 MMAs per
//               K = 16 slice into the two halves of O at TMEM [128, 256)
// 256 TMEM columns and 97 KB of shared memory per CTA: two CTAs per SM.  Two of every eight exponentials run as the
// FMA-pipe polynomial: MUFU (128 x 64 / 16 = 512 cycles per tile) would otherwise tie with the tensor pipe (512).
#include "common.cuh"

namespace sdb {

constexpr int kWdThreads = 256;   // warp 0 TMA, warp 1 MMA issue, warp 2 TMEM alloc, warp 3 idle; warps 4-7: rows
constexpr int kWdBM = 128;
constexpr int kWdBN = 64;
constexpr int kWdStages = 2;
constexpr int kWdQBytes = 2 * kWdBM * 128;       // two 64-column halves, 128-byte rows: 32 KB
constexpr int kWdKBytes = 2 * kWdBN * 128;       // 16 KB
constexpr int kWdStageBytes = 2 * kWdKBytes;     // K | V
constexpr int kWdSmem = kWdQBytes + kWdStages * kWdStageBytes + 1024 + 256;
constexpr int kWdRowRegs = 200;
constexpr int kWdPoly = 2;

struct WdParams {
  CUtensorMap tmQ, tmK, tmV;      // boxes [64 columns][128 | 64 | 64 rows][1]
  __nv_bfloat16* out;
  long long ldo;
  int n, m, heads;
  float scale_log2;
};

// MN-major B operand, one SWIZZLE_128B atom wide: [keys][64] bf16 in 128-byte rows, 8-key groups 1024 B apart
__device__ __forceinline__ uint64_t umma_smem_desc_mn128(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3ffff) >> 4);
  d |= 1ull << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;  // SWIZZLE_128B
  return d;
}

__global__ void __launch_bounds__(kWdThreads, 2) attention_d128_umma_kernel(const __grid_constant__ WdParams p) {
  extern __shared__ __align__(1024) uint8_t wd_smem_raw[];
  const uint32_t base = (smem_u32(wd_smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base;
  const uint32_t sKV = base + kWdQBytes;
  const uint32_t bars = sKV + kWdStages * kWdStageBytes;
  const uint32_t bar_q = bars;
  auto bar_kv_full = [&](int s) { return bars + 8u * (1 + s); };
  auto bar_kv_empty = [&](int s) { return bars + 8u * (1 + kWdStages + s); };
  const uint32_t bar_s_full = bars + 8u * (1 + 2 * kWdStages);
  const uint32_t bar_s_free = bar_s_full + 8u;
  const uint32_t bar_p_full = bar_s_full + 16u;
  const uint32_t bar_o_full = bar_s_full + 24u;
  const uint32_t tmem_slot = bar_s_full + 32u;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(wd_smem_raw + (tmem_slot - smem_u32(wd_smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kWdBM;
  const int T = (p.m + kWdBN - 1) / kWdBN;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmQ);
    prefetch_tmap(&p.tmK);
    prefetch_tmap(&p.tmV);
    mbar_init(bar_q, 1);
    for (int s = 0; s < kWdStages; ++s) {
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
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;      // S [0, 64), P [64, 96), O [128, 256)

  if (warp < 4) {
    setmaxnreg_dec<24>();
    if (warp == 0 && lane == 0) {
      // ===================== TMA producer =====================
      mbar_arrive_expect_tx(bar_q, kWdQBytes);
      tma_load_3d(sQ, &p.tmQ, bar_q, h * 128, q0, b);
      tma_load_3d(sQ + kWdBM * 128, &p.tmQ, bar_q, h * 128 + 64, q0, b);
      bool ok = true;
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kWdStages;
        ok = mbar_wait(bar_kv_empty(s), (((uint32_t)(j / kWdStages)) & 1u) ^ 1u);
        if (!ok) break;
        const uint32_t sk = sKV + (uint32_t)s * kWdStageBytes;
        mbar_arrive_expect_tx(bar_kv_full(s), kWdStageBytes);
        tma_load_3d(sk, &p.tmK, bar_kv_full(s), h * 128, j * kWdBN, b);
        tma_load_3d(sk + kWdBN * 128, &p.tmK, bar_kv_full(s), h * 128 + 64, j * kWdBN, b);
        tma_load_3d(sk + kWdKBytes, &p.tmV, bar_kv_full(s), h * 128, j * kWdBN, b);
        tma_load_3d(sk + kWdKBytes + kWdBN * 128, &p.tmV, bar_kv_full(s), h * 128 + 64, j * kWdBN, b);
      }
    } else if (warp == 1) {
      // ===================== MMA issuer =====================
      const uint32_t idesc_s = umma_idesc_bf16(kWdBM, kWdBN);
      const uint32_t idesc_o = umma_idesc_bf16(kWdBM, 64) | (1u << 16);   // B (= a V half) is MN-major
      const uint64_t qd0 = umma_smem_desc(sQ, 128), qd1 = umma_smem_desc(sQ + kWdBM * 128, 128);
      const uint32_t tcol = tmem_base;
      auto issue_s = [&](int stage) {
        const uint32_t sk = sKV + (uint32_t)stage * kWdStageBytes;
        const uint64_t kd0 = umma_smem_desc(sk, 128), kd1 = umma_smem_desc(sk + kWdBN * 128, 128);
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) umma_f16(tcol, qd0 + 2 * ks, kd0 + 2 * ks, idesc_s, ks ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) umma_f16(tcol, qd1 + 2 * ks, kd1 + 2 * ks, idesc_s, 1u);
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
        const int s = j % kWdStages, sn = (j + 1) % kWdStages;
        const uint32_t par = (uint32_t)j & 1u;
        if (j + 1 < T) {
          ok = mbar_wait(bar_kv_full(sn), ((uint32_t)((j + 1) / kWdStages)) & 1u) && mbar_wait(bar_s_free, par);
          if (!ok) break;
          tc_fence_after();
          issue_s(sn);
        }
        ok = mbar_wait(bar_p_full, par);
        if (!ok) break;
        tc_fence_after();
        const uint32_t sv = sKV + (uint32_t)s * kWdStageBytes + kWdKBytes;
        const uint64_t vd0 = umma_smem_desc_mn128(sv), vd1 = umma_smem_desc_mn128(sv + kWdBN * 128);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kWdBN / 16; ++kk) {
            // 16 keys further along K: 16 rows of 128 B
            umma_f16_ts(tcol + 128u, tcol + 64u + (uint32_t)kk * 8u, vd0 + (uint64_t)(kk * 128), idesc_o, (j | kk) ? 1u : 0u);
            umma_f16_ts(tcol + 192u, tcol + 64u + (uint32_t)kk * 8u, vd1 + (uint64_t)(kk * 128), idesc_o, (j | kk) ? 1u : 0u);
          }
          umma_commit(bar_o_full);
          umma_commit(bar_kv_empty(s));
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== softmax warpgroup: one query row per thread =====================
    setmaxnreg_inc<kWdRowRegs>();
    const int quarter = warp & 3;
    const uint32_t tw = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const float c = p.scale_log2;
    constexpr float kLazyLog2 = 8.0f;   // the reference maximum may lag by 2^8: P stays <= 256, far inside bf16 / fp32 range
    float m_run = -INFINITY, l_run = 0.f;
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
      if (lane == 0) mbar_arrive(bar_s_free);
      if (j == T - 1 && (p.m & (kWdBN - 1)) != 0) {     // ragged key tail
        const int valid = p.m - j * kWdBN;
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= valid) su[i] = 0xff800000u;
      }
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 64; i += 8) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          mx[k] = max3(mx[k], __uint_as_float(su[i + 2 * k]), __uint_as_float(su[i + 2 * k + 1]));
      }
      const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
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
        l_run *= alpha;
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          uint32_t o[32];
          tmem_ld_32x32(tw + 128 + 32 * ch, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st_32x16(tw + 128 + 32 * ch, *reinterpret_cast<uint32_t(*)[16]>(&o[0]));
          tmem_st_32x16(tw + 144 + 32 * ch, *reinterpret_cast<uint32_t(*)[16]>(&o[16]));
        }
      }
      const float negm = -m_run * c;
      float ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t pk[16];
#pragma unroll
        for (int g = 0; g < 2; ++g) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int i = ch * 32 + g * 16 + 2 * k;
            float x0, x1;
            ffma2(x0, x1, __uint_as_float(su[i]), __uint_as_float(su[i + 1]), c, c, negm, negm);
            if ((k * kWdPoly) % 8 < kWdPoly) {
              ex2_poly2(x0, x1);
            } else {
              x0 = ex2_mufu(x0);
              x1 = ex2_mufu(x1);
            }
            fadd2(ls[(k & 1) * 2], ls[(k & 1) * 2 + 1], ls[(k & 1) * 2], ls[(k & 1) * 2 + 1], x0, x1);
            pk[g * 8 + k] = pack_bf16x2(x0, x1);          // round to nearest: the fp32 row sum stays unbiased
          }
        }
        tmem_st_32x16(tw + 64 + ch * 16, pk);
      }
      l_run += (ls[0] + ls[1]) + (ls[2] + ls[3]);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full);
    }
    if (ok) ok = mbar_wait(bar_o_full, ((uint32_t)(T - 1)) & 1u);
    tc_fence_after();
    const int row = q0 + quarter * 32 + lane;
    const float inv = 1.0f / l_run;
    __nv_bfloat16* og = p.out + ((long long)b * p.n + row) * p.ldo + h * 128;
#pragma unroll 1
    for (int ch = 0; ch < 4; ++ch) {
      uint32_t o[32];
      tmem_ld_32x32(tw + 128 + 32 * ch, o);
      tmem_ld_wait();
      if (row < p.n && ok) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 v = make_uint4(pack_bf16x2(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv),
                               pack_bf16x2(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv),
                               pack_bf16x2(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv),
                               pack_bf16x2(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv));
          *reinterpret_cast<uint4*>(og + ch * 32 + i) = v;
        }
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

// q / k / v / out: bf16 [B, rows, >= heads * 128] with row strides ld* (multiples of 8), 16-byte aligned.
// Returns SDB_EUNSUPPORTED when the operands do not meet the TMA constraints (the caller keeps the mma.sync kernel).
int sdb_attention_d128_dispatch(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                void* out, int64_t ldo, int B, int heads, int n, int m, float scale, cudaStream_t stream) {
  if (ldq % 8 || ldk % 8 || ldv % 8 || ldo % 8 || ((uintptr_t)q & 15) || ((uintptr_t)k & 15) || ((uintptr_t)v & 15) ||
      ((uintptr_t)out & 15) || ldq < heads * 128 || ldk < heads * 128 || ldv < heads * 128)
    return SDB_EUNSUPPORTED;
  WdParams p;
  memset(&p, 0, sizeof(p));
  const uint32_t qbox[3] = {64, 128, 1}, kvbox[3] = {64, 64, 1};
  uint64_t dims[3] = {(uint64_t)heads * 128, (uint64_t)n, (uint64_t)B};
  uint64_t st[3] = {0, (uint64_t)ldq * 2, (uint64_t)n * ldq * 2};
  int rc = sdb_make_tensor_map(&p.tmQ, q, 2, 3, dims, st, qbox, 128);
  if (rc) return rc;
  dims[1] = (uint64_t)m;
  st[1] = (uint64_t)ldk * 2; st[2] = (uint64_t)m * ldk * 2;
  rc = sdb_make_tensor_map(&p.tmK, k, 2, 3, dims, st, kvbox, 128);
  if (rc) return rc;
  st[1] = (uint64_t)ldv * 2; st[2] = (uint64_t)m * ldv * 2;
  rc = sdb_make_tensor_map(&p.tmV, v, 2, 3, dims, st, kvbox, 128);
  if (rc) return rc;
  p.out = (__nv_bfloat16*)out;
  p.ldo = ldo;
  p.n = n; p.m = m; p.heads = heads;
  p.scale_log2 = scale * 1.4426950408889634f;
  static bool attr_set[64] = {false};
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    SDB_CUDA_OK(cudaFuncSetAttribute(attention_d128_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWdSmem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  dim3 grid((n + kWdBM - 1) / kWdBM, heads, B);
  attention_d128_umma_kernel<<<grid, kWdThreads, kWdSmem, stream>>>(p);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

}  // namespace sdb
