// Attention backward, dQ on the tensor cores (tcgen05 / TMEM / TMA), head-padded operands, d <= 20 (row T1).
//   S = Q K^T,  P = exp2(S c - lse2),  dP = dO V^T,  dS = P o (dP - delta),  dQ = scale * dS K
// The pipeline of the forward kernel (attention_umma.cu) with the roles re-cast: per 64-key tile the MMA warp issues
// S = Q.K^T and dP = dO.V^T into TMEM (both operands K-major as they land from TMA; V's ones column meets the zero pad
// column of dO), each softmax thread owns one query row -- its lse2 and delta are two scalars -- reads the two 64-column
// rows back, forms dS in registers, writes it to TMEM as packed bf16 and the MMA warp consumes it as the A operand of
// dQ += dS.K with the SAME K tile read as an MN-major B operand.  dQ stays in TMEM across the key loop.
// No running maximum, no rescale: the forward kernel handed over the row log-sum-exp.
// TMEM columns per warpgroup w (256 each): S [0,64), dP [64,128), dS [128,160), dQ [192,224).
#include "common.cuh"
#include "umma_helpers.cuh"

namespace sdb {

constexpr int kBdThreads = 384;   // warp 0: TMA, 1-2: MMA issue, 3: TMEM alloc; 4-7, 8-11: the two row warpgroups
constexpr int kBdStages = 6;
constexpr int kBdBM = 128;
constexpr int kBdBN = 64;         // keys per tile (TMEM budget: S + dP + dS + dQ per warpgroup)
constexpr int kBdQBytes = kBdBM * 32 * 2;       // 8 KB
constexpr int kBdKVBytes = kBdBN * 32 * 2;      // 4 KB each for K and V
constexpr int kBdStageBytes = 2 * kBdKVBytes;
constexpr int kBdSmem = 4 * kBdQBytes + kBdStages * kBdStageBytes + 1024 + 256;

// pairs of every 8 exponentiated on the FMA pipe; SDB_BWD_POLY=0|2|3 overrides the default (2) for A/B timing
static int bwd_poly_pairs() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("SDB_BWD_POLY");
    v = e ? atoi(e) : 2;
  }
  return v;
}

struct BdParams {
  CUtensorMap tmQ, tmdO, tmK, tmV;
  const float* lse;     // [B, heads, n] log2-domain log-sum-exp of the scaled scores
  const float* delta;   // [B, heads, n] rowsum(dO o O)
  __nv_bfloat16* dq;    // [B, n, >= heads*32] head-padded
  long long lddq;
  int n, m, heads;
  float scale, scale_log2;
};

// kPoly = pairs out of every 8 whose exp2 runs as a polynomial on the FMA pipe (the MUFU unit is the bound otherwise)
template <int kPoly>
__global__ void __launch_bounds__(kBdThreads, 1) attn_bwd_dq_umma_kernel(const __grid_constant__ BdParams p) {
  extern __shared__ __align__(1024) uint8_t bd_smem[];
  const uint32_t base = (smem_u32(bd_smem) + 1023u) & ~1023u;
  const uint32_t sQ = base;                                  // 2 x 8 KB
  const uint32_t sdO = base + 2 * kBdQBytes;                 // 2 x 8 KB
  const uint32_t sKV = base + 4 * kBdQBytes;                 // stages x (K 4 KB | V 4 KB)
  const uint32_t bars = sKV + kBdStages * kBdStageBytes;
  const uint32_t bar_q = bars;
  auto bar_kv_full = [&](int s) { return bars + 8u * (1 + s); };
  auto bar_kv_empty = [&](int s) { return bars + 8u * (1 + kBdStages + s); };
  auto bar_s_full = [&](int w) { return bars + 8u * (1 + 2 * kBdStages + w); };
  auto bar_s_free = [&](int w) { return bars + 8u * (3 + 2 * kBdStages + w); };
  auto bar_p_full = [&](int w) { return bars + 8u * (5 + 2 * kBdStages + w); };
  auto bar_o_full = [&](int w) { return bars + 8u * (7 + 2 * kBdStages + w); };
  const uint32_t tmem_slot = bars + 8u * (9 + 2 * kBdStages);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(bd_smem + (tmem_slot - smem_u32(bd_smem)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * (2 * kBdBM);
  const int T = (p.m + kBdBN - 1) / kBdBN;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmQ);
    prefetch_tmap(&p.tmdO);
    prefetch_tmap(&p.tmK);
    prefetch_tmap(&p.tmV);
    mbar_init(bar_q, 1);
    for (int s = 0; s < kBdStages; ++s) {
      mbar_init(bar_kv_full(s), 1);
      mbar_init(bar_kv_empty(s), 2);
    }
    for (int w = 0; w < 2; ++w) {
      mbar_init(bar_s_full(w), 1);
      mbar_init(bar_s_free(w), 4);
      mbar_init(bar_p_full(w), 4);
      mbar_init(bar_o_full(w), 1);
    }
    fence_barrier_init();
  }
  if (warp == 3) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp < 4) {
    setmaxnreg_dec<56>();
    if (warp == 0 && lane == 0) {
      // ===================== TMA producer =====================
      mbar_arrive_expect_tx(bar_q, 4 * kBdQBytes);
      tma_load_3d(sQ, &p.tmQ, bar_q, h * 32, q0, b);
      tma_load_3d(sQ + kBdQBytes, &p.tmQ, bar_q, h * 32, q0 + kBdBM, b);
      tma_load_3d(sdO, &p.tmdO, bar_q, h * 32, q0, b);
      tma_load_3d(sdO + kBdQBytes, &p.tmdO, bar_q, h * 32, q0 + kBdBM, b);
      bool ok = true;
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kBdStages;
        ok = mbar_wait(bar_kv_empty(s), (((uint32_t)(j / kBdStages)) & 1u) ^ 1u);
        if (!ok) break;
        const uint32_t sk = sKV + (uint32_t)s * kBdStageBytes;
        mbar_arrive_expect_tx(bar_kv_full(s), kBdStageBytes);
        tma_load_3d(sk, &p.tmK, bar_kv_full(s), h * 32, j * kBdBN, b);
        tma_load_3d(sk + kBdKVBytes, &p.tmV, bar_kv_full(s), h * 32, j * kBdBN, b);
      }
    } else if (warp == 1 || warp == 2) {
      // ===================== MMA issuers: one warp per row warpgroup =====================
      const int w = warp - 1;
      const uint32_t idesc_s = umma_idesc_bf16(kBdBM, kBdBN);
      const uint32_t idesc_q = umma_idesc_bf16(kBdBM, 32) | (1u << 16);   // B (= K tile) read MN-major
      const uint64_t qdesc = umma_smem_desc(sQ + (uint32_t)w * kBdQBytes, 64);
      const uint64_t dodesc = umma_smem_desc(sdO + (uint32_t)w * kBdQBytes, 64);
      const uint32_t tcol = tmem_base + (uint32_t)w * 256u;
      auto issue_s = [&](int stage) {
        const uint32_t sk = sKV + (uint32_t)stage * kBdStageBytes;
        const uint64_t kdesc = umma_smem_desc(sk, 64);
        const uint64_t vdesc = umma_smem_desc(sk + kBdKVBytes, 64);
        if (elect_one()) {
          umma_f16(tcol, qdesc, kdesc, idesc_s, 0u);                  // S = Q K^T
          umma_f16(tcol, qdesc + 2, kdesc + 2, idesc_s, 1u);
          umma_f16(tcol + 64u, dodesc, vdesc, idesc_s, 0u);           // dP = dO V^T
          umma_f16(tcol + 64u, dodesc + 2, vdesc + 2, idesc_s, 1u);
          umma_commit(bar_s_full(w));
        }
        __syncwarp();
      };
      bool ok = mbar_wait(bar_q, 0) && mbar_wait(bar_kv_full(0), 0);
      if (ok) {
        tc_fence_after();
        issue_s(0);
      }
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kBdStages, sn = (j + 1) % kBdStages;
        const uint32_t par = (uint32_t)j & 1u;
        if (j + 1 < T) {
          ok = mbar_wait(bar_kv_full(sn), ((uint32_t)((j + 1) / kBdStages)) & 1u) && mbar_wait(bar_s_free(w), par);
          if (!ok) break;
          tc_fence_after();
          issue_s(sn);
        }
        ok = mbar_wait(bar_p_full(w), par);
        if (!ok) break;
        tc_fence_after();
        const uint64_t kmn = umma_smem_desc_mn64(sKV + (uint32_t)s * kBdStageBytes);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kBdBN / 16; ++kk)
            umma_f16_ts(tcol + 192u, tcol + 128u + (uint32_t)kk * 8u, kmn + (uint64_t)(kk * 64), idesc_q,
                        (j | kk) ? 1u : 0u);   // dQ accumulates in TMEM across the key tiles
          umma_commit(bar_o_full(w));
          umma_commit(bar_kv_empty(s));
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== row warpgroups: one query row per thread =====================
    setmaxnreg_inc<224>();
    const int w = (warp - 4) >> 2, quarter = warp & 3;
    const uint32_t tw = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)w * 256u;
    const float c = p.scale_log2;
    const int row = q0 + w * kBdBM + quarter * 32 + lane;
    float neg_lse = 0.f, neg_delta = 0.f;
    if (row < p.n) {
      const long long o = ((long long)b * p.heads + h) * p.n + row;
      neg_lse = -p.lse[o];
      neg_delta = -p.delta[o];
    }
    bool ok = true;
    for (int j = 0; j < T; ++j) {
      const uint32_t par = (uint32_t)j & 1u;
      if (ok) ok = mbar_wait(bar_s_full(w), par);
      tc_fence_after();
      uint32_t su[128];                       // [0,64) = S row, [64,128) = dP row
      tmem_ld_32x32_at<0>(tw + 0, su);
      tmem_ld_32x32_at<32>(tw + 32, su);
      tmem_ld_32x32_at<64>(tw + 64, su);
      tmem_ld_32x32_at<96>(tw + 96, su);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_s_free(w));
      if (j == T - 1 && (p.m & (kBdBN - 1)) != 0) {     // ragged key tail: P = exp2(-inf) = 0
        const int valid = p.m - j * kBdBN;
#pragma unroll
        for (int i = 0; i < kBdBN; ++i)
          if (i >= valid) su[i] = 0xff800000u;
      }
      // the previous tile's dS.K must have finished reading dS before it is overwritten
      if (j > 0) {
        if (ok) ok = mbar_wait(bar_o_full(w), par ^ 1u);
        tc_fence_after();
      }
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t pk[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const int i = ch * 32 + 2 * k;
          float x0, x1, d0, d1;
          ffma2(x0, x1, __uint_as_float(su[i]), __uint_as_float(su[i + 1]), c, c, neg_lse, neg_lse);
          if ((k & 7) < kPoly) {
            ex2_poly2(x0, x1);
          } else {
            x0 = ex2_mufu(x0);
            x1 = ex2_mufu(x1);
          }
          fadd2(d0, d1, __uint_as_float(su[64 + i]), __uint_as_float(su[64 + i + 1]), neg_delta, neg_delta);
          fmul2(d0, d1, d0, d1, x0, x1);
          pk[k] = pack_bf16x2(d0, d1);
        }
        tmem_st_32x16(tw + 128 + ch * 16, pk);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full(w));
    }
    if (ok) ok = mbar_wait(bar_o_full(w), ((uint32_t)(T - 1)) & 1u);
    tc_fence_after();
    uint32_t o[32];
    tmem_ld_32x32(tw + 192, o);
    tmem_ld_wait();
    if (row < p.n && ok) {
      uint4* dst = reinterpret_cast<uint4*>(p.dq + ((long long)b * p.n + row) * p.lddq + h * 32);
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        dst[v] = make_uint4(
            pack_bf16x2(__uint_as_float(o[8 * v]) * p.scale, __uint_as_float(o[8 * v + 1]) * p.scale),
            pack_bf16x2(__uint_as_float(o[8 * v + 2]) * p.scale, __uint_as_float(o[8 * v + 3]) * p.scale),
            pack_bf16x2(__uint_as_float(o[8 * v + 4]) * p.scale, __uint_as_float(o[8 * v + 5]) * p.scale),
            pack_bf16x2(__uint_as_float(o[8 * v + 6]) * p.scale, __uint_as_float(o[8 * v + 7]) * p.scale));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace sdb

// dQ of the head-padded attention on tcgen05 (see the kernel).  q / k / v / dO_pad / dq: bf16, head h at columns
// [32h, 32h+32) (q, k, dO_pad zero beyond d; v with its ones column at d); lse / delta: fp32 [B, heads, n].
extern "C" int sdb_attention_bwd_dq_p32(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                        const void* dO_pad, int64_t lddo, const float* lse, const float* delta, void* dq,
                                        int64_t lddq, int B, int heads, int n, int m, int d, float scale, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(q && k && v && dO_pad && lse && delta && dq && B > 0 && heads > 0 && n > 0 && m > 0,
                "attention_bwd_dq_p32: bad args");
  SDB_CHECK_ARG(d > 0 && d <= 20 && d % 4 == 0, "attention_bwd_dq_p32: head dim (%d) must be a multiple of 4, <= 20", d);
  SDB_CHECK_ARG(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && lddo % 8 == 0 && lddq % 8 == 0 &&
                    ldq >= heads * 32 && ldk >= heads * 32 && ldv >= heads * 32 && lddo >= heads * 32 && lddq >= heads * 32,
                "attention_bwd_dq_p32: row strides must be multiples of 8 covering heads*32 padded columns");
  SDB_CHECK_ARG(((uintptr_t)q & 15) == 0 && ((uintptr_t)k & 15) == 0 && ((uintptr_t)v & 15) == 0 &&
                    ((uintptr_t)dO_pad & 15) == 0 && ((uintptr_t)dq & 15) == 0,
                "attention_bwd_dq_p32: operands must be 16-byte aligned");
  SDB_CHECK_ARG(heads <= 65535 && B <= 65535, "attention_bwd_dq_p32: grid too large");
  BdParams p;
  memset(&p, 0, sizeof(p));
  auto make = [&](CUtensorMap* tm, const void* base, int64_t ld, int rows, uint32_t box_rows) {
    const uint32_t box[3] = {32, box_rows, 1};
    uint64_t dims[3] = {(uint64_t)heads * 32, (uint64_t)rows, (uint64_t)B};
    uint64_t st[3] = {0, (uint64_t)ld * 2, (uint64_t)rows * ld * 2};
    return sdb_make_tensor_map(tm, base, 2, 3, dims, st, box, 64);
  };
  int rc;
  if ((rc = make(&p.tmQ, q, ldq, n, kBdBM))) return rc;
  if ((rc = make(&p.tmdO, dO_pad, lddo, n, kBdBM))) return rc;
  if ((rc = make(&p.tmK, k, ldk, m, kBdBN))) return rc;
  if ((rc = make(&p.tmV, v, ldv, m, kBdBN))) return rc;
  p.lse = lse; p.delta = delta; p.dq = (__nv_bfloat16*)dq; p.lddq = lddq;
  p.n = n; p.m = m; p.heads = heads;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
  static bool attr_set[64] = {false};
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    SDB_CUDA_OK(cudaFuncSetAttribute(attn_bwd_dq_umma_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBdSmem));
    SDB_CUDA_OK(cudaFuncSetAttribute(attn_bwd_dq_umma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBdSmem));
    SDB_CUDA_OK(cudaFuncSetAttribute(attn_bwd_dq_umma_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBdSmem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  dim3 grid((n + 2 * kBdBM - 1) / (2 * kBdBM), heads, B);
  switch (bwd_poly_pairs()) {
    case 0: attn_bwd_dq_umma_kernel<0><<<grid, kBdThreads, kBdSmem, sdb_stream(stream)>>>(p); break;
    case 3: attn_bwd_dq_umma_kernel<3><<<grid, kBdThreads, kBdSmem, sdb_stream(stream)>>>(p); break;
    default: attn_bwd_dq_umma_kernel<2><<<grid, kBdThreads, kBdSmem, sdb_stream(stream)>>>(p); break;
  }
  SDB_LAUNCH_OK();
  return SDB_OK;
}

// ================================================================================================================
// dK / dV on the tensor cores: the transposed problem.  One CTA = 256 keys of one (batch, head) (K and V tiles resident),
// streaming 64-query tiles of Q, dO (head-padded) and the rows' lse / delta:
//   S^T = K Q^T,  dP^T = V dO^T  (UMMA into TMEM, all operands K-major as they land),
//   P^T = exp2(S^T c - lse[query]),  dS^T = P^T o (dP^T - delta[query])      (thread = one KEY row; lse / delta are
//   per COLUMN here and come from shared memory as broadcasts),
//   dV += P^T dO,  dK += dS^T Q   (P^T / dS^T as TMEM A operands, the dO / Q tiles re-read as MN-major B operands).
// TMEM columns per warpgroup (256): S^T [0,64), dP^T [64,128), P^T [128,160), dS^T [160,192), dV [192,224), dK [224,256).
namespace sdb {

constexpr int kBkStages = 6;
constexpr int kBkTileBytes = 64 * 32 * 2;                    // 4 KB: a 64-query tile of Q or dO
constexpr int kBkStageBytes = 2 * kBkTileBytes;              // Q | dO
constexpr int kBkLdBytes = 2 * 64 * 4;                       // lse | delta of the 64 queries
constexpr int kBkSmem = 4 * kBdQBytes + kBkStages * (kBkStageBytes + kBkLdBytes) + 1024 + 256;

struct BkParams {
  CUtensorMap tmK, tmV, tmQ, tmdO, tmL, tmD;
  __nv_bfloat16* dk; long long lddk;
  __nv_bfloat16* dv; long long lddv;
  int n, m, heads;
  float scale, scale_log2;
};

template <int kPoly>
__global__ void __launch_bounds__(kBdThreads, 1) attn_bwd_dkv_umma_kernel(const __grid_constant__ BkParams p) {
  extern __shared__ __align__(1024) uint8_t bk_smem[];
  const uint32_t base = (smem_u32(bk_smem) + 1023u) & ~1023u;
  const uint32_t sK = base;                                   // 2 x 8 KB (one 128-key tile per warpgroup)
  const uint32_t sV = base + 2 * kBdQBytes;                   // 2 x 8 KB
  const uint32_t sQd = base + 4 * kBdQBytes;                  // stages x (Q 4 KB | dO 4 KB)
  const uint32_t sLD = sQd + kBkStages * kBkStageBytes;       // stages x (lse 256 B | delta 256 B)
  const uint32_t bars = sLD + kBkStages * kBkLdBytes;
  const uint32_t bar_kv = bars;
  auto bar_q_full = [&](int s) { return bars + 8u * (1 + s); };
  auto bar_q_empty = [&](int s) { return bars + 8u * (1 + kBkStages + s); };
  auto bar_s_full = [&](int w) { return bars + 8u * (1 + 2 * kBkStages + w); };
  auto bar_s_free = [&](int w) { return bars + 8u * (3 + 2 * kBkStages + w); };
  auto bar_p_full = [&](int w) { return bars + 8u * (5 + 2 * kBkStages + w); };
  auto bar_o_full = [&](int w) { return bars + 8u * (7 + 2 * kBkStages + w); };
  const uint32_t tmem_slot = bars + 8u * (9 + 2 * kBkStages);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(bk_smem + (tmem_slot - smem_u32(bk_smem)));
  const float* ld_generic = reinterpret_cast<const float*>(bk_smem + (sLD - smem_u32(bk_smem)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, k0 = blockIdx.x * (2 * kBdBM);
  const int T = (p.n + 63) / 64;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmK); prefetch_tmap(&p.tmV); prefetch_tmap(&p.tmQ);
    prefetch_tmap(&p.tmdO); prefetch_tmap(&p.tmL); prefetch_tmap(&p.tmD);
    mbar_init(bar_kv, 1);
    for (int s = 0; s < kBkStages; ++s) {
      mbar_init(bar_q_full(s), 1);
      mbar_init(bar_q_empty(s), 2);
    }
    for (int w = 0; w < 2; ++w) {
      mbar_init(bar_s_full(w), 1);
      mbar_init(bar_s_free(w), 4);
      mbar_init(bar_p_full(w), 4);
      mbar_init(bar_o_full(w), 1);
    }
    fence_barrier_init();
  }
  if (warp == 3) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp < 4) {
    setmaxnreg_dec<56>();
    if (warp == 0 && lane == 0) {
      // ===================== TMA producer =====================
      mbar_arrive_expect_tx(bar_kv, 4 * kBdQBytes);
      tma_load_3d(sK, &p.tmK, bar_kv, h * 32, k0, b);
      tma_load_3d(sK + kBdQBytes, &p.tmK, bar_kv, h * 32, k0 + kBdBM, b);
      tma_load_3d(sV, &p.tmV, bar_kv, h * 32, k0, b);
      tma_load_3d(sV + kBdQBytes, &p.tmV, bar_kv, h * 32, k0 + kBdBM, b);
      bool ok = true;
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kBkStages;
        ok = mbar_wait(bar_q_empty(s), (((uint32_t)(j / kBkStages)) & 1u) ^ 1u);
        if (!ok) break;
        const uint32_t sq = sQd + (uint32_t)s * kBkStageBytes;
        const uint32_t sl = sLD + (uint32_t)s * kBkLdBytes;
        mbar_arrive_expect_tx(bar_q_full(s), kBkStageBytes + kBkLdBytes);
        tma_load_3d(sq, &p.tmQ, bar_q_full(s), h * 32, j * 64, b);
        tma_load_3d(sq + kBkTileBytes, &p.tmdO, bar_q_full(s), h * 32, j * 64, b);
        tma_load_2d(sl, &p.tmL, bar_q_full(s), j * 64, b * p.heads + h);
        tma_load_2d(sl + 256u, &p.tmD, bar_q_full(s), j * 64, b * p.heads + h);
      }
    } else if (warp == 1 || warp == 2) {
      // ===================== MMA issuers =====================
      const int w = warp - 1;
      const uint32_t idesc_s = umma_idesc_bf16(kBdBM, 64);
      const uint32_t idesc_o = umma_idesc_bf16(kBdBM, 32) | (1u << 16);   // B read MN-major
      const uint64_t kdesc = umma_smem_desc(sK + (uint32_t)w * kBdQBytes, 64);
      const uint64_t vdesc = umma_smem_desc(sV + (uint32_t)w * kBdQBytes, 64);
      const uint32_t tcol = tmem_base + (uint32_t)w * 256u;
      auto issue_s = [&](int stage) {
        const uint32_t sq = sQd + (uint32_t)stage * kBkStageBytes;
        const uint64_t qdesc = umma_smem_desc(sq, 64);
        const uint64_t dodesc = umma_smem_desc(sq + kBkTileBytes, 64);
        if (elect_one()) {
          umma_f16(tcol, kdesc, qdesc, idesc_s, 0u);                  // S^T = K Q^T
          umma_f16(tcol, kdesc + 2, qdesc + 2, idesc_s, 1u);
          umma_f16(tcol + 64u, vdesc, dodesc, idesc_s, 0u);           // dP^T = V dO^T
          umma_f16(tcol + 64u, vdesc + 2, dodesc + 2, idesc_s, 1u);
          umma_commit(bar_s_full(w));
        }
        __syncwarp();
      };
      bool ok = mbar_wait(bar_kv, 0) && mbar_wait(bar_q_full(0), 0);
      if (ok) {
        tc_fence_after();
        issue_s(0);
      }
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kBkStages, sn = (j + 1) % kBkStages;
        const uint32_t par = (uint32_t)j & 1u;
        if (j + 1 < T) {
          ok = mbar_wait(bar_q_full(sn), ((uint32_t)((j + 1) / kBkStages)) & 1u) && mbar_wait(bar_s_free(w), par);
          if (!ok) break;
          tc_fence_after();
          issue_s(sn);
        }
        ok = mbar_wait(bar_p_full(w), par);
        if (!ok) break;
        tc_fence_after();
        const uint32_t sq = sQd + (uint32_t)s * kBkStageBytes;
        const uint64_t qmn = umma_smem_desc_mn64(sq);
        const uint64_t domn = umma_smem_desc_mn64(sq + kBkTileBytes);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_f16_ts(tcol + 192u, tcol + 128u + (uint32_t)kk * 8u, domn + (uint64_t)(kk * 64), idesc_o,
                        (j | kk) ? 1u : 0u);   // dV += P^T dO
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_f16_ts(tcol + 224u, tcol + 160u + (uint32_t)kk * 8u, qmn + (uint64_t)(kk * 64), idesc_o,
                        (j | kk) ? 1u : 0u);   // dK += dS^T Q
          umma_commit(bar_o_full(w));
          umma_commit(bar_q_empty(s));
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== row warpgroups: one KEY row per thread =====================
    setmaxnreg_inc<224>();
    const int w = (warp - 4) >> 2, quarter = warp & 3;
    const uint32_t tw = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)w * 256u;
    const float c = p.scale_log2;
    const int row = k0 + w * kBdBM + quarter * 32 + lane;
    bool ok = true;
    for (int j = 0; j < T; ++j) {
      const uint32_t par = (uint32_t)j & 1u;
      const int s = j % kBkStages;
      if (ok) ok = mbar_wait(bar_s_full(w), par);
      tc_fence_after();
      uint32_t su[128];                       // [0,64) = S^T row, [64,128) = dP^T row
      tmem_ld_32x32_at<0>(tw + 0, su);
      tmem_ld_32x32_at<32>(tw + 32, su);
      tmem_ld_32x32_at<64>(tw + 64, su);
      tmem_ld_32x32_at<96>(tw + 96, su);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_s_free(w));
      // lse / delta of this tile's queries (TMA-written: acquire through the stage barrier)
      if (ok) ok = mbar_wait(bar_q_full(s), ((uint32_t)(j / kBkStages)) & 1u);
      const float4* sl = reinterpret_cast<const float4*>(ld_generic + s * (kBkLdBytes / 4));
      const float4* sd = sl + 16;
      if (j == T - 1 && (p.n & 63) != 0) {     // ragged query tail: P^T = 0
        const int valid = p.n - j * 64;
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= valid) su[i] = 0xff800000u;
      }
      if (j > 0) {
        if (ok) ok = mbar_wait(bar_o_full(w), par ^ 1u);
        tc_fence_after();
      }
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t pp[16], ps[16];
#pragma unroll
        for (int g4 = 0; g4 < 8; ++g4) {       // four queries per shared-memory broadcast (-lse, -delta)
          const float4 l = sl[ch * 8 + g4], dl = sd[ch * 8 + g4];
          const int i = ch * 32 + 4 * g4;
          float x0, x1, x2, x3, d0, d1, d2, d3;
          ffma2(x0, x1, __uint_as_float(su[i]), __uint_as_float(su[i + 1]), c, c, l.x, l.y);
          ffma2(x2, x3, __uint_as_float(su[i + 2]), __uint_as_float(su[i + 3]), c, c, l.z, l.w);
          if (((2 * g4) & 7) < kPoly) {
            ex2_poly2(x0, x1);
          } else {
            x0 = ex2_mufu(x0);
            x1 = ex2_mufu(x1);
          }
          if (((2 * g4 + 1) & 7) < kPoly) {
            ex2_poly2(x2, x3);
          } else {
            x2 = ex2_mufu(x2);
            x3 = ex2_mufu(x3);
          }
          pp[2 * g4] = pack_bf16x2(x0, x1);
          pp[2 * g4 + 1] = pack_bf16x2(x2, x3);
          fadd2(d0, d1, __uint_as_float(su[64 + i]), __uint_as_float(su[64 + i + 1]), dl.x, dl.y);
          fadd2(d2, d3, __uint_as_float(su[64 + i + 2]), __uint_as_float(su[64 + i + 3]), dl.z, dl.w);
          fmul2(d0, d1, d0, d1, x0, x1);
          fmul2(d2, d3, d2, d3, x2, x3);
          ps[2 * g4] = pack_bf16x2(d0, d1);
          ps[2 * g4 + 1] = pack_bf16x2(d2, d3);
        }
        tmem_st_32x16(tw + 128 + ch * 16, pp);
        tmem_st_32x16(tw + 160 + ch * 16, ps);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full(w));
    }
    if (ok) ok = mbar_wait(bar_o_full(w), ((uint32_t)(T - 1)) & 1u);
    tc_fence_after();
    uint32_t ov[32], ok_[32];
    tmem_ld_32x32(tw + 192, ov);
    tmem_ld_32x32(tw + 224, ok_);
    tmem_ld_wait();
    if (row < p.m && ok) {
      uint4* dvp = reinterpret_cast<uint4*>(p.dv + ((long long)b * p.m + row) * p.lddv + h * 32);
      uint4* dkp = reinterpret_cast<uint4*>(p.dk + ((long long)b * p.m + row) * p.lddk + h * 32);
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        dvp[v] = make_uint4(pack_bf16x2(__uint_as_float(ov[8 * v]), __uint_as_float(ov[8 * v + 1])),
                            pack_bf16x2(__uint_as_float(ov[8 * v + 2]), __uint_as_float(ov[8 * v + 3])),
                            pack_bf16x2(__uint_as_float(ov[8 * v + 4]), __uint_as_float(ov[8 * v + 5])),
                            pack_bf16x2(__uint_as_float(ov[8 * v + 6]), __uint_as_float(ov[8 * v + 7])));
        dkp[v] = make_uint4(
            pack_bf16x2(__uint_as_float(ok_[8 * v]) * p.scale, __uint_as_float(ok_[8 * v + 1]) * p.scale),
            pack_bf16x2(__uint_as_float(ok_[8 * v + 2]) * p.scale, __uint_as_float(ok_[8 * v + 3]) * p.scale),
            pack_bf16x2(__uint_as_float(ok_[8 * v + 4]) * p.scale, __uint_as_float(ok_[8 * v + 5]) * p.scale),
            pack_bf16x2(__uint_as_float(ok_[8 * v + 6]) * p.scale, __uint_as_float(ok_[8 * v + 7]) * p.scale));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace sdb

// dK / dV of the head-padded attention on tcgen05.  dk / dv head-padded [B, m, >= heads*32]; n % 4 == 0 (the lse /
// delta rows are TMA boxes); SDB_EUNSUPPORTED otherwise (the caller then uses sdb_attention_bwd_dkv).
// NOTE: lse and delta are passed NEGATED ([B, heads, n] fp32 of -lse2 and -delta): they are per-column operands here and
// feed packed f32x2 adds straight from shared memory.
extern "C" int sdb_attention_bwd_dkv_p32(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                         const void* dO_pad, int64_t lddo, const float* neg_lse, const float* neg_delta,
                                         void* dk, int64_t lddk, void* dv, int64_t lddv, int B, int heads, int n, int m,
                                         int d, float scale, void* stream) {
  using namespace sdb;
  const float *lse = neg_lse, *delta = neg_delta;
  SDB_CHECK_ARG(q && k && v && dO_pad && lse && delta && dk && dv && B > 0 && heads > 0 && n > 0 && m > 0,
                "attention_bwd_dkv_p32: bad args");
  SDB_CHECK_ARG(d > 0 && d <= 20 && d % 4 == 0, "attention_bwd_dkv_p32: head dim (%d) must be a multiple of 4, <= 20", d);
  if (n % 4 != 0 || ((uintptr_t)lse & 15) || ((uintptr_t)delta & 15)) {
    sdb_set_error("attention_bwd_dkv_p32: n (%d) must be a multiple of 4 and lse / delta 16-byte aligned", n);
    return SDB_EUNSUPPORTED;
  }
  SDB_CHECK_ARG(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && lddo % 8 == 0 && lddk % 8 == 0 && lddv % 8 == 0 &&
                    ldq >= heads * 32 && ldk >= heads * 32 && ldv >= heads * 32 && lddo >= heads * 32 &&
                    lddk >= heads * 32 && lddv >= heads * 32,
                "attention_bwd_dkv_p32: row strides must be multiples of 8 covering heads*32 padded columns");
  SDB_CHECK_ARG(((uintptr_t)q & 15) == 0 && ((uintptr_t)k & 15) == 0 && ((uintptr_t)v & 15) == 0 &&
                    ((uintptr_t)dO_pad & 15) == 0 && ((uintptr_t)dk & 15) == 0 && ((uintptr_t)dv & 15) == 0,
                "attention_bwd_dkv_p32: operands must be 16-byte aligned");
  SDB_CHECK_ARG(heads <= 65535 && B <= 65535, "attention_bwd_dkv_p32: grid too large");
  BkParams p;
  memset(&p, 0, sizeof(p));
  auto make = [&](CUtensorMap* tm, const void* base, int64_t ld, int rows, uint32_t box_rows) {
    const uint32_t box[3] = {32, box_rows, 1};
    uint64_t dims[3] = {(uint64_t)heads * 32, (uint64_t)rows, (uint64_t)B};
    uint64_t st[3] = {0, (uint64_t)ld * 2, (uint64_t)rows * ld * 2};
    return sdb_make_tensor_map(tm, base, 2, 3, dims, st, box, 64);
  };
  auto make_f32 = [&](CUtensorMap* tm, const float* base) {
    const uint32_t box[2] = {64, 1};
    uint64_t dims[2] = {(uint64_t)n, (uint64_t)B * heads};
    uint64_t st[2] = {0, (uint64_t)n * 4};
    return sdb_make_tensor_map(tm, base, 4, 2, dims, st, box, 0);
  };
  int rc;
  if ((rc = make(&p.tmK, k, ldk, m, kBdBM))) return rc;
  if ((rc = make(&p.tmV, v, ldv, m, kBdBM))) return rc;
  if ((rc = make(&p.tmQ, q, ldq, n, 64))) return rc;
  if ((rc = make(&p.tmdO, dO_pad, lddo, n, 64))) return rc;
  if ((rc = make_f32(&p.tmL, lse))) return rc;
  if ((rc = make_f32(&p.tmD, delta))) return rc;
  p.dk = (__nv_bfloat16*)dk; p.lddk = lddk; p.dv = (__nv_bfloat16*)dv; p.lddv = lddv;
  p.n = n; p.m = m; p.heads = heads;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
  static bool attr_set[64] = {false};
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    SDB_CUDA_OK(cudaFuncSetAttribute(attn_bwd_dkv_umma_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBkSmem));
    SDB_CUDA_OK(cudaFuncSetAttribute(attn_bwd_dkv_umma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBkSmem));
    SDB_CUDA_OK(cudaFuncSetAttribute(attn_bwd_dkv_umma_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBkSmem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  dim3 grid((m + 2 * kBdBM - 1) / (2 * kBdBM), heads, B);
  switch (bwd_poly_pairs()) {
    case 0: attn_bwd_dkv_umma_kernel<0><<<grid, kBdThreads, kBkSmem, sdb_stream(stream)>>>(p); break;
    case 3: attn_bwd_dkv_umma_kernel<3><<<grid, kBdThreads, kBkSmem, sdb_stream(stream)>>>(p); break;
    default: attn_bwd_dkv_umma_kernel<2><<<grid, kBdThreads, kBkSmem, sdb_stream(stream)>>>(p); break;
  }
  SDB_LAUNCH_OK();
  return SDB_OK;
}

// ================================================================================================================
// One pass over dO and O ahead of the two tcgen05 backward kernels: head-pads dO (32 columns per head, pad = 0), forms
// delta = rowsum(dO o O) per (batch, head, query) and the negated copies the transposed dK/dV problem consumes.
namespace sdb {

__global__ void attn_bwd_prep_kernel(const __nv_bfloat16* __restrict__ dO, long long lddo, const __nv_bfloat16* __restrict__ O,
                                     long long ldo, const float* __restrict__ lse, __nv_bfloat16* __restrict__ do_pad,
                                     float* __restrict__ delta, float* __restrict__ neg_delta, float* __restrict__ neg_lse,
                                     long long rows, int n, int heads, int d) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (row, head), head fastest
  if (t >= rows * heads) return;
  const long long r = t / heads;
  const int h = (int)(t - r * heads);
  const uint2* a = reinterpret_cast<const uint2*>(dO + r * lddo + h * d);
  const uint2* o = reinterpret_cast<const uint2*>(O + r * ldo + h * d);
  uint32_t w[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) w[i] = 0u;
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    if (4 * i < d) {
      const uint2 x = __ldg(a + i), y = __ldg(o + i);
      w[2 * i] = x.x;
      w[2 * i + 1] = x.y;
      acc = fmaf(__uint_as_float(x.x << 16), __uint_as_float(y.x << 16), acc);
      acc = fmaf(__uint_as_float(x.x & 0xffff0000u), __uint_as_float(y.x & 0xffff0000u), acc);
      acc = fmaf(__uint_as_float(x.y << 16), __uint_as_float(y.y << 16), acc);
      acc = fmaf(__uint_as_float(x.y & 0xffff0000u), __uint_as_float(y.y & 0xffff0000u), acc);
    }
  }
  uint4* dst = reinterpret_cast<uint4*>(do_pad + t * 32);
#pragma unroll
  for (int v = 0; v < 4; ++v) dst[v] = make_uint4(w[4 * v], w[4 * v + 1], w[4 * v + 2], w[4 * v + 3]);
  const long long b = r / n;
  const long long idx = (b * heads + h) * n + (r - b * n);
  delta[idx] = acc;
  neg_delta[idx] = -acc;
  neg_lse[idx] = -lse[idx];
}

}  // namespace sdb

// dO, O: [B, n, heads*d] bf16 (row strides lddo / ldo, multiples of 4); lse: [B, heads, n] fp32 from the forward.
// Writes do_pad [B, n, heads, 32] bf16 and delta, neg_delta, neg_lse [B, heads, n] fp32.  d % 4 == 0, d <= 20.
extern "C" int sdb_attention_bwd_prep(const void* dO, int64_t lddo, const void* O, int64_t ldo, const float* lse, void* do_pad,
                                      float* delta, float* neg_delta, float* neg_lse, int B, int heads, int n, int d,
                                      void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(dO && O && lse && do_pad && delta && neg_delta && neg_lse && B > 0 && heads > 0 && n > 0,
                "attention_bwd_prep: bad args");
  SDB_CHECK_ARG(d > 0 && d <= 20 && d % 4 == 0, "attention_bwd_prep: head dim (%d) must be a multiple of 4, <= 20", d);
  SDB_CHECK_ARG(lddo % 4 == 0 && ldo % 4 == 0 && lddo >= heads * d && ldo >= heads * d && ((uintptr_t)dO & 7) == 0 &&
                    ((uintptr_t)O & 7) == 0 && ((uintptr_t)do_pad & 15) == 0,
                "attention_bwd_prep: rows must be 8-byte aligned (strides multiples of 4), do_pad 16-byte aligned");
  const long long rows = (long long)B * n, total = rows * heads;
  const int threads = 256;
  attn_bwd_prep_kernel<<<(unsigned)((total + threads - 1) / threads), threads, 0, sdb_stream(stream)>>>(
      (const __nv_bfloat16*)dO, lddo, (const __nv_bfloat16*)O, ldo, lse, (__nv_bfloat16*)do_pad, delta, neg_delta, neg_lse,
      rows, n, heads, d);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
