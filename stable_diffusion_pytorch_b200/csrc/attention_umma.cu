// K4 (tensor-core path): flash-style attention for small head dims on tcgen05 / TMEM / TMA.
// (reference: modules/transformer.py:93-114 -- softmax(Q K^T * scale) V per (batch, head))
//
// The UNet's head dim is 20 (SURVEY top item 7): per (query, key) pair there are 80 useful flops
// and one exp, so the kernel is bound by the MUFU ex2 pipe (16/clk/SM measured) and the issue
// slots of the softmax warps, never by the tensor pipe.  Design:
//   * operands arrive HEAD-PADDED from the projection GEMM: head h of Q / K / V lives at columns
//     [32h, 32h+32) with zeros beyond d for Q and K, and V carries a column of ones at index d
//     so that the P.V product also yields the softmax denominator (no per-element adds);
//   * one CTA = 256 queries of one (batch, head): two softmax warpgroups (128 rows each, ONE ROW
//     PER THREAD -> no shuffles), one TMA warp, one MMA-issuing warp;
//   * S = Q K^T (128 x 128 fp32) is produced in TMEM by two K=16 tcgen05.mma, read back with
//     tcgen05.ld, exponentiated, written back to TMEM as packed bf16 P and consumed directly as
//     the A operand of the P.V tcgen05.mma (B = V tile, MN-major straight from the TMA box);
//   * ~3/8 of the exponentials are evaluated on the FMA pipe (Cody-Waite split + degree-3
//     polynomial in packed f32x2 arithmetic) so MUFU and FMA pipes finish together;
//   * O (128 x 32 fp32) is read back per key tile and folded into registers with the usual
//     online-softmax rescale (cheap: 21 live columns).
#include <stdlib.h>

#include "common.cuh"
#include "umma_helpers.cuh"

namespace sdb {

constexpr int kFaThreads = 384;  // warp 0: TMA, 1-2: MMA issue (one per warpgroup), 3: TMEM alloc; 4-7, 8-11: softmax
constexpr int kFaStages = 4;
constexpr int kFaBM = 128;       // query rows per softmax warpgroup
constexpr int kFaBN = 128;       // keys per tile
constexpr int kFaDP = 32;        // padded head dim (64-byte rows, SWIZZLE_64B)
constexpr int kFaQBytes = kFaBM * kFaDP * 2;       // 8 KB
constexpr int kFaKVBytes = kFaBN * kFaDP * 2;      // 8 KB each for K and V
constexpr int kFaStageBytes = 2 * kFaKVBytes;
constexpr int kFaSmem = 2 * kFaQBytes + kFaStages * kFaStageBytes + 1024 /*align*/ + 256 /*barriers*/;

struct FaParams {
  CUtensorMap tmQ, tmK, tmV;
  __nv_bfloat16* out;
  long long ldo;
  int n, m, d, heads;
  float scale_log2;
  float* lse;   // optional [B, heads, n]: log2-domain log-sum-exp of the scaled scores (kept for the backward)
};

// kPoly = number of element PAIRS out of every 8 pairs exponentiated on the FMA pipe
template <int kPoly>
__global__ void __launch_bounds__(kFaThreads, 1)
attention_umma_kernel(const __grid_constant__ FaParams p) {
  extern __shared__ __align__(1024) uint8_t fa_smem[];
  const uint32_t base = (smem_u32(fa_smem) + 1023u) & ~1023u;
  const uint32_t sQ = base;                                  // 2 x 8 KB
  const uint32_t sKV = base + 2 * kFaQBytes;                 // stages x (K 8 KB | V 8 KB)
  const uint32_t bars = sKV + kFaStages * kFaStageBytes;
  const uint32_t bar_q = bars;
  auto bar_kv_full = [&](int s) { return bars + 8u * (1 + s); };
  auto bar_kv_empty = [&](int s) { return bars + 8u * (1 + kFaStages + s); };
  auto bar_s_full = [&](int w) { return bars + 8u * (1 + 2 * kFaStages + w); };
  auto bar_s_free = [&](int w) { return bars + 8u * (3 + 2 * kFaStages + w); };
  auto bar_p_full = [&](int w) { return bars + 8u * (5 + 2 * kFaStages + w); };
  auto bar_o_full = [&](int w) { return bars + 8u * (7 + 2 * kFaStages + w); };
  const uint32_t tmem_slot = bars + 8u * (9 + 2 * kFaStages);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(fa_smem + (tmem_slot - smem_u32(fa_smem)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * (2 * kFaBM);
  const int T = (p.m + kFaBN - 1) / kFaBN;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmQ);
    prefetch_tmap(&p.tmK);
    prefetch_tmap(&p.tmV);
    mbar_init(bar_q, 1);
    for (int s = 0; s < kFaStages; ++s) {
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
  // TMEM columns per warpgroup w: S at 256w + [0,128), P at 256w + [128,192), O at 256w + [192,224)

  if (warp < 4) {
    setmaxnreg_dec<56>();
    if (warp == 0 && lane == 0) {
      // ===================== TMA producer =====================
      mbar_arrive_expect_tx(bar_q, 2 * kFaQBytes);
      tma_load_3d(sQ, &p.tmQ, bar_q, h * kFaDP, q0, b);
      tma_load_3d(sQ + kFaQBytes, &p.tmQ, bar_q, h * kFaDP, q0 + kFaBM, b);
      bool ok = true;
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kFaStages;
        ok = mbar_wait(bar_kv_empty(s), (((uint32_t)(j / kFaStages)) & 1u) ^ 1u);
        if (!ok) break;
        const uint32_t sk = sKV + (uint32_t)s * kFaStageBytes;
        mbar_arrive_expect_tx(bar_kv_full(s), kFaStageBytes);
        tma_load_3d(sk, &p.tmK, bar_kv_full(s), h * kFaDP, j * kFaBN, b);
        tma_load_3d(sk + kFaKVBytes, &p.tmV, bar_kv_full(s), h * kFaDP, j * kFaBN, b);
      }
    } else if (warp == 1 || warp == 2) {
      // ===================== MMA issuers: one warp per softmax warpgroup =====================
      // The whole warp runs the loop (warp-uniform control flow keeps descriptors and addresses in
      // uniform registers; measured 2x faster issue than a divergent lane-0 path, and one thread
      // serving both warpgroups was a bottleneck); tcgen05 instructions go out from one elected lane.
      const int w = warp - 1;
      const uint32_t idesc_s = umma_idesc_bf16(kFaBM, kFaBN);
      const uint32_t idesc_o = umma_idesc_bf16(kFaBM, kFaDP) | (1u << 16);  // B (= V) is MN-major
      const uint64_t qdesc = umma_smem_desc(sQ + (uint32_t)w * kFaQBytes, 64);
      const uint32_t tcol = tmem_base + (uint32_t)w * 256u;
      auto issue_s = [&](int stage) {
        const uint64_t bdesc = umma_smem_desc(sKV + (uint32_t)stage * kFaStageBytes, 64);
        if (elect_one()) {
          umma_f16(tcol, qdesc, bdesc, idesc_s, 0u);
          umma_f16(tcol, qdesc + 2, bdesc + 2, idesc_s, 1u);
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
        const int s = j % kFaStages, sn = (j + 1) % kFaStages;
        const uint32_t par = (uint32_t)j & 1u;
        if (j + 1 < T) {
          ok = mbar_wait(bar_kv_full(sn), ((uint32_t)((j + 1) / kFaStages)) & 1u) && mbar_wait(bar_s_free(w), par);
          if (!ok) break;
          tc_fence_after();
          issue_s(sn);
        }
        ok = mbar_wait(bar_p_full(w), par);
        if (!ok) break;
        tc_fence_after();
        const uint64_t vdesc = umma_smem_desc_mn64(sKV + (uint32_t)s * kFaStageBytes + kFaKVBytes);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kFaBN / 16; ++kk)
            umma_f16_ts(tcol + 192u, tcol + 128u + (uint32_t)kk * 8u, vdesc + (uint64_t)(kk * 64), idesc_o,
                        (j | kk) ? 1u : 0u);   // O accumulates in TMEM across the key tiles
          umma_commit(bar_o_full(w));
          umma_commit(bar_kv_empty(s));   // one arrival per warpgroup
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== softmax warpgroups: one query row per thread =====================
    setmaxnreg_inc<224>();
    const int w = (warp - 4) >> 2, quarter = warp & 3;
    const uint32_t tw = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)w * 256u;
    const float c = p.scale_log2;
    // O stays in TMEM for the whole key loop, relative to a LAGGING reference maximum m_run: it is only
    // rescaled (TMEM load / multiply / store) when some row of the warp saw its maximum grow by more than
    // 2^kLazyLog2 -- until then P = exp2((s - m_run) c) may exceed 1 by at most that factor, harmless in
    // bf16 / fp32.  Typical rows settle after the first few tiles, so the per-tile cost is one barrier probe.
    constexpr float kLazyLog2 = 8.0f;
    float m_run = -INFINITY;
    bool ok = true;
    for (int j = 0; j < T; ++j) {
      const uint32_t par = (uint32_t)j & 1u;
      if (ok) ok = mbar_wait(bar_s_full(w), par);
      tc_fence_after();
      uint32_t su[128];
      tmem_ld_32x32_at<0>(tw + 0, su);
      tmem_ld_32x32_at<32>(tw + 32, su);
      tmem_ld_32x32_at<64>(tw + 64, su);
      tmem_ld_32x32_at<96>(tw + 96, su);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_s_free(w));
      // ragged key tail: keys >= m contribute exp2(-inf) = 0
      if (j == T - 1 && (p.m & (kFaBN - 1)) != 0) {
        const int valid = p.m - j * kFaBN;
#pragma unroll
        for (int i = 0; i < 128; ++i)
          if (i >= valid) su[i] = 0xff800000u;
      }
      // row max (3-input max, 4 independent chains)
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 128; i += 8) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          mx[k] = max3(mx[k], __uint_as_float(su[i + 2 * k]), __uint_as_float(su[i + 2 * k + 1]));
      }
      const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      // the previous tile's P.V must have finished reading P (and writing O) before either is touched
      if (j > 0) {
        if (ok) ok = mbar_wait(bar_o_full(w), par ^ 1u);
        tc_fence_after();
      }
      const bool grow = (m_tile - m_run) * c > kLazyLog2;   // first tile: m_run = -inf -> true
      if (j == 0) {
        m_run = m_tile;
      } else if (__any_sync(0xffffffffu, grow)) {
        const float m_new = grow ? m_tile : m_run;
        const float alpha = ex2_mufu((m_run - m_new) * c);
        m_run = m_new;
        uint32_t oa[16], ob[16];
        tmem_ld_32x16(tw + 192, oa);
        tmem_ld_32x16(tw + 208, ob);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          oa[i] = __float_as_uint(__uint_as_float(oa[i]) * alpha);
          ob[i] = __float_as_uint(__uint_as_float(ob[i]) * alpha);
        }
        tmem_st_32x16(tw + 192, oa);
        tmem_st_32x16(tw + 208, ob);
      }
      const float negm = -m_run * c;
      // P = exp2(s*c - m*c) -> bf16 by truncation (one PRMT per pair instead of a half-rate F2FP; numerator
      // and denominator -- the ones column of V -- see the same P, so the bias cancels) -> TMEM (A operand)
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        uint32_t pk[16];
#pragma unroll
        for (int g = 0; g < 2; ++g) {        // 8 pairs per group
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int i = ch * 32 + g * 16 + 2 * k;
            float x0, x1;
            ffma2(x0, x1, __uint_as_float(su[i]), __uint_as_float(su[i + 1]), c, c, negm, negm);
            if (k < kPoly) {
              ex2_poly2(x0, x1);
            } else {
              x0 = ex2_mufu(x0);
              x1 = ex2_mufu(x1);
            }
            pk[g * 8 + k] = __byte_perm(__float_as_uint(x0), __float_as_uint(x1), 0x7632);
          }
        }
        tmem_st_32x16(tw + 128 + ch * 16, pk);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full(w));
    }
    // ---- last tile's P.V, normalise, store ---------------------------------------------------
    if (ok) ok = mbar_wait(bar_o_full(w), ((uint32_t)(T - 1)) & 1u);
    tc_fence_after();
    float acc[kFaDP];
    {
      uint32_t o[32];
      tmem_ld_32x32(tw + 192, o);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < kFaDP; ++i) acc[i] = __uint_as_float(o[i]);
    }
    const int row = q0 + w * kFaBM + quarter * 32 + lane;
    if (row < p.n && ok) {
      float den = acc[0];
#pragma unroll
      for (int i = 1; i < 24; ++i)
        if (i == p.d) den = acc[i];
      const float inv = 1.0f / den;
      // O and its denominator are relative to the lagging maximum m_run: lse2 = m_run c + log2(den)
      if (p.lse) p.lse[((long long)b * p.heads + h) * p.n + row] = fmaf(m_run, c, log2f(den));
      __nv_bfloat16* og = p.out + ((long long)b * p.n + row) * p.ldo + h * p.d;
#pragma unroll
      for (int i = 0; i < 24; i += 4) {
        if (i < p.d) {
          uint2 v = make_uint2(pack_bf16x2(acc[i] * inv, acc[i + 1] * inv),
                               pack_bf16x2(acc[i + 2] * inv, acc[i + 3] * inv));
          *reinterpret_cast<uint2*>(og + i) = v;
        }
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

#include "attention_umma_wg1.cu"   // the one-warpgroup-per-CTA shape of the same pipeline (uses the helpers above)

static int attention_p32_impl(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                              int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                              float scale, float* lse, int* lse_written, void* stream) {
  using namespace sdb;
  if (lse_written) *lse_written = 0;
  SDB_CHECK_ARG(q && k && v && out && B > 0 && heads > 0 && n > 0 && m > 0, "attention_p32: bad args");
  SDB_CHECK_ARG(d > 0 && d <= 20 && d % 4 == 0, "attention_p32: head dim (%d) must be a multiple of 4, <= 20", d);
  SDB_CHECK_ARG(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 4 == 0,
                "attention_p32: ldq/ldk/ldv must be multiples of 8 (16-byte TMA rows), ldo of 4");
  SDB_CHECK_ARG(ldq >= heads * 32 && ldk >= heads * 32 && ldv >= heads * 32,
                "attention_p32: row strides must cover heads*32 padded columns");
  SDB_CHECK_ARG(((uintptr_t)q & 15) == 0 && ((uintptr_t)k & 15) == 0 && ((uintptr_t)v & 15) == 0 &&
                    ((uintptr_t)out & 7) == 0,
                "attention_p32: q/k/v must be 16-byte aligned, out 8-byte aligned");
  SDB_CHECK_ARG(heads <= 65535 && B <= 65535, "attention_p32: grid too large");
  // a context of <= 128 tokens (cross-attention over CLIP's 77) is one key tile: the streaming mma.sync
  // kernel beats this pipeline's per-CTA set-up by an order of magnitude there
  static int no_ctx = -1;
  if (no_ctx < 0) { const char* e = getenv("SDB_FA_NO_CTX"); no_ctx = e ? atoi(e) : 0; }
  if (m <= 128 && !no_ctx)
    return sdb_attention_ctx_dispatch(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, sdb_stream(stream));
  FaParams p;
  memset(&p, 0, sizeof(p));
  // kernel shape: 0 = two warpgroups per CTA, 128-key tiles; 3 / 4 = one warpgroup per CTA, 64-key tiles, that many
  // CTAs per SM.  Default 4 (config 2: 1.105 vs 1.199 ms per layer, 831 vs 795 latent-steps/s); SDB_FA_WG1 overrides
  static int wg1 = -1;
  if (wg1 < 0) {
    const char* e = getenv("SDB_FA_WG1");
    wg1 = e ? atoi(e) : 4;
    if (wg1 != 3 && wg1 != 4) wg1 = 0;
  }
  const uint32_t box[3] = {32, 128, 1};
  const uint32_t box_kv[3] = {32, wg1 ? 64u : 128u, 1};
  {
    uint64_t dims[3] = {(uint64_t)heads * 32, (uint64_t)n, (uint64_t)B};
    uint64_t st[3] = {0, (uint64_t)ldq * 2, (uint64_t)n * ldq * 2};
    int rc = sdb_make_tensor_map(&p.tmQ, q, 2, 3, dims, st, box, 64);
    if (rc) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)heads * 32, (uint64_t)m, (uint64_t)B};
    uint64_t st[3] = {0, (uint64_t)ldk * 2, (uint64_t)m * ldk * 2};
    int rc = sdb_make_tensor_map(&p.tmK, k, 2, 3, dims, st, box_kv, 64);
    if (rc) return rc;
    st[1] = (uint64_t)ldv * 2;
    st[2] = (uint64_t)m * ldv * 2;
    rc = sdb_make_tensor_map(&p.tmV, v, 2, 3, dims, st, box_kv, 64);
    if (rc) return rc;
  }
  p.out = (__nv_bfloat16*)out;
  p.ldo = ldo;
  p.n = n; p.m = m; p.d = d; p.heads = heads;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.lse = lse;
  if (lse && lse_written) *lse_written = 1;
  // share of exponentials evaluated on the FMA pipe: kPoly pairs out of 8 (tuning aid: SDB_FA_POLY)
  static int poly = -1;
  if (poly < 0) {
    const char* e = getenv("SDB_FA_POLY");
    poly = e ? atoi(e) : (wg1 ? 2 : 3);   // measured best per shape
    if (poly < 0 || poly > 6) poly = wg1 ? 2 : 3;
  }
  if (wg1) {
    dim3 grid1((n + kF1BM - 1) / kF1BM, heads, B);
    int rc;
    static int spread = -1;
    if (spread < 0) { const char* e = getenv("SDB_FA_SPREAD"); spread = e ? atoi(e) : 1; }
    // instantiated shapes: the default and its A/B neighbours (tools/experiments/README.md has the full sweep)
    if (wg1 == 4 && spread) rc = poly == 3 ? launch_attention_wg1<3, 4, true>(p, grid1, sdb_stream(stream))
                                           : launch_attention_wg1<2, 4, true>(p, grid1, sdb_stream(stream));
    else if (wg1 == 4) rc = launch_attention_wg1<2, 4>(p, grid1, sdb_stream(stream));
    else rc = launch_attention_wg1<2, 3>(p, grid1, sdb_stream(stream));
    if (rc) return rc;
    SDB_LAUNCH_OK();
    return SDB_OK;
  }
  dim3 grid((n + 2 * kFaBM - 1) / (2 * kFaBM), heads, B);
#define SDB_FA_LAUNCH(P)                                                                                   \
  case P: {                                                                                                \
    static bool attr_set[64] = {false};                                                                    \
    int dev = 0;                                                                                           \
    SDB_CUDA_OK(cudaGetDevice(&dev));                                                                      \
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {                                                          \
      SDB_CUDA_OK(cudaFuncSetAttribute(attention_umma_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                       kFaSmem));                                                          \
      if (dev >= 0 && dev < 64) attr_set[dev] = true;                                                      \
    }                                                                                                      \
    attention_umma_kernel<P><<<grid, kFaThreads, kFaSmem, sdb_stream(stream)>>>(p);                        \
  } break;
  switch (poly) {
    SDB_FA_LAUNCH(0)
    SDB_FA_LAUNCH(2)
    SDB_FA_LAUNCH(3)
    SDB_FA_LAUNCH(4)
    default: sdb_set_error("attention_p32: SDB_FA_POLY=%d is not instantiated for the two-warpgroup shape (0, 2, 3, 4)", poly);
      return SDB_EUNSUPPORTED;
  }
#undef SDB_FA_LAUNCH
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_attention_p32(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                                 int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                                 float scale, void* stream) {
  return attention_p32_impl(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, nullptr, nullptr, stream);
}

extern "C" int sdb_attention_p32_lse(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                                     int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                                     float scale, float* lse, int* lse_written, void* stream) {
  SDB_CHECK_ARG(lse && lse_written, "attention_p32_lse: null lse buffer");
  return attention_p32_impl(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, lse, lse_written, stream);
}
