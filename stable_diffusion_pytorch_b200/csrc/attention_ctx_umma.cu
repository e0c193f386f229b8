// K4d: cross-attention over a SHORT context (CLIP's 77 tokens; any m <= 96) on tcgen05 / TMEM / TMA.
// (reference: modules/transformer.py:89-114 with context_emb = the text embedding)
//
// The long-sequence kernel (attention_umma_wg1.cu) streams key tiles past a fixed query tile; here the whole context
// is ONE key tile, so the roles flip: K and V of a (batch, head) are loaded once per CTA and QUERY tiles stream
// through a TMA ring.  Per 128-query tile: S = Q K^T (128 x NS fp32, NS = m rounded up to 16) by two K=16
// tcgen05.mma into TMEM, one query row per softmax thread (single pass: the row maximum is exact, no online rescale),
// P written back as packed bf16 over the columns S just vacated, O = P V by NS/16 MMAs (V MN-major straight from the
// TMA box, ones column -> denominator), normalise, store.  The MMA warp issues S of tile t+1 right behind P.V of
// tile t (in-order tensor pipe: no barrier needed for the aliased columns).  Two softmax warpgroups take alternate
// tiles, each with its own score buffer and accumulator, so one exponentiates while the other waits for its P.V and
// writes out.  TMEM per CTA: S/P [0, 96) [96, 192) + O [192, 224) [224, 256) -> two CTAs (16 row warps) per SM.  Replaces the mma.sync kernel attention_ctx.cu on the inference path; operands are the
// head-padded ones of sdb_attention_p32.
#include "common.cuh"

namespace sdb {

constexpr int kCuThreads = 384;     // warp 0: TMA, warp 1: MMA issue, warp 2: TMEM alloc, warp 3: idle;
                                    // warps 4-7: rows of the even tiles, warps 8-11: rows of the odd tiles
constexpr int kCuBM = 128;
constexpr int kCuQStages = 3;
constexpr int kCuQBytes = kCuBM * 32 * 2;                 // 8 KB per query tile
constexpr int kCuMaxNS = 96;
constexpr int kCuKVBytes = kCuMaxNS * 32 * 2;             // 6 KB each for K and V
constexpr int kCuStageBytes = 32 * 48;                    // one row warp's output rows (d <= 24), staged for coalescing
constexpr int kCuSmem = kCuQStages * kCuQBytes + 2 * kCuKVBytes + 256 + 8 * kCuStageBytes + 1024;
constexpr int kCuRowRegs = 104;   // 2 CTAs x (128 x 24 + 256 x 104) <= the 2 x 384 x 80 registers the launch reserves
constexpr int kCuTmemCols = 256;

struct CuParams {
  CUtensorMap tmQ, tmK, tmV;
  __nv_bfloat16* out;
  long long ldo;
  int n, m, d, heads;
  int tiles_per_cta;
  float scale_log2;
};

__device__ __forceinline__ void tmem_st_32x8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
      : "memory");
}

// NS = keys per tile (multiple of 16, <= 96); kPoly = element pairs out of every 8 exponentiated on the FMA pipe
#ifndef CU_WAIT
#define CU_WAIT mbar_wait_spin
#endif
template <int NS, int kPoly>
__global__ void __launch_bounds__(kCuThreads, 2) attention_ctx_umma_kernel(const __grid_constant__ CuParams p) {
  extern __shared__ __align__(1024) uint8_t cu_smem_raw[];
  const uint32_t base = (smem_u32(cu_smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = cu_smem_raw + (base - smem_u32(cu_smem_raw));
  const uint32_t sQ = base;                                        // ring of query tiles
  const uint32_t sK = base + kCuQStages * kCuQBytes;
  const uint32_t sV = sK + kCuKVBytes;
  const uint32_t bars = sV + kCuKVBytes;
  const uint32_t bar_kv = bars;
  auto bar_q_full = [&](int s) { return bars + 8u * (1 + s); };
  auto bar_q_empty = [&](int s) { return bars + 8u * (1 + kCuQStages + s); };
  const uint32_t bar_s0 = bars + 8u * (1 + 2 * kCuQStages);
  auto bar_s_full = [&](int i) { return bar_s0 + 8u * i; };          // two score buffers, two accumulators
  auto bar_p_full = [&](int i) { return bar_s0 + 16u + 8u * i; };
  auto bar_o_full = [&](int i) { return bar_s0 + 32u + 8u * i; };
  const uint32_t tmem_slot = bar_s0 + 48u;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heads fastest: the CTAs that read the eight 64-byte head segments of the same 512-byte query rows run together
  const int b = blockIdx.z, h = blockIdx.x;
  const int n_tiles = (p.n + kCuBM - 1) / kCuBM;
  const int t0 = blockIdx.y * p.tiles_per_cta;
  const int T = n_tiles - t0 < p.tiles_per_cta ? n_tiles - t0 : p.tiles_per_cta;   // query tiles of this CTA

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmQ);
    prefetch_tmap(&p.tmK);
    prefetch_tmap(&p.tmV);
    mbar_init(bar_kv, 1);
    for (int s = 0; s < kCuQStages; ++s) {
      mbar_init(bar_q_full(s), 1);
      mbar_init(bar_q_empty(s), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_s_full(i), 1);
      mbar_init(bar_p_full(i), 4);
      mbar_init(bar_o_full(i), 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, kCuTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;      // S/P buffers [0, 96) [96, 192), O buffers [192, 224) [224, 256)

  if (warp < 4) {
    setmaxnreg_dec<24>();
    if (warp == 0 && lane == 0) {
      // ===================== TMA producer =====================
      mbar_arrive_expect_tx(bar_kv, 2u * NS * 64u);
      tma_load_3d(sK, &p.tmK, bar_kv, h * 32, 0, b);       // rows >= m: zero-filled by the TMA unit
      tma_load_3d(sV, &p.tmV, bar_kv, h * 32, 0, b);
      bool ok = true;
      for (int j = 0; j < T && ok; ++j) {
        const int s = j % kCuQStages;
        ok = mbar_wait(bar_q_empty(s), (((uint32_t)(j / kCuQStages)) & 1u) ^ 1u);
        if (!ok) break;
        mbar_arrive_expect_tx(bar_q_full(s), kCuQBytes);
        tma_load_3d(sQ + (uint32_t)s * kCuQBytes, &p.tmQ, bar_q_full(s), h * 32, (t0 + j) * kCuBM, b);
      }
    } else if (warp == 1) {
      // ===================== MMA issuer =====================
      const uint32_t idesc_s = umma_idesc_bf16(kCuBM, NS);
      const uint32_t idesc_o = umma_idesc_bf16(kCuBM, 32) | (1u << 16);   // B (= V) is MN-major
      const uint64_t kdesc = umma_smem_desc(sK, 64);
      const uint64_t vdesc = umma_smem_desc_mn64(sV);
      bool ok = mbar_wait(bar_kv, 0);
      // S(j) goes to score buffer j & 1, over the columns P(j-2) occupied: the tensor pipe runs MMAs in issue order
      // and P.V(j-2) is issued before it
      auto issue_s = [&](int j) -> bool {
        const int s = j % kCuQStages;
        if (!CU_WAIT(bar_q_full(s), ((uint32_t)(j / kCuQStages)) & 1u)) return false;
        tc_fence_after();
        const uint64_t qdesc = umma_smem_desc(sQ + (uint32_t)s * kCuQBytes, 64);
        if (elect_one()) {
          const uint32_t ts = tmem_base + 96u * (uint32_t)(j & 1);
          umma_f16(ts, qdesc, kdesc, idesc_s, 0u);
          umma_f16(ts, qdesc + 2, kdesc + 2, idesc_s, 1u);
          umma_commit(bar_s_full(j & 1));
          umma_commit(bar_q_empty(s));
        }
        __syncwarp();
        return true;
      };
      if (ok && T > 0) ok = issue_s(0);
      if (ok && T > 1) ok = issue_s(1);
      for (int j = 0; j < T && ok; ++j) {
        // P(j) arrives after the row warps have read O(j-2), the previous tenant of accumulator j & 1
        ok = CU_WAIT(bar_p_full(j & 1), ((uint32_t)(j >> 1)) & 1u);
        if (!ok) break;
        tc_fence_after();
        if (elect_one()) {
          const uint32_t tp = tmem_base + 96u * (uint32_t)(j & 1);
          const uint32_t to = tmem_base + 192u + 32u * (uint32_t)(j & 1);
#pragma unroll
          for (int kk = 0; kk < NS / 16; ++kk)
            umma_f16_ts(to, tp + (uint32_t)kk * 8u, vdesc + (uint64_t)(kk * 64), idesc_o, kk ? 1u : 0u);
          umma_commit(bar_o_full(j & 1));
        }
        __syncwarp();
        if (j + 2 < T) ok = issue_s(j + 2);     // scores of tile j+2 are ready long before the row warps get there
      }
    }
  } else {
    // ===================== softmax warpgroup: one query row per thread, the whole context in one pass =============
    setmaxnreg_inc<kCuRowRegs>();
    const int quarter = warp & 3;
    const uint32_t tw = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const float c = p.scale_log2;
    bool ok = true;
    // tile range again, from a volatile read: the copy computed above lives in a spill slot of the 24-register warps
    uint32_t by;
    asm volatile("mov.u32 %0, %%ctaid.y;" : "=r"(by));
    const int t0r = (int)by * p.tiles_per_cta;
    const int Tr = min((p.n + kCuBM - 1) / kCuBM - t0r, p.tiles_per_cta);
    const int dq = p.d >> 2;                            // 8-byte chunks per output row
    const float rdq = 1.0f / (float)dq;
    uint2* stage = reinterpret_cast<uint2*>(gen + (bars + 256u - base)) + (warp - 4) * (kCuStageBytes / 8);
    // normalise and store tile jj from accumulator jj & 1
    auto emit = [&](int jj) {
      if (ok) ok = CU_WAIT(bar_o_full(jj & 1), ((uint32_t)(jj >> 1)) & 1u);
      tc_fence_after();
      uint32_t o[32];
      tmem_ld_32x32(tw + 192u + 32u * (uint32_t)(jj & 1), o);
      tmem_ld_wait();
      tc_fence_before();
      float den = __uint_as_float(o[0]);        // the ones column of V sits at column d; selp keeps o[] in registers
#pragma unroll
      for (int i = 1; i < 24; ++i)
        asm("{\n.reg .pred q;\nsetp.eq.s32 q, %2, %3;\nselp.f32 %0, %1, %0, q;\n}"
            : "+f"(den) : "f"(__uint_as_float(o[i])), "r"(p.d), "r"(i));
      const float inv = 1.0f / den;
      // a thread owns a row, but rows are ldo apart: 32 lanes x 8 bytes would be 32 separate lines per store.  Through
      // shared memory the warp writes its 32 x 2d bytes as runs of whole rows instead (5x fewer LSU wavefronts at d=20)
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 24; i += 4)
        if (i < p.d)
          stage[lane * dq + (i >> 2)] =
              make_uint2(pack_bf16x2(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv),
                         pack_bf16x2(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv));
      __syncwarp();
      const int row0 = (t0r + jj) * kCuBM + quarter * 32;
      for (int it = 0; it < dq; ++it) {
        const int idx = it * 32 + lane;
        const int r = __float2int_rz(((float)idx + 0.5f) * rdq);
        const int cc = idx - r * dq;
        if (row0 + r < p.n && ok)
          *reinterpret_cast<uint2*>(p.out + ((long long)b * p.n + row0 + r) * p.ldo + h * p.d + 4 * cc) = stage[idx];
      }
    };
    const int wg = (warp >> 2) - 1;                     // this warpgroup owns score buffer / accumulator wg
    for (int j = wg; j < Tr; j += 2) {
      const uint32_t ts = tw + 96u * (uint32_t)(j & 1);
      if (ok) ok = CU_WAIT(bar_s_full(j & 1), ((uint32_t)(j >> 1)) & 1u);
      tc_fence_after();
      uint32_t su[NS];
#pragma unroll
      for (int g = 0; g < NS / 16; ++g)
        tmem_ld_32x16(ts + 16u * g, *reinterpret_cast<uint32_t(*)[16]>(&su[16 * g]));
      tmem_ld_wait();
      if (p.m < NS) {                                   // padded keys (only in the last 16): zero rows of K gave s = 0
#pragma unroll
        for (int i = NS - 16; i < NS; ++i)
          if (i >= p.m) su[i] = 0xff800000u;
      }
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < NS; i += 8) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          mx[k] = max3(mx[k], __uint_as_float(su[i + 2 * k]), __uint_as_float(su[i + 2 * k + 1]));
      }
      const float negm = -fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * c;
#pragma unroll
      for (int g = 0; g < NS / 16; ++g) {
        uint32_t pk[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = g * 16 + 2 * k;
          float x0, x1;
          ffma2(x0, x1, __uint_as_float(su[i]), __uint_as_float(su[i + 1]), c, c, negm, negm);
          if ((k * kPoly) % 8 < kPoly) {
            ex2_poly2(x0, x1);
          } else {
            x0 = ex2_mufu(x0);
            x1 = ex2_mufu(x1);
          }
          pk[k] = __byte_perm(__float_as_uint(x0), __float_as_uint(x1), 0x7632);
        }
        tmem_st_32x8(ts + 8u * g, pk);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full(j & 1));
      // the other warpgroup's softmax fills the issue slots while P.V(j) runs; S(j+2) is issued right behind it
      emit(j);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kCuTmemCols);
  }
}

template <int NS>
static int launch_attention_ctx_umma(const CuParams& p, dim3 grid, cudaStream_t stream) {
  static bool attr_set[64] = {false};
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    SDB_CUDA_OK(cudaFuncSetAttribute(attention_ctx_umma_kernel<NS, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCuSmem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  attention_ctx_umma_kernel<NS, 2><<<grid, kCuThreads, kCuSmem, stream>>>(p);
  return SDB_OK;
}

}  // namespace sdb
