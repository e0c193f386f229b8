// K1 / K3: persistent warp-specialised tcgen05 GEMM with TMA operand staging and TMEM
// accumulators.  One kernel serves
//   * GEMM  out[M,N] = A[M,K] W[N,K]^T           (A = one or two K segments, 2-D tensor maps)
//   * 3x3 convolution as implicit GEMM over NHWC  (A = 9 shifted 4-D boxes per channel chunk;
//     out-of-image taps are zero-filled by the TMA unit, which is exactly Conv2d(padding=1))
// Roles: warp 0 = TMA producer, warp 1 = MMA issuer (owns TMEM), warps 2..9 = epilogue
// (TMEM -> registers -> fused bias/temb/residual/activation -> global).  Two TMEM accumulator
// stages let the epilogue of tile i overlap the main loop of tile i+1.
#include <stdlib.h>

#include "common.cuh"

namespace sdb {

constexpr int kTileM = 128;
constexpr int kMaxBN = 256;
constexpr int kGemmThreads = 320;  // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int kMaxStages = 12;
constexpr int kSmemBudget = 224 * 1024;  // operand ring + epilogue staging; + barriers / alignment slack

struct UmmaParams {
  CUtensorMap tmA[2];
  CUtensorMap tmB;
  CUtensorMap tmO, tmOt;   // output: 64-column slabs (SWIZZLE_128B) and the narrower tail slab
  CUtensorMap tmR, tmRt;   // bf16 residual, same boxes
  // main loop
  int mode;  // 0 = GEMM, 1 = conv3x3
  int num_kb, kb1, K1;
  int BK, swz, BN;
  int num_stages, stage_bytes, a_bytes, b_bytes;
  int tiles_m, tiles_n;
  // conv geometry (GEMM-M space = input-resolution pixels)
  int kb_per_tap, Cin;
  int TW, TH, TN, tiles_w, tiles_h;
  int B, H, W;
  int out_mode;  // 0 = same resolution, 1 = keep even rows/cols (conv then nearest x0.5)
  signed char tap_map[9], tap_dh[9], tap_dw[9];
  // epilogue
  int M, N;
  const float* bias;
  const float* temb;
  long long ld_temb;
  const void* residual;
  long long ldr;
  int res_dtype;
  void* out;
  long long ldo;
  int out_dtype;
  int act;
  // staged epilogue (bf16 output through swizzled smem slabs + TMA store): 0 = direct stores
  int staged;
  int bno;          // output columns per tile (BN, or BN/2 for GEGLU)
  int nslab;        // full 64-column slabs per tile
  int tail_cols;    // columns of the last, narrower slab (0, 16 or 32)
  int staging_bytes;
};

// exact-erf GELU (transformer.py:177 uses nn.GELU()) with erf from Abramowitz-Stegun 7.1.26
// (|error| <= 1.5e-7, far inside the bf16 rounding of the result): 2 MUFU + ~10 FMA-pipe ops
// instead of the ~30-instruction branchy erff -- the GEGLU epilogue is issue-bound.
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float e = exp2f(-z * z * 1.4426950408889634f);
  const float erf_abs = fmaf(-poly * t, e, 1.0f);          // erf(|x|/sqrt2)
  const float erf_v = copysignf(erf_abs, x);
  return 0.5f * x * (1.0f + erf_v);
}

// ---- epilogue for 16 consecutive accumulator columns of one row -------------------------------
// `res` holds the row's 16 residual values (bf16, prefetched before the accumulator was ready)
// when kResPre, else the residual is read here (fp32 residual: rare, un-prefetched path).
template <bool kResPre>
__device__ __forceinline__ void epilogue16(const UmmaParams& p, const uint32_t (&acc)[16],
                                           long long out_row, int img, int col, const uint4 (&res)[2]) {
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(acc[j]);
  if (p.bias) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col + j));
      v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
    }
  }
  if (p.temb) {
    const float* t = p.temb + (long long)img * p.ld_temb + col;
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      float4 b = __ldg(reinterpret_cast<const float4*>(t + j));
      v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
    }
  }
  if (p.act == SDB_ACT_GEGLU) {
    // interleaved columns (a_j, g_j): 16 accumulator columns -> 8 outputs
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = v[2 * j] * gelu_erf_fast(v[2 * j + 1]);
    const long long off = out_row * p.ldo + (col >> 1);
    if (p.out_dtype == SDB_BF16) {
      uint4 w = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]),
                           pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
      *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off) = w;
    } else {
      float4* d = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + off);
      d[0] = make_float4(o[0], o[1], o[2], o[3]);
      d[1] = make_float4(o[4], o[5], o[6], o[7]);
    }
    return;
  }
  if (p.act == SDB_ACT_SILU) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = silu_f(v[j]);
  }
  if (p.residual) {
    if (kResPre) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint4 q = res[h];
        float2 a = unpack_bf16x2(q.x), b = unpack_bf16x2(q.y), c = unpack_bf16x2(q.z),
               d = unpack_bf16x2(q.w);
        v[8 * h + 0] += a.x; v[8 * h + 1] += a.y; v[8 * h + 2] += b.x; v[8 * h + 3] += b.y;
        v[8 * h + 4] += c.x; v[8 * h + 5] += c.y; v[8 * h + 6] += d.x; v[8 * h + 7] += d.y;
      }
    } else {
      const float4* r = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.residual) +
                                                        out_row * p.ldr + col);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float4 q = __ldg(r + j);
        v[4 * j] += q.x; v[4 * j + 1] += q.y; v[4 * j + 2] += q.z; v[4 * j + 3] += q.w;
      }
    }
  }
  const long long off = out_row * p.ldo + col;
  if (p.out_dtype == SDB_BF16) {
    uint4* d = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off);
    d[0] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                      pack_bf16x2(v[6], v[7]));
    d[1] = make_uint4(pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]),
                      pack_bf16x2(v[12], v[13]), pack_bf16x2(v[14], v[15]));
  } else {
    float4* d = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + off);
#pragma unroll
    for (int j = 0; j < 4; ++j) d[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  }
}

__device__ __forceinline__ void tma_store_2d(const void* tmap, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const void* tmap, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Shared-memory address of the 16-byte chunk holding output columns [oc, oc+8) of tile row r.  The tile is
// staged as 64-column slabs of 128 rows x 128 B (SWIZZLE_128B: chunk ^= r & 7) followed by one narrower tail
// slab (64-byte rows: chunk ^= (r >> 1) & 3; 32-byte rows: chunk ^= (r >> 2) & 1) -- exactly the layouts
// the TMA store / residual-load boxes use, and conflict-free for row-per-thread 16-byte accesses.
__device__ __forceinline__ uint32_t stage_addr(const UmmaParams& p, uint32_t stg, int r, int oc) {
  const int slab = oc >> 6;
  if (slab < p.nslab) {
    const int chunk = (oc & 63) >> 3;
    return stg + (uint32_t)slab * 16384u + (uint32_t)r * 128u + (uint32_t)((chunk ^ (r & 7)) << 4);
  }
  const int chunk = (oc - 64 * p.nslab) >> 3;
  const uint32_t base = stg + (uint32_t)p.nslab * 16384u;
  if (p.tail_cols == 32) return base + (uint32_t)r * 64u + (uint32_t)((chunk ^ ((r >> 1) & 3)) << 4);
  return base + (uint32_t)r * 32u + (uint32_t)((chunk ^ ((r >> 2) & 1)) << 4);
}

// epilogue of 16 accumulator columns into the staged tile: bias / temb / activation, residual read from
// (and result written back to) the swizzled slab the TMA unit loaded / will store.
__device__ __forceinline__ void epilogue16_staged(const UmmaParams& p, const uint32_t (&acc)[16], uint32_t stg,
                                                  int r, int img, int col, int col_in_tile) {
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(acc[j]);
  if (p.bias) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col + j));
      v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
    }
  }
  if (p.temb) {
    const float* t = p.temb + (long long)img * p.ld_temb + col;
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      float4 b = __ldg(reinterpret_cast<const float4*>(t + j));
      v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
    }
  }
  if (p.act == SDB_ACT_GEGLU) {
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = v[2 * j] * gelu_erf_fast(v[2 * j + 1]);
    sts128(stage_addr(p, stg, r, col_in_tile >> 1),
           make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                      pack_bf16x2(o[6], o[7])));
    return;
  }
  if (p.act == SDB_ACT_SILU) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = silu_f(v[j]);
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const uint32_t a = stage_addr(p, stg, r, col_in_tile + 8 * h);
    if (p.residual) {
      const uint4 q = lds128(a);
      float2 x = unpack_bf16x2(q.x), y = unpack_bf16x2(q.y), z = unpack_bf16x2(q.z), w = unpack_bf16x2(q.w);
      v[8 * h + 0] += x.x; v[8 * h + 1] += x.y; v[8 * h + 2] += y.x; v[8 * h + 3] += y.y;
      v[8 * h + 4] += z.x; v[8 * h + 5] += z.y; v[8 * h + 6] += w.x; v[8 * h + 7] += w.y;
    }
    sts128(a, make_uint4(pack_bf16x2(v[8 * h], v[8 * h + 1]), pack_bf16x2(v[8 * h + 2], v[8 * h + 3]),
                         pack_bf16x2(v[8 * h + 4], v[8 * h + 5]), pack_bf16x2(v[8 * h + 6], v[8 * h + 7])));
  }
}

__global__ void __launch_bounds__(kGemmThreads, 1)
umma_gemm_kernel(const __grid_constant__ UmmaParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-B aligned operand ring (SWIZZLE_128B atoms), then the barriers
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring_bytes = (uint32_t)p.num_stages * (uint32_t)p.stage_bytes;
  const uint32_t stg_base = smem_base + ring_bytes;                 // epilogue staging (1024-B aligned)
  const uint32_t bar_base = stg_base + (uint32_t)p.staging_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kMaxStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kMaxStages + 2 + a); };
  const uint32_t res_bar = bar_base + 8u * (2 * kMaxStages + 4);
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 5);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmA[0]);
    prefetch_tmap(&p.tmB);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 256);
    }
    mbar_init(res_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  const int total_tiles = p.tiles_m * p.tiles_n;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // (whole warp runs the loop -- warp-uniform control flow keeps coordinates and descriptors in
    // uniform registers; one elected lane issues the TMA / tcgen05 instructions)
    {
      int stage = 0;
      uint32_t phase = 0;
      bool ok = true;
      for (int tile = blockIdx.x; tile < total_tiles && ok; tile += gridDim.x) {
        const int n_tile = tile % p.tiles_n, m_tile = tile / p.tiles_n;
        const int n0 = n_tile * p.BN;
        int w0 = 0, h0 = 0, i0 = 0;
        if (p.mode == 1) {
          w0 = (m_tile % p.tiles_w) * p.TW;
          h0 = ((m_tile / p.tiles_w) % p.tiles_h) * p.TH;
          i0 = (m_tile / (p.tiles_w * p.tiles_h)) * p.TN;
        }
        for (int kb = 0; kb < p.num_kb; ++kb) {
          ok = mbar_wait(empty_bar(stage), phase ^ 1u);
          if (!ok) break;
          const uint32_t sa = smem_base + (uint32_t)stage * p.stage_bytes;
          const uint32_t sb = sa + p.a_bytes;
          if (elect_one()) {
          mbar_arrive_expect_tx(full_bar(stage), (uint32_t)(p.a_bytes + p.b_bytes));
          if (p.mode == 0) {
            if (kb < p.kb1) {
              tma_load_2d(sa, &p.tmA[0], full_bar(stage), kb * p.BK, m_tile * kTileM);
              tma_load_2d(sb, &p.tmB, full_bar(stage), kb * p.BK, n0);
            } else {
              tma_load_2d(sa, &p.tmA[1], full_bar(stage), (kb - p.kb1) * p.BK, m_tile * kTileM);
              tma_load_2d(sb, &p.tmB, full_bar(stage), p.K1 + (kb - p.kb1) * p.BK, n0);
            }
          } else {
            const int tap = kb / p.kb_per_tap;
            const int c0 = (kb - tap * p.kb_per_tap) * p.BK;
            tma_load_4d(sa, &p.tmA[p.tap_map[tap]], full_bar(stage), c0, w0 + p.tap_dw[tap],
                        h0 + p.tap_dh[tap], i0);
            tma_load_2d(sb, &p.tmB, full_bar(stage), tap * p.Cin + c0, n0);
          }
          }
          __syncwarp();
          if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    {
      const uint32_t idesc = umma_idesc_bf16(kTileM, (uint32_t)p.BN);
      const int k_sub = p.BK / 16;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      bool ok = true;
      for (int tile = blockIdx.x; tile < total_tiles && ok; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (uint32_t)(it >> 1) & 1u;
        ok = mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        if (!ok) break;
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)acc * kMaxBN;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          ok = mbar_wait(full_bar(stage), phase);
          if (!ok) break;
          tc_fence_after();
          const uint32_t sa = smem_base + (uint32_t)stage * p.stage_bytes;
          const uint32_t sb = sa + p.a_bytes;
          const uint64_t adesc = umma_smem_desc(sa, (uint32_t)p.swz);
          const uint64_t bdesc = umma_smem_desc(sb, (uint32_t)p.swz);
          if (elect_one()) {
            for (int k = 0; k < k_sub; ++k) {
              // advance 16 bf16 = 32 B along K inside the swizzle atom: +2 in the (addr >> 4) field
              umma_f16(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                       (kb | k) != 0 ? 1u : 0u);
            }
            umma_commit(empty_bar(stage));
          }
          __syncwarp();
          if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
        }
        if (elect_one()) umma_commit(tfull_bar(acc));
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    // Two warps per TMEM lane quarter; each thread owns one output row and every other 32-column
    // pair of the tile.  The row's bf16 residual is fetched BEFORE the accumulator is waited for,
    // so its latency hides behind the main loop instead of serialising 32 bytes at a time.
    const int q = warp & 3;            // TMEM lane quarter this warp may access
    const int hf = (warp - 2) >> 2;    // which 32-column pairs: hf, hf + 2, ...
    const int r = q * 32 + lane;
    if (p.staged) {
      // ---- staged epilogue: TMEM -> registers -> swizzled smem slabs -> TMA store ----------------
      // Row-per-thread global stores touch 32 lines per warp instruction; staging the tile and letting
      // the TMA unit write full lines (and fetch the bf16 residual the same way) removes that bound.
      int it = 0;
      bool ok = true;
      const bool leader = (warp == 2);
      const uint32_t res_bytes = 128u * (uint32_t)p.bno * 2u;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int n_tile = tile % p.tiles_n, m_tile = tile / p.tiles_n;
        const int n0 = n_tile * p.BN;
        const int o0 = (p.act == SDB_ACT_GEGLU) ? (n0 >> 1) : n0;   // first output column of the tile
        const int acc = it & 1;
        const uint32_t acc_phase = (uint32_t)(it >> 1) & 1u;
        int img = 0, w0 = 0, h0 = 0, i0 = 0;
        if (p.mode == 1) {
          w0 = (m_tile % p.tiles_w) * p.TW;
          h0 = ((m_tile / p.tiles_w) % p.tiles_h) * p.TH;
          i0 = (m_tile / (p.tiles_w * p.tiles_h)) * p.TN;
          img = i0 + r / (p.TW * p.TH);
          if (img >= p.B) img = p.B - 1;   // clipped rows: any valid temb row will do
        }
        if (leader) {
          if (elect_one()) {
            tma_store_wait_read();          // the previous tile's stores have drained the staging slabs
            if (p.residual) {
              mbar_arrive_expect_tx(res_bar, res_bytes);
              for (int sl = 0; sl <= p.nslab; ++sl) {
                if (sl == p.nslab && p.tail_cols == 0) break;
                const void* tm = sl < p.nslab ? (const void*)&p.tmR : (const void*)&p.tmRt;
                const uint32_t dst = stg_base + (uint32_t)sl * 16384u;
                if (p.mode == 0) tma_load_2d(dst, tm, res_bar, o0 + sl * 64, m_tile * kTileM);
                else tma_load_4d(dst, tm, res_bar, o0 + sl * 64, w0, h0, i0);
              }
            }
          }
          __syncwarp();
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");   // staging is free (and the residual is on its way)
        if (ok) ok = mbar_wait(tfull_bar(acc), acc_phase);
        if (p.residual && ok) ok = mbar_wait(res_bar, (uint32_t)it & 1u);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * kMaxBN;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int c0 = (hf + 2 * i) * 32;
          if (c0 < p.BN) {
            uint32_t v0[16], v1[16];
            const bool two = (c0 + 16 < p.BN);
            tmem_ld_32x16(taddr + c0, v0);
            if (two) tmem_ld_32x16(taddr + c0 + 16, v1);
            tmem_ld_wait();
            // columns beyond N carry zero-filled weights; they are clipped by the TMA store
            const int cb = (n0 + c0 < p.N) ? n0 + c0 : 0, cb1 = (n0 + c0 + 16 < p.N) ? n0 + c0 + 16 : 0;
            epilogue16_staged(p, v0, stg_base, r, img, cb, c0);
            if (two) epilogue16_staged(p, v1, stg_base, r, img, cb1, c0 + 16);
          }
        }
        tc_fence_before();
        mbar_arrive(tempty_bar(acc));
        fence_proxy_async();                               // generic-proxy smem writes -> async proxy
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (leader) {
          if (elect_one() && ok) {
            for (int sl = 0; sl <= p.nslab; ++sl) {
              if (sl == p.nslab && p.tail_cols == 0) break;
              const void* tm = sl < p.nslab ? (const void*)&p.tmO : (const void*)&p.tmOt;
              const uint32_t src = stg_base + (uint32_t)sl * 16384u;
              if (p.mode == 0) tma_store_2d(tm, src, o0 + sl * 64, m_tile * kTileM);
              else tma_store_4d(tm, src, o0 + sl * 64, w0, h0, i0);
            }
            tma_store_commit();
          }
          __syncwarp();
        }
      }
      if (leader) {
        if (elect_one()) tma_store_wait_read();            // smem must outlive the last store's reads
        __syncwarp();
      }
    } else {
    const bool res_pre = p.residual != nullptr && p.res_dtype == SDB_BF16;
    int it = 0;
    bool ok = true;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int n_tile = tile % p.tiles_n, m_tile = tile / p.tiles_n;
      const int n0 = n_tile * p.BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (uint32_t)(it >> 1) & 1u;
      long long out_row = -1;
      int img = 0;
      if (p.mode == 0) {
        const long long m = (long long)m_tile * kTileM + r;
        if (m < p.M) out_row = m;
      } else {
        const int tw = r % p.TW, th = (r / p.TW) % p.TH, tn = r / (p.TW * p.TH);
        const int w = (m_tile % p.tiles_w) * p.TW + tw;
        const int h = ((m_tile / p.tiles_w) % p.tiles_h) * p.TH + th;
        img = (m_tile / (p.tiles_w * p.tiles_h)) * p.TN + tn;
        if (img < p.B && h < p.H && w < p.W) {
          if (p.out_mode == 0) out_row = ((long long)img * p.H + h) * p.W + w;
          else if (((h | w) & 1) == 0)
            out_row = ((long long)img * (p.H >> 1) + (h >> 1)) * (p.W >> 1) + (w >> 1);
        }
      }
      uint4 res[4][4];   // up to 4 pairs x 32 columns of bf16 residual
      if (res_pre && out_row >= 0) {
        const __nv_bfloat16* rrow = reinterpret_cast<const __nv_bfloat16*>(p.residual) + out_row * p.ldr + n0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int c0 = (hf + 2 * i) * 32;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (c0 + 8 * j < p.BN && n0 + c0 + 8 * j < p.N)
              res[i][j] = __ldg(reinterpret_cast<const uint4*>(rrow + c0 + 8 * j));
        }
      }
      if (ok) ok = mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * kMaxBN;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c0 = (hf + 2 * i) * 32;
        if (c0 < p.BN) {
          uint32_t v0[16], v1[16];
          const bool two = (c0 + 16 < p.BN);
          tmem_ld_32x16(taddr + c0, v0);
          if (two) tmem_ld_32x16(taddr + c0 + 16, v1);
          tmem_ld_wait();
          if (out_row >= 0 && ok) {
            const uint4(&r0)[2] = *reinterpret_cast<const uint4(*)[2]>(&res[i][0]);
            const uint4(&r1)[2] = *reinterpret_cast<const uint4(*)[2]>(&res[i][2]);
            if (res_pre) {
              if (n0 + c0 < p.N) epilogue16<true>(p, v0, out_row, img, n0 + c0, r0);
              if (two && n0 + c0 + 16 < p.N) epilogue16<true>(p, v1, out_row, img, n0 + c0 + 16, r1);
            } else {
              if (n0 + c0 < p.N) epilogue16<false>(p, v0, out_row, img, n0 + c0, r0);
              if (two && n0 + c0 + 16 < p.N) epilogue16<false>(p, v1, out_row, img, n0 + c0 + 16, r1);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(tempty_bar(acc));
    }
      }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int pick_bn(int N, int cap = kMaxBN) {
  // largest multiple of 16 (<= 256) that divides N; if only a narrow one exists, take a wide tile
  // and let the last N tile be partial (TMA zero-fills the missing weight rows, stores are masked)
  int best = 16;
  for (int bn = 16; bn <= cap; bn += 16)
    if (N % bn == 0) best = bn;
  if (best < 96 && N > best) best = N >= cap ? cap : ((N + 15) / 16) * 16;
  return best;
}

static void pick_bk(int K, int K2, int* BK, int* swz) {
  // 128-byte rows (BK = 64, SWIZZLE_128B) even when K is not a multiple of 64: the K tail of the last
  // block is zero-filled by the TMA unit (operands AND weights), which costs a few idle MMA columns
  // but halves the number of TMA row requests per byte compared with 64-byte rows.
  static int force32 = -1;
  if (force32 < 0) { const char* e = getenv("SDB_GEMM_BK32"); force32 = e ? atoi(e) : 0; }
  const bool d64 = (K % 64 == 0) && (K2 % 64 == 0);
  const bool d32 = (K % 32 == 0) && (K2 % 32 == 0);
  if (d64 || !d32 || !force32) { *BK = 64; *swz = 128; }
  else { *BK = 32; *swz = 64; }
}

// Decides whether the epilogue can be staged through smem + TMA stores (bf16 output at the tile's own
// resolution, bf16 or no residual, a tail slab the swizzle modes can express) and sizes the staging area.
static void plan_staging(UmmaParams& p, bool bf16_out, bool res_ok, bool same_res) {
  static int off = -1;
  if (off < 0) { const char* e = getenv("SDB_GEMM_DIRECT"); off = e ? atoi(e) : 0; }
  p.bno = p.act == SDB_ACT_GEGLU ? p.BN / 2 : p.BN;
  p.nslab = p.bno / 64;
  p.tail_cols = p.bno % 64;
  p.staged = (!off && bf16_out && res_ok && same_res &&
              (p.tail_cols == 0 || p.tail_cols == 16 || p.tail_cols == 32)) ? 1 : 0;
  p.staging_bytes = p.staged ? ((kTileM * p.bno * 2 + 1023) & ~1023) : 0;
}

static void finish_stages(UmmaParams& p) {
  p.a_bytes = kTileM * p.BK * 2;
  p.b_bytes = p.BN * p.BK * 2;
  p.stage_bytes = (p.a_bytes + p.b_bytes + 1023) & ~1023;
  int s = (kSmemBudget - p.staging_bytes) / p.stage_bytes;
  p.num_stages = s > kMaxStages ? kMaxStages : s;
}

// output / residual tensor maps of the staged epilogue: `rank` 2 (GEMM rows) or 4 (conv pixel boxes)
static int make_slab_maps(UmmaParams& p, const void* out, const void* res, int rank, const uint64_t* dims,
                          const uint64_t* so, const uint64_t* sr, const uint32_t* box_rows) {
  uint32_t box[4] = {64, box_rows[0], box_rows[1], box_rows[2]};
  int rc = 0;
  if (p.nslab > 0) {
    rc = sdb_make_tensor_map(&p.tmO, out, 2, rank, dims, so, box, 128);
    if (rc) return rc;
    if (res) rc = sdb_make_tensor_map(&p.tmR, res, 2, rank, dims, sr, box, 128);
    if (rc) return rc;
  }
  if (p.tail_cols > 0) {
    box[0] = (uint32_t)p.tail_cols;
    rc = sdb_make_tensor_map(&p.tmOt, out, 2, rank, dims, so, box, p.tail_cols * 2);
    if (rc) return rc;
    if (res) rc = sdb_make_tensor_map(&p.tmRt, res, 2, rank, dims, sr, box, p.tail_cols * 2);
  }
  return rc;
}

static int launch_umma(const UmmaParams& p, cudaStream_t st) {
  static bool attr_set[64] = {false};
  const int smem = kSmemBudget + 2048;
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    SDB_CUDA_OK(cudaFuncSetAttribute(umma_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     smem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  const int total = p.tiles_m * p.tiles_n;
  const int sms = sdb_num_sms();
  const int grid = total < sms ? total : sms;
  umma_gemm_kernel<<<grid, kGemmThreads, smem, st>>>(p);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

}  // namespace sdb

extern "C" int sdb_gemm(const SdbGemmArgs* a, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(a && a->a && a->w && a->out, "gemm: null pointer");
  SDB_CHECK_ARG(a->M > 0 && a->N > 0 && a->K1 > 0, "gemm: bad shape M=%d N=%d K1=%d", a->M, a->N, a->K1);
  SDB_CHECK_ARG(a->N % 16 == 0, "gemm: N (%d) must be a multiple of 16", a->N);
  const int K2 = a->a2 ? a->K2 : 0;
  const int K = a->K1 + K2;
  SDB_CHECK_ARG(a->K1 % 8 == 0 && K2 % 8 == 0 && a->lda % 8 == 0 && (!a->a2 || a->lda2 % 8 == 0),
                "gemm: K and lda must be multiples of 8 (16-byte TMA rows)");
  SDB_CHECK_ARG(a->ldo % 8 == 0 && (!a->residual || a->ldr % 8 == 0), "gemm: ldo/ldr must be multiples of 8");
  SDB_CHECK_ARG(((uintptr_t)a->a & 15) == 0 && ((uintptr_t)a->w & 15) == 0 && ((uintptr_t)a->out & 15) == 0 &&
                    ((uintptr_t)a->a2 & 15) == 0 && ((uintptr_t)a->residual & 15) == 0 &&
                    ((uintptr_t)a->bias & 15) == 0,
                "gemm: pointers must be 16-byte aligned");
  SDB_CHECK_ARG(a->act != SDB_ACT_GEGLU || (a->N % 32 == 0 && !a->residual), "gemm: GEGLU needs N %% 32 == 0 and no residual");

  UmmaParams p;
  memset(&p, 0, sizeof(p));
  p.mode = 0;
  pick_bk(a->K1, K2, &p.BK, &p.swz);
  // short-K GEMMs are bound by the output stream, not the tensor pipe: 128-wide tiles keep the
  // staging area small (deeper operand ring) at no cost
  p.BN = pick_bn(a->N, (K <= 320 && a->N > 256) ? 128 : kMaxBN);
  p.act = a->act;
  plan_staging(p, a->out_dtype == SDB_BF16, !a->residual || a->res_dtype == SDB_BF16, true);
  finish_stages(p);
  p.kb1 = (a->K1 + p.BK - 1) / p.BK;
  p.K1 = a->K1;
  p.num_kb = p.kb1 + (K2 + p.BK - 1) / p.BK;
  p.tiles_m = (a->M + kTileM - 1) / kTileM;
  p.tiles_n = (a->N + p.BN - 1) / p.BN;
  p.M = a->M;
  p.N = a->N;
  p.bias = a->bias;
  p.residual = a->residual;
  p.ldr = a->ldr;
  p.res_dtype = a->res_dtype;
  p.out = a->out;
  p.ldo = a->ldo;
  p.out_dtype = a->out_dtype;
  p.act = a->act;

  uint64_t dims[2], strides[2];
  uint32_t box[2];
  dims[0] = (uint64_t)a->K1; dims[1] = (uint64_t)a->M;
  strides[0] = 0; strides[1] = (uint64_t)a->lda * 2;
  box[0] = (uint32_t)p.BK; box[1] = kTileM;
  int rc = sdb_make_tensor_map(&p.tmA[0], a->a, 2, 2, dims, strides, box, p.swz);
  if (rc) return rc;
  if (a->a2 && K2 > 0) {
    dims[0] = (uint64_t)K2; strides[1] = (uint64_t)a->lda2 * 2;
    rc = sdb_make_tensor_map(&p.tmA[1], a->a2, 2, 2, dims, strides, box, p.swz);
    if (rc) return rc;
  }
  dims[0] = (uint64_t)K; dims[1] = (uint64_t)a->N;
  strides[1] = (uint64_t)K * 2;
  box[0] = (uint32_t)p.BK; box[1] = (uint32_t)p.BN;
  rc = sdb_make_tensor_map(&p.tmB, a->w, 2, 2, dims, strides, box, p.swz);
  if (rc) return rc;
  if (p.staged) {
    const int n_out = a->act == SDB_ACT_GEGLU ? a->N / 2 : a->N;
    uint64_t od[2] = {(uint64_t)n_out, (uint64_t)a->M};
    uint64_t so[2] = {0, (uint64_t)a->ldo * 2}, sr[2] = {0, (uint64_t)a->ldr * 2};
    uint32_t rows[3] = {kTileM, 1, 1};
    rc = make_slab_maps(p, a->out, a->residual, 2, od, so, sr, rows);
    if (rc) return rc;
  }
  return launch_umma(p, sdb_stream(stream));
}

extern "C" int sdb_conv3x3(const SdbConvArgs* a, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(a && a->x && a->w && a->out, "conv3x3: null pointer");
  SDB_CHECK_ARG(a->B > 0 && a->H > 0 && a->W > 0, "conv3x3: bad shape");
  SDB_CHECK_ARG(a->Cin % 8 == 0 && a->Cin >= 16, "conv3x3: Cin (%d) must be a multiple of 8, >= 16 (use sdb_conv3x3_in)", a->Cin);
  SDB_CHECK_ARG(a->Cout % 16 == 0, "conv3x3: Cout (%d) must be a multiple of 16 (use sdb_conv3x3_out)", a->Cout);
  SDB_CHECK_ARG(!a->up2, "conv3x3: fused up2 not built; call sdb_upsample2x first");
  SDB_CHECK_ARG(!a->down2 || (a->H % 2 == 0 && a->W % 2 == 0), "conv3x3: down2 needs even H, W");
  SDB_CHECK_ARG(((uintptr_t)a->x & 15) == 0 && ((uintptr_t)a->w & 15) == 0 && ((uintptr_t)a->out & 15) == 0 &&
                    ((uintptr_t)a->residual & 15) == 0 && ((uintptr_t)a->bias & 15) == 0 &&
                    ((uintptr_t)a->temb & 15) == 0 && a->ld_temb % 4 == 0,
                "conv3x3: pointers must be 16-byte aligned");

  UmmaParams p;
  memset(&p, 0, sizeof(p));
  p.mode = 1;
  pick_bk(a->Cin, 0, &p.BK, &p.swz);
  p.BN = pick_bn(a->Cout);
  p.act = SDB_ACT_NONE;
  plan_staging(p, a->out_dtype == SDB_BF16, !a->residual || a->res_dtype == SDB_BF16, !a->down2);
  finish_stages(p);
  p.Cin = a->Cin;
  p.kb_per_tap = (a->Cin + p.BK - 1) / p.BK;
  p.num_kb = 9 * p.kb_per_tap;
  // tile = TW x TH x TN pixels with TW*TH*TN = 128
  int tw = 1;
  while (tw < 128 && a->W % (tw * 2) == 0) tw *= 2;
  SDB_CHECK_ARG(tw >= 8 || tw == a->W, "conv3x3: W (%d) needs a power-of-two factor >= 8", a->W);
  int th = 1;
  while (tw * th < 128 && th < a->H) th *= 2;
  int tn = 128 / (tw * th);
  if (tw * th * tn != 128) {
    sdb_set_error("conv3x3: cannot tile W=%d H=%d into 128-pixel boxes", a->W, a->H);
    return SDB_EUNSUPPORTED;
  }
  p.TW = tw; p.TH = th; p.TN = tn;
  p.tiles_w = a->W / tw;
  p.tiles_h = (a->H + th - 1) / th;
  const int tiles_b = (a->B + tn - 1) / tn;
  p.tiles_m = p.tiles_w * p.tiles_h * tiles_b;
  p.tiles_n = (a->Cout + p.BN - 1) / p.BN;
  p.B = a->B; p.H = a->H; p.W = a->W;
  p.out_mode = a->down2 ? 1 : 0;
  for (int t = 0; t < 9; ++t) {
    p.tap_map[t] = 0;
    p.tap_dh[t] = (signed char)(t / 3 - 1);
    p.tap_dw[t] = (signed char)(t % 3 - 1);
  }
  p.M = 0;
  p.N = a->Cout;
  p.bias = a->bias;
  p.temb = a->temb;
  p.ld_temb = a->ld_temb;
  p.residual = a->residual;
  p.ldr = a->Cout;
  p.res_dtype = a->res_dtype;
  p.out = a->out;
  p.ldo = a->Cout;
  p.out_dtype = a->out_dtype;
  p.act = SDB_ACT_NONE;

  uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->B};
  uint64_t strides[4] = {0, (uint64_t)a->Cin * 2, (uint64_t)a->W * a->Cin * 2,
                         (uint64_t)a->H * a->W * a->Cin * 2};
  uint32_t box[4] = {(uint32_t)p.BK, (uint32_t)tw, (uint32_t)th, (uint32_t)tn};
  int rc = sdb_make_tensor_map(&p.tmA[0], a->x, 2, 4, dims, strides, box, p.swz);
  if (rc) return rc;
  uint64_t wd[2] = {(uint64_t)9 * a->Cin, (uint64_t)a->Cout};
  uint64_t ws[2] = {0, (uint64_t)9 * a->Cin * 2};
  uint32_t wb[2] = {(uint32_t)p.BK, (uint32_t)p.BN};
  rc = sdb_make_tensor_map(&p.tmB, a->w, 2, 2, wd, ws, wb, p.swz);
  if (rc) return rc;
  if (p.staged) {
    uint64_t od[4] = {(uint64_t)a->Cout, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->B};
    uint64_t so[4] = {0, (uint64_t)a->Cout * 2, (uint64_t)a->W * a->Cout * 2, (uint64_t)a->H * a->W * a->Cout * 2};
    uint32_t rows[3] = {(uint32_t)tw, (uint32_t)th, (uint32_t)tn};
    rc = make_slab_maps(p, a->out, a->residual, 4, od, so, so, rows);
    if (rc) return rc;
  }
  return launch_umma(p, sdb_stream(stream));
}
