// K1 / K3: persistent warp-specialised tcgen05 GEMM with TMA operand staging and TMEM
// accumulators.  One kernel serves
//   * GEMM  out[M,N] = A[M,K] W[N,K]^T           (A = one or two K segments, 2-D tensor maps)
//   * 3x3 convolution as implicit GEMM over NHWC  (A = 9 shifted 4-D boxes per channel chunk;
//     out-of-image taps are zero-filled by the TMA unit, which is exactly Conv2d(padding=1))
// Roles (384 threads):
//   warp 0      TMA producer (operand ring; in "resident-B" mode the weight tile is loaded once)
//   warp 1      MMA issuer (owns TMEM: two accumulator stages of up to 256 columns)
//   warp 2      epilogue DMA: bias (+temb) rows and the bf16 residual tile into the staging buffers
//               (TMA loads), finished tiles out of them (TMA stores)
//   warp 3      idle (keeps the epilogue warps on warpgroup boundaries)
//   warps 4-7   epilogue warpgroup 0: tiles 0, 2, 4, ... of this CTA (accumulator stage 0)
//   warps 8-11  epilogue warpgroup 1: tiles 1, 3, 5, ... (accumulator stage 1)
// Every hand-off is an mbarrier; the two epilogue warpgroups ping-pong so that the epilogue of
// tile i, the main loop of tile i+1 and the store / residual fetch of tiles i-1 / i+1 overlap.
#include <stdlib.h>

#include "common.cuh"
#include "umma_helpers.cuh"

namespace sdb {

constexpr int kTileM = 128;
constexpr int kMaxBN = 256;
constexpr int kGemmThreads = 384;
constexpr int kMaxStages = 12;
constexpr int kSmemBudget = 222 * 1024;  // resident B + operand ring + epilogue staging
constexpr int kMaxBuf = 3;
constexpr int kBiasBytes = kMaxBuf * kMaxBN * 4;

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
  // resident-B scheduling (GEMM with a short K): each CTA keeps one n-tile of the weights in smem for
  // its whole life and streams only A; CTA c owns n-tile c % tiles_n and m-tiles c / tiles_n + j * m_groups
  int resB, bres_bytes, m_groups;
  int walk_dm, walk_dn;   // TileWalk increments (host-computed from the grid size)
  // CTA-pair mode (conv, cta_group::2): the two CTAs of a cluster own the m-tiles 2k and 2k + 1 of ONE n-tile and run
  // them as a single M = 256 MMA issued by the leader; each stages its own A tile and HALF of the weight tile, so a
  // weight byte crosses L2 -> SM once per 256 output rows (the kernel is bound by that path: 90 % of the LTS cap).
  // Tiles are enumerated n-major over tiles_m padded to even; every CTA runs `tiles_per_cta` iterations, ghosts
  // included (zero-filled loads, clipped stores).
  int pair, tiles_m_pad, tiles_per_cta;
  // halo loader (conv, W % 128 == 0: a tile = 128 pixels of ONE image row): instead of one shifted 128 x BK box per tap,
  // ONE (rows x halo_w)-pixel block per channel block is fetched and every tap's A operand is a view into it -- a UMMA
  // descriptor may start at any 128-byte row of a SWIZZLE_128B tile (the swizzle is a function of the address bits;
  // tools/ubench_halo.cu) -- so an activation byte crosses L2 -> SM 3.05x instead of 9x per output tile.  Two rings:
  // stages [0, halo_nA) hold halo blocks (warp 3 loads them), stages [halo_nA, halo_nA + halo_nB) weight tiles (warp 0).
  int halo, halo_w, halo_nA, halo_nB, halo_stage_bytes, halo_box_bytes, halo_dw0, halo_dh0;
  short tap_row[12];  // first halo-block row of the tap's view (padded: the issue loop reads two ahead)
  int ring_bytes;     // operand ring(s) in shared memory
  int halo_tg;        // halo mode: taps per weight-ring stage (1, 2 or 3; divides the tap count)
  // conv geometry (GEMM-M space = input-resolution pixels)
  int kb_per_tap, Cin;
  int TW, TH, TN, tiles_w, tiles_h;
  int B, H, W;
  int out_mode;  // 0 = same resolution, 1 = keep even rows/cols (conv then nearest x0.5),
                 // 2 = phase (ph_a, ph_b) of a nearest-x2 upsampled output: pixel (h, w) -> (2h + ph_a, 2w + ph_b)
  int ph_a, ph_b;
  signed char tap_map[9], tap_dh[9], tap_dw[9];
  // epilogue
  int M, N;
  const float* bias;
  const float* temb;
  long long ld_temb;
  int temb_in_bias;  // tile lies inside one image: temb row folded into the staged bias row
  const void* residual;
  long long ldr;
  int res_dtype;
  void* out;
  long long ldo;
  int out_dtype;
  int u8_channels;   // SDB_U8 output: channels kept per pixel
  int act;
  // LayerNorm over the output row in the epilogue (kEpiResLN: one n-tile holds whole rows)
  const float* ln_gamma;
  const float* ln_beta;
  float ln_eps;
  // staged epilogue (bf16 output through swizzled smem slabs + TMA store): 0 = direct stores
  int staged;
  int nbuf;         // staging buffers, tile j uses buffer j % nbuf (3 = a store never delays the next tile,
                    // 2 = one per epilogue warpgroup, 1 = shared)
  int bno;          // output columns per tile (BN, or BN/2 for GEGLU)
  int nslab;        // full 64-column slabs per tile
  int tail_cols;    // columns of the last, narrower slab (0, 16 or 32)
  int stg_bytes;    // bytes of one staging buffer
  // per-channel sums of the finished (bf16-rounded) output for the GroupNorm that consumes it: the epilogue warpgroup
  // re-reads its staged tile column-wise and adds sum / sum of squares into chsum[img][N][2] (64-bit fixed-point
  // atomics, flushed only when the CTA moves to another image or n-tile).  Needs the staged epilogue and tiles that lie
  // inside one image.
  unsigned long long* chsum;
  int st_tiles_img;   // m-tiles per image
  int st_splits;      // row ranges a column pair is split over (threads = bno / 2 * st_splits <= 128)
  int st_rows;        // rows per range = 128 / st_splits
};

// ---- packed fp32 pair helpers (one issue slot for two elements) ---------------------------------
// (fmul2 / ffma2 / fadd2 live in attention_umma.cu, included first in the unity build)

// a * gelu(g) for two (a, g) pairs, exact-erf GELU (transformer.py:177 uses nn.GELU()).
//   erfc(u/sqrt2) = 2^(-u P(u)),  u = min(|g|, 6),  P = degree-5 fit of -log2(erfc(u/sqrt2))/u on [0, 6]
// Absolute error of Phi(g) <= 3e-6 and -- because the error sits in the exponent -- relative error of the
// negative-side tail <= 1.5e-4, an order below the bf16 rounding of the result.  One MUFU (ex2) per element;
// the polynomial runs in nu = -u (odd coefficients negated) so no negation is needed.
__device__ __forceinline__ void geglu2(float a0, float a1, float g0, float g1, float& o0, float& o1) {
  const float n0 = fmaxf(-fabsf(g0), -6.0f), n1 = fmaxf(-fabsf(g1), -6.0f);
  float p0, p1;
  ffma2(p0, p1, n0, n1, 2.7593860068009235e-05f, 2.7593860068009235e-05f, 0.0006911802338436246f,
        0.0006911802338436246f);
  ffma2(p0, p1, p0, p1, n0, n1, 0.00771324010565877f, 0.00771324010565877f);
  ffma2(p0, p1, p0, p1, n0, n1, 0.05270770937204361f, 0.05270770937204361f);
  ffma2(p0, p1, p0, p1, n0, n1, -0.4592883586883545f, -0.4592883586883545f);
  ffma2(p0, p1, p0, p1, n0, n1, 1.1510891914367676f, 1.1510891914367676f);
  fmul2(p0, p1, p0, p1, n0, n1);
  const float e0 = ex2_mufu(p0), e1 = ex2_mufu(p1);
  // 0.5 g (1 + erf(g/sqrt2)) = h + h * copysign(1 - e, g),  h = g / 2
  const float s0 = copysignf(1.0f - e0, g0), s1 = copysignf(1.0f - e1, g1);
  float h0, h1;
  fmul2(h0, h1, g0, g1, 0.5f, 0.5f);
  ffma2(h0, h1, h0, h1, s0, s1, h0, h1);
  fmul2(o0, o1, h0, h1, a0, a1);
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
// all but the most recent committed store have finished reading shared memory
__device__ __forceinline__ void tma_store_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float4 lds128f(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts32f(uint32_t a, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
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

// acc[0..16) += bias row (fp32 in shared memory; every lane reads the same address: one wavefront)
__device__ __forceinline__ void add_bias16(float (&v)[16], const uint32_t (&acc)[16], uint32_t bias_s) {
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    const float4 b = lds128f(bias_s + 4u * j);
    fadd2(v[j], v[j + 1], __uint_as_float(acc[j]), __uint_as_float(acc[j + 1]), b.x, b.y);
    fadd2(v[j + 2], v[j + 3], __uint_as_float(acc[j + 2]), __uint_as_float(acc[j + 3]), b.z, b.w);
  }
}

__device__ __forceinline__ void add_bf16x8(float* v, uint4 q) {
  const float2 x = unpack_bf16x2(q.x), y = unpack_bf16x2(q.y), z = unpack_bf16x2(q.z), w = unpack_bf16x2(q.w);
  fadd2(v[0], v[1], v[0], v[1], x.x, x.y);
  fadd2(v[2], v[3], v[2], v[3], y.x, y.y);
  fadd2(v[4], v[5], v[4], v[5], z.x, z.y);
  fadd2(v[6], v[7], v[6], v[7], w.x, w.y);
}
__device__ __forceinline__ uint4 pack_bf16x8(const float* v) {
  return make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                    pack_bf16x2(v[6], v[7]));
}

// Epilogue flavours the kernel is instantiated for: the common ones are straight-line code (two warps
// per scheduler cannot hide fetch bubbles of a branchy epilogue), the generic one branches at run time.
// kEpiResLN = kEpiRes followed by LayerNorm of the finished row (post-norm transformer block, transformer.py:274-278).
enum { kEpiPlain = 0, kEpiRes = 1, kEpiGeglu = 2, kEpiGeneric = 3, kEpiResLN = 4 };

// ---- 16 accumulator columns of one row -> staged tile ------------------------------------------
// `tcol` = accumulator column inside the tile, `temb_row` = per-element temb (generic flavour only, when
// it could not be folded into the staged bias row).  The residual is read from (and the result written
// back to) the swizzled slab the TMA unit loaded / will store.
template <int kEpi>
__device__ __forceinline__ void epilogue16_staged(const UmmaParams& p, const uint32_t (&acc)[16], uint32_t stg,
                                                  uint32_t bias_s, const float* temb_row, int r, int tcol,
                                                  float& ln_s, float& ln_ss) {
  float v[16];
  add_bias16(v, acc, bias_s + 4u * tcol);
  if (kEpi == kEpiGeglu) {
    // pair-interleaved columns (a0, a1, g0, g1, ...): 16 accumulator columns -> 8 outputs, and every
    // operand of the packed arithmetic is an adjacent register pair
    float o[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) geglu2(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3], o[2 * k], o[2 * k + 1]);
    sts128(stage_addr(p, stg, r, tcol >> 1), pack_bf16x8(o));
    return;
  }
  if (kEpi == kEpiGeneric) {
    if (temb_row) {
#pragma unroll
      for (int j = 0; j < 16; j += 4) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(temb_row + tcol + j));
        v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
      }
    }
    if (p.act == SDB_ACT_SILU) {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = silu_f(v[j]);
    }
  }
  const uint32_t a = stage_addr(p, stg, r, tcol);   // 16 columns never straddle a slab: chunks c, c ^ 1
  const uint32_t a1 = a ^ 16u;
  if (kEpi == kEpiRes || kEpi == kEpiResLN || (kEpi == kEpiGeneric && p.residual)) {
    add_bf16x8(v, lds128(a));
    add_bf16x8(v + 8, lds128(a1));
  }
  if (kEpi == kEpiResLN) {
    // row statistics from the fp32 values, the row itself parked as bf16 in the staging slab
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      ln_s += v[j];
      ln_ss = fmaf(v[j], v[j], ln_ss);
    }
  }
  sts128(a, pack_bf16x8(v));
  sts128(a1, pack_bf16x8(v + 8));
}

// second pass of kEpiResLN over 16 parked columns of row r: y = (h - mean) * rstd * gamma + beta
__device__ __forceinline__ void layernorm16_staged(const UmmaParams& p, uint32_t stg, int r, int tcol, float mean,
                                                   float rstd) {
  const uint32_t a = stage_addr(p, stg, r, tcol);
  const uint32_t a1 = a ^ 16u;
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = 0.f;
  add_bf16x8(v, lds128(a));
  add_bf16x8(v + 8, lds128(a1));
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    const float4 g = __ldg(reinterpret_cast<const float4*>(p.ln_gamma + tcol + j));
    const float4 b = __ldg(reinterpret_cast<const float4*>(p.ln_beta + tcol + j));
    v[j] = fmaf((v[j] - mean) * rstd, g.x, b.x);
    v[j + 1] = fmaf((v[j + 1] - mean) * rstd, g.y, b.y);
    v[j + 2] = fmaf((v[j + 2] - mean) * rstd, g.z, b.z);
    v[j + 3] = fmaf((v[j + 3] - mean) * rstd, g.w, b.w);
  }
  sts128(a, pack_bf16x8(v));
  sts128(a1, pack_bf16x8(v + 8));
}

// ---- 16 accumulator columns of one row -> global memory (fp32 output / fp32 residual / strided rows) ----
__device__ __forceinline__ void epilogue16_direct(const UmmaParams& p, const uint32_t (&acc)[16], uint32_t bias_s,
                                                  const float* temb_row, long long out_row, int n0, int tcol) {
  float v[16];
  add_bias16(v, acc, bias_s + 4u * tcol);
  const int col = n0 + tcol;
  if (p.out_dtype == SDB_U8) {
    // image head: utils/prepare_dataset.py:145-150 -- (x + 1) / 2, clip to [0, 1], * 255, truncate -- on the fp32
    // accumulator, same arithmetic as to_uint8_kernel; uint8 NHWC with only the real channels
    uint8_t* d = reinterpret_cast<uint8_t*>(p.out) + out_row * p.u8_channels;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (col + j < p.u8_channels) {
        float y = (v[j] + 1.0f) / 2.0f;
        y = fminf(fmaxf(y, 0.0f), 1.0f);
        d[col + j] = (uint8_t)(y * 255.0f);
      }
    }
    return;
  }
  if (temb_row) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(temb_row + tcol + j));
      v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
    }
  }
  if (p.act == SDB_ACT_GEGLU) {
    float o[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) geglu2(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3], o[2 * k], o[2 * k + 1]);
    const long long off = out_row * p.ldo + (col >> 1);
    if (p.out_dtype == SDB_BF16) {
      *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off) = pack_bf16x8(o);
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
    if (p.res_dtype == SDB_BF16) {
      const uint4* rr = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.residual) +
                                                       out_row * p.ldr + col);
      add_bf16x8(v, __ldg(rr));
      add_bf16x8(v + 8, __ldg(rr + 1));
    } else {
      const float4* rr = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.residual) +
                                                         out_row * p.ldr + col);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float4 q = __ldg(rr + j);
        v[4 * j] += q.x; v[4 * j + 1] += q.y; v[4 * j + 2] += q.z; v[4 * j + 3] += q.w;
      }
    }
  }
  const long long off = out_row * p.ldo + col;
  if (p.out_dtype == SDB_BF16) {
    uint4* d = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off);
    d[0] = pack_bf16x8(v);
    d[1] = pack_bf16x8(v + 8);
  } else {
    float4* d = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + off);
#pragma unroll
    for (int j = 0; j < 4; ++j) d[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  }
}

// Optional pipeline trace (tools/trace_gemm.py builds a private copy of the library with -DSDB_GEMM_TRACE):
// clock64 stamps of CTA 0's roles for its first 16 tiles.
#ifdef SDB_GEMM_TRACE
__device__ long long g_gemm_trace[5 * 16 * 8];
__device__ long long g_gemm_cta_ns[160 * 4];   // per CTA: globaltimer at start / end, clock64 at start / end
#define GEMM_TRACE(role, j, ev)                                                                   \
  do {                                                                                            \
    if (blockIdx.x == 0 && (j) < 16 && (threadIdx.x & 31) == 0 && ((role) < 3 || (threadIdx.x & 127) == 0)) \
      g_gemm_trace[(((role) * 16 + (j)) * 8) + (ev)] = clock64();                                 \
  } while (0)
#else
#define GEMM_TRACE(role, j, ev) do { } while (0)
#endif

// The tiles of this CTA, in order, without per-tile divisions.  Resident-B: fixed n-tile, m-tiles m_groups
// apart; otherwise tile t = blockIdx.x + j * gridDim.x of the row-major (m, n) enumeration.
struct TileWalk {
  int m, n, dm, dn, tiles_m, tiles_n;
  bool pr;                      // CTA-pair mode
  int left, t, pad, stride;     // iterations left, linear tile index, padded m-tile count, grid size
  __device__ __forceinline__ explicit TileWalk(const UmmaParams& p) {
    tiles_m = p.tiles_m;
    tiles_n = p.tiles_n;
    dm = p.walk_dm;
    dn = p.walk_dn;
    pr = p.pair != 0;
    left = t = pad = stride = 0;
    if (pr) {
      left = p.tiles_per_cta;
      t = (int)blockIdx.x;
      pad = p.tiles_m_pad;
      stride = (int)gridDim.x;
      n = t / pad;
      m = t - n * pad;
    } else {
      n = (int)blockIdx.x % tiles_n;
      m = (int)blockIdx.x / tiles_n;
    }
  }
  // pair mode: true for ghost tiles too (m >= tiles_m or n >= tiles_n) -- both CTAs of a pair keep walking together
  __device__ __forceinline__ bool valid() const { return pr ? left > 0 : m < tiles_m; }
  __device__ __forceinline__ void next() {
    if (pr) {
      --left;
      t += stride;
      n = t / pad;
      m = t - n * pad;
      return;
    }
    m += dm;
    n += dn;
    if (n >= tiles_n) { n -= tiles_n; ++m; }
  }
};

template <int kEpi, bool k2 = false>
__global__ void __launch_bounds__(kGemmThreads, 1)
umma_gemm_kernel(const __grid_constant__ UmmaParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-B aligned: [resident B][operand ring][staging buffers][bias rows][barriers]
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring_base = smem_base + (uint32_t)p.bres_bytes;
  const uint32_t stg_base = ring_base + (uint32_t)p.ring_bytes;
  const uint32_t bias_base = stg_base + (uint32_t)p.nbuf * (uint32_t)p.stg_bytes;
  const uint32_t bar_base = bias_base + (uint32_t)kBiasBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kMaxStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kMaxStages + 2 + a); };
  auto res_full_bar = [&](int b) { return bar_base + 8u * (2 * kMaxStages + 4 + b); };   // bias row (+ residual) landed
  auto stg_full_bar = [&](int b) { return bar_base + 8u * (2 * kMaxStages + 6 + b); };   // tile staged / bias row consumed
  const uint32_t bfull_bar = bar_base + 8u * (2 * kMaxStages + 8);
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 9);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
#ifdef SDB_GEMM_TRACE
  if (threadIdx.x == 0 && blockIdx.x < 160) {
    g_gemm_cta_ns[blockIdx.x * 4] = (long long)globaltimer_ns();
    g_gemm_cta_ns[blockIdx.x * 4 + 2] = clock64();
  }
#endif

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmA[0]);
    prefetch_tmap(&p.tmB);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), k2 ? 256 : 128);   // pair: the leader's MMA warp waits for both CTAs' epilogues
      mbar_init(res_full_bar(a), 1);
      mbar_init(stg_full_bar(a), 128);
    }
    mbar_init(bfull_bar, 1);
    fence_barrier_init();
  }
  if (k2) sdb_cluster_sync();   // the peer's barriers exist before anything is signalled at them
  if (warp == 1) {
    if (k2) {
      tmem_alloc_2cta(tmem_slot, 512);
      tmem_relinquish_2cta();
    } else {
      tmem_alloc(tmem_slot, 512);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  const uint32_t pair_rank = k2 ? sdb_cluster_ctarank() : 0u;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // (whole warp runs the loop -- warp-uniform control flow keeps coordinates and descriptors in
    // uniform registers; one elected lane issues the TMA instructions)
    int stage = 0;
    uint32_t phase = 0;
    bool ok = true;
    TileWalk tw(p);
    if (p.resB && tw.valid()) {
      if (elect_one()) {
        mbar_arrive_expect_tx(bfull_bar, (uint32_t)p.bres_bytes);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          const int kc = kb < p.kb1 ? kb * p.BK : p.K1 + (kb - p.kb1) * p.BK;
          tma_load_2d(smem_base + (uint32_t)kb * p.b_bytes, &p.tmB, bfull_bar, kc, tw.n * p.BN);
        }
      }
      __syncwarp();
    }
    if (p.halo) {
      // halo mode: this warp streams the weight tiles, `halo_tg` taps of one channel block per load (the weights seen as
      // [Cin][Cout][tap]: one 3-D box lands as halo_tg consecutive K-major tiles) in the MMA warp's order
      const uint32_t sbytes = (uint32_t)p.halo_tg * (uint32_t)p.b_bytes;
      const uint32_t txb = (k2 ? 2u : 1u) * sbytes;
      const uint32_t bring = ring_base + (uint32_t)p.halo_nA * (uint32_t)p.halo_stage_bytes;
      const int n_tap = p.num_kb / p.kb_per_tap;
      for (; tw.valid() && ok; tw.next()) {
        const int n0 = tw.n * p.BN;
        for (int kb = 0; kb < p.kb_per_tap && ok; ++kb) {
          for (int tap = 0; tap < n_tap; tap += p.halo_tg) {
            ok = mbar_wait(empty_bar(p.halo_nA + stage), phase ^ 1u);
            if (!ok) break;
            const uint32_t sb = bring + (uint32_t)stage * sbytes;
            if (elect_one()) {
              if (k2) {
                const uint32_t lead_full = sdb_mapa(full_bar(p.halo_nA + stage), 0u);
                if (pair_rank == 0) mbar_arrive_expect_tx(full_bar(p.halo_nA + stage), txb);
                tma_load_3d_2cta(sb, &p.tmB, lead_full, kb * p.BK, n0 + (int)pair_rank * (p.BN >> 1), tap);
              } else {
                mbar_arrive_expect_tx(full_bar(p.halo_nA + stage), txb);
                tma_load_3d(sb, &p.tmB, full_bar(p.halo_nA + stage), kb * p.BK, n0, tap);
              }
            }
            __syncwarp();
            if (++stage == p.halo_nB) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
    // pair: the leader's barrier collects both CTAs' A tile and weight half
    const uint32_t tx = k2 ? 2u * (uint32_t)(p.a_bytes + p.b_bytes)
                           : (uint32_t)(p.resB ? p.a_bytes : p.a_bytes + p.b_bytes);
    for (int j = 0; !p.halo && tw.valid() && ok; tw.next(), ++j) {
      GEMM_TRACE(0, j, 0);
      const int m_tile = tw.m;
      const int n0 = tw.n * p.BN;
      int w0 = 0, h0 = 0, i0 = 0;
      if (p.mode == 1) {
        w0 = (m_tile % p.tiles_w) * p.TW;
        h0 = ((m_tile / p.tiles_w) % p.tiles_h) * p.TH;
        i0 = (m_tile / (p.tiles_w * p.tiles_h)) * p.TN;
      }
      if (p.mode == 1) {
        // conv: this warp fetches the activation tiles (and posts the stage's byte count), warp 3 the weight tiles --
        // one producer thread's wait / post / two-TMA round trip per stage was as long as the stage's four MMAs
        const int n_tap = p.num_kb / p.kb_per_tap;
        for (int tap = 0; tap < n_tap && ok; ++tap) {
          const int wt = w0 + p.tap_dw[tap], ht = h0 + p.tap_dh[tap];
          for (int c0 = 0; c0 < p.kb_per_tap * p.BK; c0 += p.BK) {
            ok = mbar_wait(empty_bar(stage), phase ^ 1u);
            if (!ok) break;
            const uint32_t sa = ring_base + (uint32_t)stage * p.stage_bytes;
            if (elect_one()) {
              if (k2) {
                if (pair_rank == 0) mbar_arrive_expect_tx(full_bar(stage), tx);
                tma_load_4d_2cta(sa, &p.tmA[0], sdb_mapa(full_bar(stage), 0u), c0, wt, ht, i0);
              } else {
                mbar_arrive_expect_tx(full_bar(stage), tx);
                tma_load_4d(sa, &p.tmA[0], full_bar(stage), c0, wt, ht, i0);
              }
            }
            __syncwarp();
            if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
          }
        }
        GEMM_TRACE(0, j, 1);
        continue;
      }
      for (int kb = 0; kb < p.num_kb; ++kb) {
        ok = mbar_wait(empty_bar(stage), phase ^ 1u);
        if (!ok) break;
        const uint32_t sa = ring_base + (uint32_t)stage * p.stage_bytes;
        const uint32_t sb = sa + p.a_bytes;
        if (elect_one()) {
          mbar_arrive_expect_tx(full_bar(stage), tx);
          if (kb < p.kb1) {
            tma_load_2d(sa, &p.tmA[0], full_bar(stage), kb * p.BK, m_tile * kTileM);
            if (!p.resB) tma_load_2d(sb, &p.tmB, full_bar(stage), kb * p.BK, n0);
          } else {
            tma_load_2d(sa, &p.tmA[1], full_bar(stage), (kb - p.kb1) * p.BK, m_tile * kTileM);
            if (!p.resB) tma_load_2d(sb, &p.tmB, full_bar(stage), p.K1 + (kb - p.kb1) * p.BK, n0);
          }
        }
        __syncwarp();
        if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
      }
      GEMM_TRACE(0, j, 1);
    }
  } else if (warp == 1 && (!k2 || pair_rank == 0)) {
    // ===================== MMA issuer (pair mode: the leader CTA only) =====================
    // Descriptors are kept as (constant high word | 14-bit address field) and advanced by adds: the loop
    // body is a barrier wait, four MMAs and a commit.
    const uint32_t idesc = umma_idesc_bf16(k2 ? 2 * kTileM : kTileM, (uint32_t)p.BN);
    const uint64_t desc_hi = umma_smem_desc(0u, (uint32_t)p.swz);
    const uint32_t a_step = (uint32_t)p.stage_bytes >> 4;
    const uint32_t a_lo0 = (ring_base & 0x3ffffu) >> 4;
    const uint32_t b_step = p.resB ? ((uint32_t)p.b_bytes >> 4) : 0u;
    const uint32_t b_lo0 = p.resB ? ((smem_base & 0x3ffffu) >> 4) : a_lo0 + ((uint32_t)p.a_bytes >> 4);
    const bool bk64 = p.BK == 64;
    int stage = 0, hb_stage = 0;
    uint32_t phase = 0, hb_phase = 0;
    uint32_t a_lo = a_lo0;
    bool ok = true;
    TileWalk tw(p);
    if (p.resB && tw.valid()) ok = mbar_wait(bfull_bar, 0);
    for (int j = 0; tw.valid() && ok; tw.next(), ++j) {
      const int acc = j & 1;
      const uint32_t acc_phase = (uint32_t)(j >> 1) & 1u;
      GEMM_TRACE(1, j, 0);
      ok = mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
      if (!ok) break;
      GEMM_TRACE(1, j, 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)acc * kMaxBN;
      uint32_t b_lo = b_lo0;
      if (p.halo) {
        // one halo block per channel block, its taps as row-shifted views; `stage` / `phase` walk the halo ring,
        // hb_stage / hb_phase the weight ring, whose stages hold `halo_tg` taps: one barrier wait, election and commit
        // per tap group -- that round trip is of the order of 400 cycles for the one issuing thread, longer than the
        // four MMAs of a single tap on tiles of <= 128 columns
        const int n_tap = p.num_kb / p.kb_per_tap;
        const int tg = p.halo_tg;
        const uint32_t sbytes = (uint32_t)tg * (uint32_t)p.b_bytes;
        const uint32_t bring = ring_base + (uint32_t)p.halo_nA * (uint32_t)p.halo_stage_bytes;
        const uint32_t bstep = (uint32_t)p.b_bytes >> 4;
        for (int kb = 0; kb < p.kb_per_tap && ok; ++kb) {
          ok = mbar_wait(full_bar(stage), phase);
          if (!ok) break;
          const uint32_t sa = ring_base + (uint32_t)stage * (uint32_t)p.halo_stage_bytes;
          for (int tap = 0; tap < n_tap; tap += tg) {
            ok = mbar_wait(full_bar(p.halo_nA + hb_stage), hb_phase);
            if (!ok) break;
            tc_fence_after();
            const uint64_t ad0 = desc_hi | (uint64_t)(((sa + (uint32_t)p.tap_row[tap] * 128u) & 0x3ffffu) >> 4);
            const uint64_t ad1 = desc_hi | (uint64_t)(((sa + (uint32_t)p.tap_row[tap + 1] * 128u) & 0x3ffffu) >> 4);
            const uint64_t ad2 = desc_hi | (uint64_t)(((sa + (uint32_t)p.tap_row[tap + 2] * 128u) & 0x3ffffu) >> 4);
            const uint64_t bd0 = desc_hi | (uint64_t)(((bring + (uint32_t)hb_stage * sbytes) & 0x3ffffu) >> 4);
            if (elect_one()) {
              const uint32_t first = (kb | tap) != 0 ? 1u : 0u;
              if (k2) {
                umma_f16_2cta(tmem_d, ad0, bd0, idesc, first);
                umma_f16_2cta(tmem_d, ad0 + 2, bd0 + 2, idesc, 1u);
                umma_f16_2cta(tmem_d, ad0 + 4, bd0 + 4, idesc, 1u);
                umma_f16_2cta(tmem_d, ad0 + 6, bd0 + 6, idesc, 1u);
                if (tg > 1) {
                  const uint64_t bd1 = bd0 + bstep;
                  umma_f16_2cta(tmem_d, ad1, bd1, idesc, 1u);
                  umma_f16_2cta(tmem_d, ad1 + 2, bd1 + 2, idesc, 1u);
                  umma_f16_2cta(tmem_d, ad1 + 4, bd1 + 4, idesc, 1u);
                  umma_f16_2cta(tmem_d, ad1 + 6, bd1 + 6, idesc, 1u);
                }
                if (tg > 2) {
                  const uint64_t bd2 = bd0 + 2u * bstep;
                  umma_f16_2cta(tmem_d, ad2, bd2, idesc, 1u);
                  umma_f16_2cta(tmem_d, ad2 + 2, bd2 + 2, idesc, 1u);
                  umma_f16_2cta(tmem_d, ad2 + 4, bd2 + 4, idesc, 1u);
                  umma_f16_2cta(tmem_d, ad2 + 6, bd2 + 6, idesc, 1u);
                }
                umma_commit_2cta(empty_bar(p.halo_nA + hb_stage));
                if (tap + tg >= n_tap) umma_commit_2cta(empty_bar(stage));
              } else {
                umma_f16(tmem_d, ad0, bd0, idesc, first);
                umma_f16(tmem_d, ad0 + 2, bd0 + 2, idesc, 1u);
                umma_f16(tmem_d, ad0 + 4, bd0 + 4, idesc, 1u);
                umma_f16(tmem_d, ad0 + 6, bd0 + 6, idesc, 1u);
                if (tg > 1) {
                  const uint64_t bd1 = bd0 + bstep;
                  umma_f16(tmem_d, ad1, bd1, idesc, 1u);
                  umma_f16(tmem_d, ad1 + 2, bd1 + 2, idesc, 1u);
                  umma_f16(tmem_d, ad1 + 4, bd1 + 4, idesc, 1u);
                  umma_f16(tmem_d, ad1 + 6, bd1 + 6, idesc, 1u);
                }
                if (tg > 2) {
                  const uint64_t bd2 = bd0 + 2u * bstep;
                  umma_f16(tmem_d, ad2, bd2, idesc, 1u);
                  umma_f16(tmem_d, ad2 + 2, bd2 + 2, idesc, 1u);
                  umma_f16(tmem_d, ad2 + 4, bd2 + 4, idesc, 1u);
                  umma_f16(tmem_d, ad2 + 6, bd2 + 6, idesc, 1u);
                }
                umma_commit(empty_bar(p.halo_nA + hb_stage));
                if (tap + tg >= n_tap) umma_commit(empty_bar(stage));
              }
            }
            __syncwarp();
            if (++hb_stage == p.halo_nB) { hb_stage = 0; hb_phase ^= 1u; }
          }
          if (++stage == p.halo_nA) { stage = 0; phase ^= 1u; }
        }
      }
      for (int kb = 0; !p.halo && kb < p.num_kb; ++kb) {
        ok = mbar_wait(full_bar(stage), phase);
        if (!ok) break;
        tc_fence_after();
        const uint64_t adesc = desc_hi | (uint64_t)a_lo;
        const uint64_t bdesc = desc_hi | (uint64_t)(p.resB ? b_lo : a_lo + ((uint32_t)p.a_bytes >> 4));
        if (elect_one()) {
          // advance 16 bf16 = 32 B along K inside the swizzle atom: +2 in the (addr >> 4) field
          if (k2) {
            umma_f16_2cta(tmem_d, adesc, bdesc, idesc, kb != 0 ? 1u : 0u);
            umma_f16_2cta(tmem_d, adesc + 2, bdesc + 2, idesc, 1u);
            umma_f16_2cta(tmem_d, adesc + 4, bdesc + 4, idesc, 1u);
            umma_f16_2cta(tmem_d, adesc + 6, bdesc + 6, idesc, 1u);
            umma_commit_2cta(empty_bar(stage));     // frees the slot in both CTAs
          } else {
            umma_f16(tmem_d, adesc, bdesc, idesc, kb != 0 ? 1u : 0u);
            umma_f16(tmem_d, adesc + 2, bdesc + 2, idesc, 1u);
            if (bk64) {
              umma_f16(tmem_d, adesc + 4, bdesc + 4, idesc, 1u);
              umma_f16(tmem_d, adesc + 6, bdesc + 6, idesc, 1u);
            }
            umma_commit(empty_bar(stage));
          }
        }
        __syncwarp();
        b_lo += b_step;
        a_lo += a_step;
        if (++stage == p.num_stages) { stage = 0; phase ^= 1u; a_lo = a_lo0; }
      }
      if (elect_one()) {
        if (k2) umma_commit_2cta(tfull_bar(acc));   // both CTAs' epilogues
        else umma_commit(tfull_bar(acc));
      }
      __syncwarp();
      GEMM_TRACE(1, j, 2);
    }
  } else if (warp == 2) {
    // ===================== epilogue DMA =====================
    // For every tile: stage its bias (+temb) row as fp32, fetch the bf16 residual tile into the staging
    // buffer (TMA), and -- once the owning warpgroup has overwritten it with the result -- store it (TMA).
    auto tile_coords = [&](int m_tile, int n_tile, int& o0, int& w0, int& h0, int& i0) {
      const int n0 = n_tile * p.BN;
      o0 = (p.act == SDB_ACT_GEGLU) ? (n0 >> 1) : n0;   // first output column of the tile
      w0 = h0 = i0 = 0;
      if (p.mode == 1) {
        w0 = (m_tile % p.tiles_w) * p.TW;
        h0 = ((m_tile / p.tiles_w) % p.tiles_h) * p.TH;
        i0 = (m_tile / (p.tiles_w * p.tiles_h)) * p.TN;
      }
    };
    // bias (+temb) row of a tile: 8 values per lane, fetched early so the latency overlaps the store drain
    auto load_bias = [&](int m_tile, int n_tile, float (&bv)[8]) {
      const int n0 = n_tile * p.BN;
      int img = 0;
      if (p.temb_in_bias) {
        img = (m_tile / (p.tiles_w * p.tiles_h)) * p.TN;
        if (img >= p.B) img = 0;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int c = lane + 32 * i;
        float b = 0.f;
        if (c < p.BN && n0 + c < p.N) {
          if (p.bias) b = __ldg(p.bias + n0 + c);
          if (p.temb_in_bias) b += __ldg(p.temb + (long long)img * p.ld_temb + n0 + c);
        }
        bv[i] = b;
      }
    };
    auto store_bias = [&](int buf, const float (&bv)[8]) {
      const uint32_t bias_s = bias_base + (uint32_t)buf * (kMaxBN * 4);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (lane + 32 * i < p.BN) sts32f(bias_s + 4u * (lane + 32 * i), bv[i]);
    };
    // hand buffer `buf` to the warpgroup of tile j: residual tile on its way (or nothing to wait for)
    auto release = [&](int j, int buf, int m_tile, int n_tile) {
      __syncwarp();
      if (elect_one()) {
        if (p.staged && p.residual) {
          int o0, w0, h0, i0;
          tile_coords(m_tile, n_tile, o0, w0, h0, i0);
          mbar_arrive_expect_tx(res_full_bar(j & 1), 128u * (uint32_t)p.bno * 2u);
          const uint32_t stg = stg_base + (uint32_t)buf * (uint32_t)p.stg_bytes;
          for (int sl = 0; sl <= p.nslab; ++sl) {
            if (sl == p.nslab && p.tail_cols == 0) break;
            const void* tm = sl < p.nslab ? (const void*)&p.tmR : (const void*)&p.tmRt;
            const uint32_t dst = stg + (uint32_t)sl * 16384u;
            if (p.mode == 0) tma_load_2d(dst, tm, res_full_bar(j & 1), o0 + sl * 64, m_tile * kTileM);
            else tma_load_4d(dst, tm, res_full_bar(j & 1), o0 + sl * 64, w0, h0, i0);
          }
        } else {
          mbar_arrive(res_full_bar(j & 1));
        }
      }
      __syncwarp();
    };
    bool ok = true;
    TileWalk tw(p), nx(p);
    const bool bias_static = p.resB != 0;   // fixed n-tile: the bias rows are written once
    float bv[8];
    if (tw.valid()) {
      load_bias(tw.m, tw.n, bv);
      for (int b = 0; b < (bias_static ? p.nbuf : 1); ++b) store_bias(b, bv);
      release(0, 0, tw.m, tw.n);
    }
    nx.next();
    int buf = 0, nbuf_next = p.nbuf > 1 ? 1 : 0;
    for (int j = 0; tw.valid(); tw.next(), nx.next(), ++j) {
      const int m_tile = tw.m, n_tile = tw.n, m2 = nx.m, n2 = nx.n;
      const uint32_t use_phase = (uint32_t)(j >> 1) & 1u;   // barriers are per warpgroup, whatever the buffer
      const bool more = nx.valid();
      GEMM_TRACE(2, j, 0);
      if (more && !bias_static) load_bias(m2, n2, bv);
      if (p.nbuf >= 2 && more) {
        // buffer (j+1) % nbuf was last used by tile j+1-nbuf: its store must have been read out
        if (p.staged) {
          if (elect_one()) {
            if (p.nbuf == 2) tma_store_wait_read(); else tma_store_wait_read1();
          }
          __syncwarp();
        }
        if (!bias_static) store_bias(nbuf_next, bv);
        release(j + 1, nbuf_next, m2, n2);
      }
      GEMM_TRACE(2, j, 1);
      if (ok) ok = mbar_wait(stg_full_bar(j & 1), use_phase);
      GEMM_TRACE(2, j, 2);
      if (p.staged) {
        if (elect_one() && ok) {
          int o0, w0, h0, i0;
          tile_coords(m_tile, n_tile, o0, w0, h0, i0);
          const uint32_t stg = stg_base + (uint32_t)buf * (uint32_t)p.stg_bytes;
          for (int sl = 0; sl <= p.nslab; ++sl) {
            if (sl == p.nslab && p.tail_cols == 0) break;
            const void* tm = sl < p.nslab ? (const void*)&p.tmO : (const void*)&p.tmOt;
            const uint32_t src = stg + (uint32_t)sl * 16384u;
            if (p.mode == 0) tma_store_2d(tm, src, o0 + sl * 64, m_tile * kTileM);
            else tma_store_4d(tm, src, o0 + sl * 64, w0, h0, i0);
          }
          tma_store_commit();
        }
        __syncwarp();
      }
      GEMM_TRACE(2, j, 3);
      if (p.nbuf == 1 && more) {
        if (p.staged) {
          if (elect_one()) tma_store_wait_read();
          __syncwarp();
        }
        if (!bias_static) store_bias(0, bv);
        release(j + 1, 0, m2, n2);
      }
      buf = nbuf_next;
      if (++nbuf_next >= p.nbuf) nbuf_next = 0;
    }
    if (p.staged) {
      if (elect_one()) tma_store_wait_read();            // smem must outlive the last store's reads
      __syncwarp();
    }
  } else if (warp == 3) {
    // ===================== second producer: conv weight tiles / halo blocks =====================
    if (p.mode == 1 && !p.halo) {
      // the weight tile of every (tap, channel block) stage; the byte count was posted by warp 0
      int stage = 0;
      uint32_t phase = 0;
      bool ok = true;
      const int n_tap = p.num_kb / p.kb_per_tap;
      for (TileWalk tw(p); tw.valid() && ok; tw.next()) {
        const int nrow = tw.n * p.BN + (k2 ? (int)pair_rank * (p.BN >> 1) : 0);
        for (int tap = 0; tap < n_tap && ok; ++tap) {
          for (int c0 = 0; c0 < p.kb_per_tap * p.BK; c0 += p.BK) {
            ok = mbar_wait(empty_bar(stage), phase ^ 1u);
            if (!ok) break;
            const uint32_t sb = ring_base + (uint32_t)stage * p.stage_bytes + (uint32_t)p.a_bytes;
            if (elect_one()) {
              if (k2) tma_load_2d_2cta(sb, &p.tmB, sdb_mapa(full_bar(stage), 0u), tap * p.Cin + c0, nrow);
              else tma_load_2d(sb, &p.tmB, full_bar(stage), tap * p.Cin + c0, nrow);
            }
            __syncwarp();
            if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
    if (p.halo) {
      int stage = 0;
      uint32_t phase = 0;
      bool ok = true;
      const uint32_t txa = (k2 ? 2u : 1u) * (uint32_t)p.halo_box_bytes;
      for (TileWalk tw(p); tw.valid() && ok; tw.next()) {
        const int m_tile = tw.m;
        const int w0 = (m_tile % p.tiles_w) * p.TW + p.halo_dw0;
        const int h0 = ((m_tile / p.tiles_w) % p.tiles_h) * p.TH + p.halo_dh0;
        const int i0 = (m_tile / (p.tiles_w * p.tiles_h)) * p.TN;
        for (int kb = 0; kb < p.kb_per_tap; ++kb) {
          ok = mbar_wait(empty_bar(stage), phase ^ 1u);
          if (!ok) break;
          const uint32_t sa = ring_base + (uint32_t)stage * (uint32_t)p.halo_stage_bytes;
          if (elect_one()) {
            if (k2) {
              const uint32_t lead_full = sdb_mapa(full_bar(stage), 0u);
              if (pair_rank == 0) mbar_arrive_expect_tx(full_bar(stage), txa);
              tma_load_4d_2cta(sa, &p.tmA[0], lead_full, kb * p.BK, w0, h0, i0);
            } else {
              mbar_arrive_expect_tx(full_bar(stage), txa);
              tma_load_4d(sa, &p.tmA[0], full_bar(stage), kb * p.BK, w0, h0, i0);
            }
          }
          __syncwarp();
          if (++stage == p.halo_nA) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warpgroups =====================
    // Warpgroup wg owns accumulator stage wg and every second tile; thread = one output row.
    const int wg = (warp - 4) >> 2;
    const int q = warp & 3;            // TMEM lane quarter this warp may access
    const int r = q * 32 + lane;
    const int nchunk = (p.BN + 31) >> 5;
    bool ok = true;
    TileWalk tw(p);
    if (wg) tw.next();
    int buf = wg < p.nbuf ? wg : 0;                       // j % nbuf
    // channel sums for the consuming GroupNorm: this thread's column pair and row range of every staged tile
    const int st_ncp = p.bno >> 1;
    const int st_cp = p.chsum ? r % st_ncp : 0, st_sp = p.chsum ? r / st_ncp : 0;
    const bool st_on = p.chsum && st_sp < p.st_splits;
    uint32_t st_off[8], st_blk = 0;
    float st_s[4] = {0.f, 0.f, 0.f, 0.f}, st_q[4] = {0.f, 0.f, 0.f, 0.f};   // two interleaved chains each
    int st_key = -1;
    if (st_on) {
      const int oc = 2 * st_cp;
#pragma unroll
      for (int i = 0; i < 8; ++i) st_off[i] = stage_addr(p, 0u, i, oc) + (uint32_t)(oc & 7) * 2u;   // swizzle period: 8 rows
      st_blk = ((oc >> 6) < p.nslab ? 128u : (uint32_t)p.tail_cols * 2u) * 8u;
    }
    auto st_flush = [&]() {
      if (st_key >= 0) {
        const int img = st_key / p.tiles_n, col = (st_key - img * p.tiles_n) * p.BN + 2 * st_cp;
        if (col < p.N) {
          // fixed point (common.cuh: kChanSumScale / kChanSqScale): integer adds commute, so the totals -- and with them
          // every downstream result -- do not depend on the order the CTAs arrive in
          unsigned long long* d = p.chsum + ((long long)img * p.N + col) * 2;
          atomicAdd(d, (unsigned long long)__float2ll_rn((st_s[0] + st_s[2]) * kChanSumScale));
          atomicAdd(d + 1, __float2ull_rn((st_q[0] + st_q[2]) * kChanSqScale));
          atomicAdd(d + 2, (unsigned long long)__float2ll_rn((st_s[1] + st_s[3]) * kChanSumScale));
          atomicAdd(d + 3, __float2ull_rn((st_q[1] + st_q[3]) * kChanSqScale));
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) st_s[i] = st_q[i] = 0.f;
    };
    for (int j = wg; tw.valid(); tw.next(), tw.next(), j += 2) {
      const int m_tile = tw.m, n_tile = tw.n;
      const int n0 = n_tile * p.BN;
      const uint32_t acc_phase = (uint32_t)(j >> 1) & 1u;
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
          else if (p.out_mode == 2)
            out_row = ((long long)img * (2 * p.H) + 2 * h + p.ph_a) * (2 * p.W) + 2 * w + p.ph_b;
          else if (((h | w) & 1) == 0)
            out_row = ((long long)img * (p.H >> 1) + (h >> 1)) * (p.W >> 1) + (w >> 1);
        }
        if (img >= p.B) img = p.B - 1;   // clipped rows: any valid temb row will do
      }
      const float* temb_row = (kEpi == kEpiGeneric && p.temb && !p.temb_in_bias)
                                  ? p.temb + (long long)img * p.ld_temb + n0 : nullptr;
      const uint32_t stg = stg_base + (uint32_t)buf * (uint32_t)p.stg_bytes;
      const uint32_t bias_s = bias_base + (uint32_t)buf * (kMaxBN * 4);
      GEMM_TRACE(3 + wg, j >> 1, 0);
      if (ok) ok = mbar_wait(res_full_bar(wg), acc_phase);
      GEMM_TRACE(3 + wg, j >> 1, 1);
      if (ok) ok = mbar_wait(tfull_bar(wg), acc_phase);
      GEMM_TRACE(3 + wg, j >> 1, 2);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)wg * kMaxBN;
      float ln_s = 0.f, ln_ss = 0.f;
      auto process16 = [&](const uint32_t (&v)[16], int tcol) {
        if (kEpi != kEpiGeneric || p.staged) {
          epilogue16_staged<kEpi>(p, v, stg, bias_s, temb_row, r, tcol, ln_s, ln_ss);
        } else if (out_row >= 0 && ok && n0 + tcol < p.N) {
          epilogue16_direct(p, v, bias_s, temb_row, out_row, n0, tcol);
        }
      };
      // both 16-column halves in one basic block: their dependency chains interleave
      auto process = [&](const uint32_t (&v)[32], int tcol) {
        const uint32_t(&lo)[16] = *reinterpret_cast<const uint32_t(*)[16]>(&v[0]);
        const uint32_t(&hi)[16] = *reinterpret_cast<const uint32_t(*)[16]>(&v[16]);
        if (tcol + 16 < p.BN) {
          process16(lo, tcol);
          process16(hi, tcol + 16);
        } else {
          process16(lo, tcol);
        }
      };
      // 32-column chunks, the TMEM load of chunk c+1 in flight while chunk c is processed
      uint32_t va[32], vb[32];
      tmem_ld_32x32(taddr, va);
#pragma unroll 1
      for (int c = 0; c < nchunk; c += 2) {
        tmem_ld_wait();
        if (c + 1 < nchunk) tmem_ld_32x32(taddr + (uint32_t)(c + 1) * 32u, vb);
        process(va, c * 32);
        if (c + 1 < nchunk) {
          tmem_ld_wait();
          if (c + 2 < nchunk) tmem_ld_32x32(taddr + (uint32_t)(c + 2) * 32u, va);
          process(vb, (c + 1) * 32);
        }
      }
      tc_fence_before();
      if (k2) mbar_arrive_cluster(sdb_mapa(tempty_bar(wg), 0u));   // the leader's barrier counts both CTAs' rows
      else mbar_arrive(tempty_bar(wg));
      if (kEpi == kEpiResLN) {
        // the accumulator stage is already handed back: the normalise pass works on the parked bf16 row only
        const float inv_n = 1.0f / (float)p.BN;
        const float mean = ln_s * inv_n;
        float var = fmaf(-mean, mean, ln_ss * inv_n);
        var = var > 0.f ? var : 0.f;
        const float rstd = rsqrtf(var + p.ln_eps);
#pragma unroll 1
        for (int tcol = 0; tcol < p.BN; tcol += 16) layernorm16_staged(p, stg, r, tcol, mean, rstd);
      }
      if (p.chsum) {
        // the whole tile is staged once all four warps are here; the store may not start before the column pass ends
        // (the buffer goes back to the DMA warp, which may refill it with a residual tile)
        asm volatile("bar.sync %0, 128;" ::"r"(1 + wg) : "memory");
        if (st_on && ok && m_tile < p.tiles_m && n_tile < p.tiles_n) {   // (pair mode walks ghost tiles too)
          const int key = (m_tile / p.st_tiles_img) * p.tiles_n + n_tile;
          if (key != st_key) {
            st_flush();
            st_key = key;
          }
          if (n0 + 2 * st_cp < p.N) {
            uint32_t a = stg + (uint32_t)((st_sp * p.st_rows) >> 3) * st_blk;
            for (int rb = 0; rb < (p.st_rows >> 3); ++rb, a += st_blk) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const uint32_t x = lds32(a + st_off[i]);
                const float lo = __uint_as_float(x << 16), hi = __uint_as_float(x & 0xffff0000u);
                const int c = (i & 1) * 2;
                fadd2(st_s[c], st_s[c + 1], st_s[c], st_s[c + 1], lo, hi);
                ffma2(st_q[c], st_q[c + 1], lo, hi, lo, hi, st_q[c], st_q[c + 1]);
              }
            }
          }
        }
      }
      if (p.staged) fence_proxy_async();                   // generic-proxy smem writes -> async proxy
      mbar_arrive(stg_full_bar(wg));
      GEMM_TRACE(3 + wg, j >> 1, 3);
      buf += 2;
      while (buf >= p.nbuf) buf -= p.nbuf;
    }
    if (p.chsum) st_flush();
  }
  tc_fence_before();
  __syncthreads();
#ifdef SDB_GEMM_TRACE
  if (threadIdx.x == 0 && blockIdx.x < 160) {
    g_gemm_cta_ns[blockIdx.x * 4 + 1] = (long long)globaltimer_ns();
    g_gemm_cta_ns[blockIdx.x * 4 + 3] = clock64();
  }
#endif
  if (k2) sdb_cluster_sync();   // neither CTA frees TMEM / leaves while the pair still works on its shared memory
  if (warp == 1) {
    tc_fence_after();
    if (k2) tmem_dealloc_2cta(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
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
// resolution, bf16 or no residual, a tail slab the swizzle modes can express) and sizes one staging buffer.
static void plan_staging(UmmaParams& p, bool bf16_out, bool res_ok, bool same_res) {
  static int off = -1;
  if (off < 0) { const char* e = getenv("SDB_GEMM_DIRECT"); off = e ? atoi(e) : 0; }
  p.bno = p.act == SDB_ACT_GEGLU ? p.BN / 2 : p.BN;
  p.nslab = p.bno / 64;
  p.tail_cols = p.bno % 64;
  p.staged = (!off && bf16_out && res_ok && same_res &&
              (p.tail_cols == 0 || p.tail_cols == 16 || p.tail_cols == 32)) ? 1 : 0;
  p.stg_bytes = p.staged ? ((kTileM * p.bno * 2 + 1023) & ~1023) : 0;
}

// Splits the shared-memory budget between the resident weight tile (short-K GEMMs), the operand ring and
// the staging buffers.  Needs mode, BK, BN, num_kb, tiles_m, tiles_n and plan_staging() done.
static void finish_stages(UmmaParams& p) {
  static int no_resb = -1;
  if (no_resb < 0) { const char* e = getenv("SDB_GEMM_NO_RESB"); no_resb = e ? atoi(e) : 0; }
  const int sms = sdb_num_sms();
  p.a_bytes = kTileM * p.BK * 2;
  p.b_bytes = (p.pair ? p.BN / 2 : p.BN) * p.BK * 2;     // pair mode: each CTA stages half of the weight tile
  p.resB = 0;
  p.bres_bytes = 0;
  p.m_groups = 1;
  static int force_nbuf = -1;
  if (force_nbuf < 0) { const char* e = getenv("SDB_GEMM_NBUF"); force_nbuf = e ? atoi(e) : 0; }
  p.nbuf = 2;
  if (p.mode == 0 && !no_resb && p.tiles_n <= sms) {
    // weights stay resident when they leave room for the staging buffers and a >= 3-deep A ring, and
    // every CTA gets enough m-tiles to amortise loading them.  Three staging buffers (the store of tile j
    // then never delays the same warpgroup's tile j+2) only if the A ring stays >= 6 deep: measured, ring
    // depth matters more than the third buffer.
    const int bres = p.num_kb * p.b_bytes;
    const int groups = sms / p.tiles_n;
    for (int nb = (p.staged ? 3 : 2); nb >= 2 && p.tiles_m >= 2 * groups; --nb) {
      if (force_nbuf && nb != force_nbuf && force_nbuf >= 2) continue;
      if (bres + nb * p.stg_bytes + (nb == 3 ? 6 : 3) * p.a_bytes <= kSmemBudget) {
        p.resB = 1;
        p.bres_bytes = bres;
        p.m_groups = groups;
        p.nbuf = nb;
        break;
      }
    }
  }
  p.stage_bytes = p.resB ? p.a_bytes : ((p.a_bytes + p.b_bytes + 1023) & ~1023);
  // without resident weights: one staging buffer per epilogue warpgroup when the tile is epilogue-heavy
  // (few k-blocks) and the ring stays >= 4 deep; long main loops hide a shared buffer's hand-off
  if (p.staged && !p.resB &&
      (p.num_kb > 12 || (kSmemBudget - 2 * p.stg_bytes) / p.stage_bytes < 4))
    p.nbuf = 1;
  int s = (kSmemBudget - p.bres_bytes - p.nbuf * p.stg_bytes) / p.stage_bytes;
  p.num_stages = s > kMaxStages ? kMaxStages : s;
  p.ring_bytes = p.num_stages * p.stage_bytes;
}

// Halo loader plan (UmmaParams::halo) for a conv whose tile is 128 pixels of one image row; call after finish_stages().
// `n_tap` taps with offsets p.tap_dh / p.tap_dw.  Keeps the classic ring when the blocks do not fit.
static void plan_halo(UmmaParams& p, int n_tap) {
  static int halo_on = -1;
  if (halo_on < 0) { const char* e = getenv("SDB_CONV_HALO"); halo_on = e ? atoi(e) : 1; }
  if (!halo_on || p.mode != 1 || p.BK != 64 || p.TW != kTileM || p.TH != 1 || p.TN != 1) return;
  int dh0 = 127, dh1 = -127, dw0 = 127, dw1 = -127;
  for (int t = 0; t < n_tap; ++t) {
    dh0 = p.tap_dh[t] < dh0 ? p.tap_dh[t] : dh0; dh1 = p.tap_dh[t] > dh1 ? p.tap_dh[t] : dh1;
    dw0 = p.tap_dw[t] < dw0 ? p.tap_dw[t] : dw0; dw1 = p.tap_dw[t] > dw1 ? p.tap_dw[t] : dw1;
  }
  const int rows = dh1 - dh0 + 1, hw = kTileM + (dw1 - dw0);
  const int box = rows * hw * 128;
  const int stage = (box + 1023) & ~1023;
  // taps per weight stage: a whole tap row (3 of 3x3, 2 of the 2x2 phases) when three such stages fit next to two halo
  // blocks -- tiles of <= 128 columns -- else one (SDB_CONV_HALO_TG=1 forces that: A/B)
  static int tg_env = -1;
  if (tg_env < 0) { const char* e = getenv("SDB_CONV_HALO_TG"); tg_env = e ? atoi(e) : 0; }
  const int tg_row = n_tap == 9 ? 3 : (n_tap == 4 ? 2 : 1);
  for (int pass = 0; pass < 2; ++pass) {
  const int tg = pass == 0 ? (tg_env > 0 ? (tg_env < tg_row ? tg_env : tg_row) : tg_row) : 1;
  if (pass == 1 && tg_row == 1) break;
  if (n_tap % tg != 0) continue;
  for (int nA = 3; nA >= 2; --nA) {
    const int rem = kSmemBudget - p.nbuf * p.stg_bytes - nA * stage;
    int nB = rem > 0 ? rem / (tg * p.b_bytes) : 0;
    if (nB > kMaxStages - nA) nB = kMaxStages - nA;
    if (nB < (tg > 1 ? 3 : (nA == 3 ? 4 : 3))) continue;
    p.halo = 1;
    p.halo_tg = tg;
    p.halo_w = hw;
    p.halo_nA = nA;
    p.halo_nB = nB;
    p.halo_stage_bytes = stage;
    p.halo_box_bytes = box;
    p.halo_dw0 = dw0;
    p.halo_dh0 = dh0;
    for (int t = 0; t < n_tap; ++t) p.tap_row[t] = (short)((p.tap_dh[t] - dh0) * hw + (p.tap_dw[t] - dw0));
    p.num_stages = nA + nB;
    p.ring_bytes = nA * stage + nB * tg * p.b_bytes;
    return;
  }
  }
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

template <int kEpi, bool k2 = false>
static int launch_umma_epi(const UmmaParams& p, int grid, int smem, cudaStream_t st) {
  static bool attr_set[64] = {false};
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    SDB_CUDA_OK(cudaFuncSetAttribute(umma_gemm_kernel<kEpi, k2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  if (k2) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid, 1, 1);
    cfg.blockDim = dim3(kGemmThreads, 1, 1);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    SDB_CUDA_OK(cudaLaunchKernelEx(&cfg, umma_gemm_kernel<kEpi, k2>, p));
  } else {
    umma_gemm_kernel<kEpi, k2><<<grid, kGemmThreads, smem, st>>>(p);
  }
  SDB_LAUNCH_OK();
  return SDB_OK;
}

static int launch_umma(UmmaParams& p, cudaStream_t st) {
  const int smem = kSmemBudget + kBiasBytes + 1024 /* alignment */ + 512 /* barriers */;
  if (p.num_stages < 2) {
    sdb_set_error("umma: tile %dx%d (BK %d) leaves no room for a 2-stage operand ring", kTileM, p.BN, p.BK);
    return SDB_EUNSUPPORTED;
  }
  const int total = p.tiles_m * p.tiles_n;
  const int sms = sdb_num_sms();
  int grid = p.resB ? p.m_groups * p.tiles_n : (total < sms ? total : sms);
  if (p.pair) {
    p.tiles_m_pad = (p.tiles_m + 1) / 2 * 2;
    const int padded = p.tiles_m_pad * p.tiles_n;
    grid = (padded < sms ? padded : sms) / 2 * 2;
    p.tiles_per_cta = (padded + grid - 1) / grid;
  }
  p.walk_dm = p.resB ? p.m_groups : grid / p.tiles_n;
  p.walk_dn = p.resB ? 0 : grid % p.tiles_n;
  const bool slow_temb = p.temb && !p.temb_in_bias;
  if (p.pair) {      // conv only: plain / residual / generic epilogues
    if (p.staged && !slow_temb)
      return p.residual ? launch_umma_epi<kEpiRes, true>(p, grid, smem, st) : launch_umma_epi<kEpiPlain, true>(p, grid, smem, st);
    return launch_umma_epi<kEpiGeneric, true>(p, grid, smem, st);
  }
  if (p.ln_gamma) return launch_umma_epi<kEpiResLN>(p, grid, smem, st);   // (checked by sdb_gemm: staged, residual, one n-tile)
  if (p.staged && p.act == SDB_ACT_GEGLU) return launch_umma_epi<kEpiGeglu>(p, grid, smem, st);
  if (p.staged && p.act == SDB_ACT_NONE && !slow_temb)
    return p.residual ? launch_umma_epi<kEpiRes>(p, grid, smem, st) : launch_umma_epi<kEpiPlain>(p, grid, smem, st);
  return launch_umma_epi<kEpiGeneric>(p, grid, smem, st);
}

// Channel sums in the epilogue (UmmaParams::chsum): possible when the tile is staged as bf16 without an activation and
// every 128-row tile lies inside one image with no clipped rows.  Returns false (and leaves p untouched) otherwise.
static bool plan_chan_sums(UmmaParams& p, unsigned long long* chsum, int tiles_per_image, bool whole_tiles) {
  if (!p.staged || p.act != SDB_ACT_NONE || p.ln_gamma || p.out_mode == 1 || !whole_tiles || tiles_per_image <= 0 ||
      ((uintptr_t)chsum & 15) != 0)
    return false;
  p.chsum = chsum;
  p.st_tiles_img = tiles_per_image;
  const int ncp = p.bno / 2;
  int splits = 1;
  while (splits * 2 * ncp <= kTileM) splits *= 2;
  p.st_splits = splits;
  p.st_rows = kTileM / splits;
  return true;
}

}  // namespace sdb

#ifdef SDB_GEMM_TRACE
extern "C" int sdb_debug_gemm_trace(long long* host_dst) {
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  return cudaMemcpyFromSymbol(host_dst, sdb::g_gemm_trace, sizeof(long long) * 5 * 16 * 8) == cudaSuccess ? 0 : -1;
}
extern "C" int sdb_debug_gemm_cta_ns(long long* host_dst) {
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  return cudaMemcpyFromSymbol(host_dst, sdb::g_gemm_cta_ns, sizeof(long long) * 160 * 4) == cudaSuccess ? 0 : -1;
}
#endif

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
  // short-K GEMMs are bound by the output stream, not the tensor pipe: tiles of <= 128 output columns keep
  // two staging buffers and the resident weight tile inside shared memory
  p.act = a->act;
  p.kb1 = (a->K1 + p.BK - 1) / p.BK;
  p.K1 = a->K1;
  p.num_kb = p.kb1 + (K2 + p.BK - 1) / p.BK;
  p.tiles_m = (a->M + kTileM - 1) / kTileM;
  static int bncap = -1;   // experiments: SDB_GEMM_BNCAP=<n> caps the tile width
  if (bncap < 0) { const char* e = getenv("SDB_GEMM_BNCAP"); bncap = e ? atoi(e) : 0; }
  int cap0 = (K <= 320 && a->act != SDB_ACT_GEGLU) ? 128 : kMaxBN;
  if (bncap && bncap < cap0) cap0 = bncap;
  for (int cap = cap0;; cap = 128) {
    p.BN = pick_bn(a->N, cap);
    if (K <= 320 && a->N <= 192 && !(bncap && bncap < a->N)) p.BN = a->N;   // one n-tile: the whole weight matrix stays resident
    plan_staging(p, a->out_dtype == SDB_BF16, !a->residual || a->res_dtype == SDB_BF16, true);
    p.tiles_n = (a->N + p.BN - 1) / p.BN;
    finish_stages(p);
    if (p.resB || cap <= 128 || K > 320) break;   // a narrower tile may make the weights fit
  }
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
  if (a->ln_gamma) {
    SDB_CHECK_ARG(a->ln_beta && ((uintptr_t)a->ln_gamma & 15) == 0 && ((uintptr_t)a->ln_beta & 15) == 0,
                  "gemm: ln_gamma needs ln_beta, both 16-byte aligned");
    if (!(p.staged && p.tiles_n == 1 && p.BN == a->N && a->residual && a->res_dtype == SDB_BF16 &&
          a->act == SDB_ACT_NONE)) {
      sdb_set_error("gemm: the LayerNorm epilogue needs bf16 output + bf16 residual, no activation and N (%d) in one "
                    "tile of <= 256 columns", a->N);
      return SDB_EUNSUPPORTED;
    }
    p.ln_gamma = a->ln_gamma;
    p.ln_beta = a->ln_beta;
    p.ln_eps = a->ln_eps;
  }
  if (a->chan_sums) {
    const long long rpi = a->rows_per_image;
    if (!(rpi > 0 && rpi % kTileM == 0 && a->M % rpi == 0 && a->out_dtype == SDB_BF16 &&
          plan_chan_sums(p, reinterpret_cast<unsigned long long*>(a->chan_sums), (int)(rpi / kTileM), true))) {
      sdb_set_error("gemm: chan_sums needs bf16 output through the staged epilogue, no activation / LayerNorm, and "
                    "rows_per_image (%lld) a multiple of 128 dividing M (%d)", rpi, a->M);
      return SDB_EUNSUPPORTED;
    }
  }

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
  SDB_CHECK_ARG(!(a->up2 && a->down2), "conv3x3: up2 and down2 are exclusive");
  SDB_CHECK_ARG(a->out_dtype != SDB_U8 || (a->Cout == 16 && a->out_channels > 0 && a->out_channels <= 16 && !a->residual &&
                                           !a->down2 && !a->up2 && !a->temb),
                "conv3x3: uint8 output is the image head only (Cout padded to 16, 1 <= out_channels <= 16, no residual)");
  SDB_CHECK_ARG(!(a->up2 && a->temb), "conv3x3: up2 takes no time embedding");
  SDB_CHECK_ARG(!a->down2 || (a->H % 2 == 0 && a->W % 2 == 0), "conv3x3: down2 needs even H, W");
  SDB_CHECK_ARG(((uintptr_t)a->x & 15) == 0 && ((uintptr_t)a->w & 15) == 0 && ((uintptr_t)a->out & 15) == 0 &&
                    ((uintptr_t)a->residual & 15) == 0 && ((uintptr_t)a->bias & 15) == 0 &&
                    ((uintptr_t)a->temb & 15) == 0 && a->ld_temb % 4 == 0,
                "conv3x3: pointers must be 16-byte aligned");

  // up2: nearest x2 THEN conv (resnet2d.py:61-62) without the 4x tensor.  Output pixel (2y+a, 2x+b) sees only a
  // 2x2 neighbourhood of the low-resolution input -- rows {y-1, y} for a = 0, {y, y+1} for a = 1, same for the
  // columns -- with the 3x3 taps that land on the same source pixel summed.  So the layer is four convolutions
  // with 2x2 taps ("phases"), 4/9 of the flops, each writing its own interleaved quarter of the output through
  // a strided TMA store.  `w` then holds the phase-collapsed weights [4 phases = 2a+b][Cout][2][2][Cin]
  // (taps summed in fp32 before the bf16 rounding: _pack.conv3x3_up_w).
  const int n_phase = a->up2 ? 4 : 1;
  const int n_tap = a->up2 ? 4 : 9;
  for (int phase = 0; phase < n_phase; ++phase) {
  const int ph_a = phase >> 1, ph_b = phase & 1;
  const __nv_bfloat16* wph = reinterpret_cast<const __nv_bfloat16*>(a->w) + (size_t)phase * a->Cout * n_tap * a->Cin;
  const int Wo = a->up2 ? 2 * a->W : a->W, Ho = a->up2 ? 2 * a->H : a->H;
  const size_t ph_off = a->up2 ? ((size_t)ph_a * Wo + ph_b) * a->Cout : 0;
  UmmaParams p;
  memset(&p, 0, sizeof(p));
  p.mode = 1;
  pick_bk(a->Cin, 0, &p.BK, &p.swz);
  p.BN = pick_bn(a->Cout);
  p.act = SDB_ACT_NONE;
  // CTA pairs (cta_group::2, M = 256): whenever the half weight tile keeps whole 8-row swizzle atoms and there is
  // more than a handful of m-tiles (SDB_CONV_PAIR=0 switches it off: A/B measurements)
  static int pair_on = -1;
  if (pair_on < 0) { const char* e = getenv("SDB_CONV_PAIR"); pair_on = e ? atoi(e) : 1; }
  p.pair = (pair_on && p.BK == 64 && p.BN % 32 == 0 && (long long)a->B * a->H * a->W >= 8 * kTileM) ? 1 : 0;
  plan_staging(p, a->out_dtype == SDB_BF16, !a->residual || a->res_dtype == SDB_BF16, !a->down2);
  p.Cin = a->Cin;
  p.kb_per_tap = (a->Cin + p.BK - 1) / p.BK;
  p.num_kb = n_tap * p.kb_per_tap;
  // tile = TW x TH x TN pixels with TW*TH*TN = 128, all powers of two.  TW = the largest power-of-two factor
  // of W when that is >= 8 (no waste); otherwise the width in {8..128} that wastes the fewest clipped
  // columns (the TMA unit zero-fills the loads and clips the stores of a box that overhangs the image).
  int tw = 1;
  while (tw < 128 && a->W % (tw * 2) == 0) tw *= 2;
  if (tw < 8) {
    int best = 8;
    double best_waste = 1e30;
    for (int c = 8; c <= 128; c *= 2) {
      const double waste = (double)((a->W + c - 1) / c * c) / a->W;
      if (waste < best_waste - 1e-9) { best_waste = waste; best = c; }
    }
    tw = best;
  }
  int th = 1;
  while (tw * th < 128 && th < a->H) th *= 2;
  int tn = 128 / (tw * th);
  if (tw * th * tn != 128) {
    sdb_set_error("conv3x3: cannot tile W=%d H=%d into 128-pixel boxes", a->W, a->H);
    return SDB_EUNSUPPORTED;
  }
  p.TW = tw; p.TH = th; p.TN = tn;
  p.tiles_w = (a->W + tw - 1) / tw;
  p.tiles_h = (a->H + th - 1) / th;
  const int tiles_b = (a->B + tn - 1) / tn;
  p.tiles_m = p.tiles_w * p.tiles_h * tiles_b;
  p.tiles_n = (a->Cout + p.BN - 1) / p.BN;
  finish_stages(p);
  p.temb_in_bias = (a->temb && tn == 1) ? 1 : 0;
  p.B = a->B; p.H = a->H; p.W = a->W;
  p.out_mode = a->down2 ? 1 : (a->up2 ? 2 : 0);
  p.ph_a = ph_a; p.ph_b = ph_b;
  for (int t = 0; t < n_tap; ++t) {
    p.tap_map[t] = 0;
    if (a->up2) {
      p.tap_dh[t] = (signed char)(ph_a - 1 + t / 2);
      p.tap_dw[t] = (signed char)(ph_b - 1 + t % 2);
    } else {
      p.tap_dh[t] = (signed char)(t / 3 - 1);
      p.tap_dw[t] = (signed char)(t % 3 - 1);
    }
  }
  plan_halo(p, n_tap);
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
  p.u8_channels = a->out_channels;
  p.act = SDB_ACT_NONE;
  if (a->chan_sums) {
    const bool whole = tn == 1 && a->H % th == 0 && a->W % tw == 0;
    if (!(a->out_dtype == SDB_BF16 && plan_chan_sums(p, reinterpret_cast<unsigned long long*>(a->chan_sums), p.tiles_w * p.tiles_h, whole))) {
      sdb_set_error("conv3x3: chan_sums needs bf16 output through the staged epilogue, no down2, and an image that "
                    "tiles into whole 128-pixel boxes (H=%d W=%d: box %dx%dx%d)", a->H, a->W, tn, th, tw);
      return SDB_EUNSUPPORTED;
    }
  }

  uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->B};
  uint64_t strides[4] = {0, (uint64_t)a->Cin * 2, (uint64_t)a->W * a->Cin * 2,
                         (uint64_t)a->H * a->W * a->Cin * 2};
  uint32_t box[4] = {(uint32_t)p.BK, (uint32_t)tw, (uint32_t)th, (uint32_t)tn};
  if (p.halo) { box[1] = (uint32_t)p.halo_w; box[2] = (uint32_t)(p.halo_box_bytes / (p.halo_w * 128)); }
  int rc = sdb_make_tensor_map(&p.tmA[0], a->x, 2, 4, dims, strides, box, p.swz);
  if (rc) return rc;
  if (p.halo) {
    // [Cin][Cout][tap]: a box of halo_tg taps lands as that many consecutive K-major weight tiles
    uint64_t wd[3] = {(uint64_t)a->Cin, (uint64_t)a->Cout, (uint64_t)n_tap};
    uint64_t ws[3] = {0, (uint64_t)n_tap * a->Cin * 2, (uint64_t)a->Cin * 2};
    uint32_t wb[3] = {(uint32_t)p.BK, (uint32_t)(p.pair ? p.BN / 2 : p.BN), (uint32_t)p.halo_tg};
    rc = sdb_make_tensor_map(&p.tmB, wph, 2, 3, wd, ws, wb, p.swz);
  } else {
    uint64_t wd[2] = {(uint64_t)n_tap * a->Cin, (uint64_t)a->Cout};
    uint64_t ws[2] = {0, (uint64_t)n_tap * a->Cin * 2};
    uint32_t wb[2] = {(uint32_t)p.BK, (uint32_t)(p.pair ? p.BN / 2 : p.BN)};
    rc = sdb_make_tensor_map(&p.tmB, wph, 2, 2, wd, ws, wb, p.swz);
  }
  if (rc) return rc;
  if (p.staged) {
    // the output box of a phase lands on every second pixel of every second row: strides doubled, base shifted
    const uint64_t ps = a->up2 ? 2 : 1;
    uint64_t od[4] = {(uint64_t)a->Cout, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->B};
    uint64_t so[4] = {0, ps * a->Cout * 2, ps * Wo * a->Cout * 2, (uint64_t)Ho * Wo * a->Cout * 2};
    uint32_t rows[3] = {(uint32_t)tw, (uint32_t)th, (uint32_t)tn};
    const void* res_ph = a->residual ? (const void*)(reinterpret_cast<const __nv_bfloat16*>(a->residual) + ph_off) : nullptr;
    rc = make_slab_maps(p, reinterpret_cast<__nv_bfloat16*>(a->out) + ph_off, res_ph, 4, od, so, so, rows);
    if (rc) return rc;
  }
  rc = launch_umma(p, sdb_stream(stream));
  if (rc) return rc;
  }  // phases
  return SDB_OK;
}
