// Attention backward (row T1), any sequence length, head dim <= 32: flash-style recomputation, nothing of size
// [n, m] is ever stored.  Reference arithmetic: autograd of modules/transformer.py:93-114
//   S = Q K^T * scale, P = softmax(S), O = P V
//   dV = P^T dO,  dP = dO V^T,  dS = P o (dP - rowsum(dO o O)) * scale,  dQ = dS K,  dK = dS^T Q.
//
// Two kernels over warp-level mma.sync m16n8k16 tiles (bf16 operands, fp32 accumulation), S/P/dS never leave
// registers -- the same structure as the forward kernel in attention.cu, whose staging / fragment helpers are used:
//   attn_bwd_dq_kernel   one CTA = 64 queries x (batch, head): pass 1 over the keys rebuilds the row log-sum-exp
//                        (the forward kernels do not store it), pass 2 accumulates dQ; also writes lse / delta
//                        = rowsum(dO o O) for the second kernel.
//   attn_bwd_dkv_kernel  one CTA = 64 keys x (batch, head), each warp owns 16 keys and computes the TRANSPOSED
//                        tiles S^T = K Q^T, dP^T = V dO^T so that P^T / dS^T come out in the A-operand layout
//                        of the dV += P^T dO and dK += dS^T Q products.
// `hp` = columns per head of q/k/v and of their gradients (32 for the head-padded tcgen05 layout, where the pad
// columns of q/k are zero and come back with zero gradient; d for dense tensors); O / dO are dense [B, n, heads*d].
#include "common.cuh"

namespace sdb {

constexpr int kAbPitch = 40;   // 32 + 8 bf16: 80-byte rows, ldmatrix conflict-free

// Staging map of a 64-row operand tile: which (row, column) granules this thread copies, computed ONCE per kernel
// (the generic stage_rows divides by the row width for every granule of every tile -- a third of the instructions
// of these kernels before).  wide = 32-column rows at 16-byte alignment (the head-padded layout): two 16-byte
// cp.async per thread and tile; otherwise 8-byte granules, up to four per thread.
struct AbRowMap {
  int rc[4];   // row << 8 | column, or -1
  int wide;
};
__device__ __forceinline__ AbRowMap ab_row_map(int cols, bool wide) {
  AbRowMap m;
  m.wide = wide ? 1 : 0;
  if (wide) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = threadIdx.x + k * kAttThreads;
      m.rc[k] = k < 2 ? (((i >> 2) << 8) | ((i & 3) * 8)) : -1;
    }
  } else {
    const int cpr = cols >> 2;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = threadIdx.x + k * kAttThreads;
      const int r = i / cpr;
      m.rc[k] = (r < 64) ? ((r << 8) | ((i - r * cpr) * 4)) : -1;
    }
  }
  return m;
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ bool ab_wide_ok(const void* g, long long ld, int cols) {
  return cols == 32 && (ld & 7) == 0 && (reinterpret_cast<uintptr_t>(g) & 15) == 0;
}
// rows [row0, row0 + 64) of g -> smem tile; rows >= valid_rows are zero-filled
__device__ __forceinline__ void ab_stage(__nv_bfloat16* s, const __nv_bfloat16* g, long long ld, int row0, int valid_rows,
                                         const AbRowMap& m) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (m.rc[k] < 0) continue;
    const int r = m.rc[k] >> 8, c = m.rc[k] & 255;
    __nv_bfloat16* dst = s + r * kAbPitch + c;
    const bool live = row0 + r < valid_rows;
    const __nv_bfloat16* src = g + (long long)(row0 + r) * ld + c;
    if (m.wide) {
      if (live) cp_async16(smem_u32(dst), src);
      else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
    } else {
      if (live) cp_async8(smem_u32(dst), src);
      else *reinterpret_cast<uint2*>(dst) = make_uint2(0u, 0u);
    }
  }
}

// A-operand fragments (16 rows x 32 cols) of the rows [row0, row0+16) of a staged tile
__device__ __forceinline__ void ab_load_a(const __nv_bfloat16* s, int row0, int lane, uint32_t (&f)[2][4]) {
#pragma unroll
  for (int ks = 0; ks < 2; ++ks) {
    const int row = row0 + (lane & 7) + ((lane >> 3) & 1) * 8;
    const int col = ks * 16 + (lane >> 4) * 8;
    ldsm_x4(smem_u32(s + row * kAbPitch + col), f[ks][0], f[ks][1], f[ks][2], f[ks][3]);
  }
}
// acc[16 x 64] = A[16 x 32] . T[64 x 32]^T   (T staged row-major: the "col" operand as it lies)
__device__ __forceinline__ void ab_mma_nt(float (&acc)[8][4], const uint32_t (&a)[2][4], const __nv_bfloat16* t, int lane) {
#pragma unroll
  for (int i = 0; i < 8; ++i) { acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f; }
#pragma unroll
  for (int np = 0; np < 4; ++np) {
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      const int row = np * 16 + (lane & 7) + ((lane >> 4) & 1) * 8;
      const int col = ks * 16 + ((lane >> 3) & 1) * 8;
      uint32_t b0, b1, b2, b3;
      ldsm_x4(smem_u32(t + row * kAbPitch + col), b0, b1, b2, b3);
      mma_bf16(acc[2 * np], a[ks][0], a[ks][1], a[ks][2], a[ks][3], b0, b1);
      mma_bf16(acc[2 * np + 1], a[ks][0], a[ks][1], a[ks][2], a[ks][3], b2, b3);
    }
  }
}
// out[16 x 32] += P[16 x 64] . T[64 x 32]   (P as packed A fragments, T staged row-major, read transposed)
__device__ __forceinline__ void ab_mma_nn(float (&out)[4][4], const uint32_t (&p)[4][4], const __nv_bfloat16* t, int lane) {
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
    for (int dp = 0; dp < 2; ++dp) {
      const int row = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
      const int col = dp * 16 + (lane >> 4) * 8;
      uint32_t b0, b1, b2, b3;
      ldsm_x4_t(smem_u32(t + row * kAbPitch + col), b0, b1, b2, b3);
      mma_bf16(out[2 * dp], p[kk][0], p[kk][1], p[kk][2], p[kk][3], b0, b1);
      mma_bf16(out[2 * dp + 1], p[kk][0], p[kk][1], p[kk][2], p[kk][3], b2, b3);
    }
  }
}

__global__ void __launch_bounds__(kAttThreads)
attn_bwd_dq_kernel(const __nv_bfloat16* __restrict__ q, long long ldq, const __nv_bfloat16* __restrict__ k, long long ldk,
                   const __nv_bfloat16* __restrict__ v, long long ldv, const __nv_bfloat16* __restrict__ o, long long ldo,
                   const __nv_bfloat16* __restrict__ dO, long long lddo, __nv_bfloat16* __restrict__ dq, long long lddq,
                   float* __restrict__ lse_out, float* __restrict__ delta_out, int n, int m, int d, int hp, float scale,
                   float scale_log2e, int have_lse) {
  __shared__ __align__(16) __nv_bfloat16 sQ[64 * kAbPitch], sdO[64 * kAbPitch], sK[2][64 * kAbPitch], sV[2][64 * kAbPitch];
  __shared__ float sDelta[64];
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * 64, heads = gridDim.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w = hp >= 32 ? 32 : d;                                 // staged columns of q/k/v
  const __nv_bfloat16* qg = q + (long long)b * n * ldq + h * hp;
  const __nv_bfloat16* kg = k + (long long)b * m * ldk + h * hp;
  const __nv_bfloat16* vg = v + (long long)b * m * ldv + h * hp;
  const __nv_bfloat16* og = o + (long long)b * n * ldo + h * d;
  const __nv_bfloat16* dog = dO + (long long)b * n * lddo + h * d;
  {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (int i = threadIdx.x; i < 64 * kAbPitch / 8; i += kAttThreads) {
      reinterpret_cast<uint4*>(sQ)[i] = z; reinterpret_cast<uint4*>(sdO)[i] = z;
      reinterpret_cast<uint4*>(sK[0])[i] = z; reinterpret_cast<uint4*>(sK[1])[i] = z;
      reinterpret_cast<uint4*>(sV[0])[i] = z; reinterpret_cast<uint4*>(sV[1])[i] = z;
    }
  }
  __syncthreads();
  const int n_tiles = (m + 63) / 64;
  const AbRowMap map_k = ab_row_map(w, ab_wide_ok(kg, ldk, w) && ab_wide_ok(vg, ldv, w));
  // O is only needed for delta: stage it where the first V tile will go later
  stage_rows<kAbPitch>(sQ, qg, ldq, q0, n, 64, w);
  stage_rows<kAbPitch>(sdO, dog, lddo, q0, n, 64, d);
  stage_rows<kAbPitch>(sV[1], og, ldo, q0, n, 64, d);
  ab_stage(sK[0], kg, ldk, 0, m, map_k);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  {  // delta = rowsum(dO o O): two threads per row
    const int r = threadIdx.x >> 1, c0 = (threadIdx.x & 1) * 16;
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c)
      acc = fmaf(__bfloat162float(sdO[r * kAbPitch + c0 + c]), __bfloat162float(sV[1][r * kAbPitch + c0 + c]), acc);
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    if ((threadIdx.x & 1) == 0) {
      sDelta[r] = acc;
      if (q0 + r < n) delta_out[((long long)b * heads + h) * n + q0 + r] = acc;
    }
  }
  __syncthreads();
  {  // the O tile is dead: clear its pad-free columns again (V tiles with fewer staged columns rely on zero pads)
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (int i = threadIdx.x; i < 64 * kAbPitch / 8; i += kAttThreads) reinterpret_cast<uint4*>(sV[1])[i] = z;
  }
  uint32_t qf[2][4], dof[2][4];
  ab_load_a(sQ, warp * 16, lane, qf);
  ab_load_a(sdO, warp * 16, lane, dof);
  __syncthreads();

  // ---- pass 1: row log-sum-exp (log2 domain, of the scaled scores) ---------------------------------
  // (skipped when the forward kernel kept the row log-sum-exp: have_lse)
  float row_max[2] = {-INFINITY, -INFINITY}, row_sum[2] = {0.f, 0.f};
  float s[8][4];
  for (int j = 0; j < (have_lse ? 0 : n_tiles); ++j) {
    const int buf = j & 1;
    if (j + 1 < n_tiles) {
      ab_stage(sK[buf ^ 1], kg, ldk, (j + 1) * 64, m, map_k);
      cp_async_commit();
    }
    ab_mma_nt(s, qf, sK[buf], lane);
    const int key_base = j * 64 + (lane & 3) * 2;
    if (j == n_tiles - 1 && (m & 63) != 0) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int key = key_base + t * 8;
        if (key >= m) { s[t][0] = -INFINITY; s[t][2] = -INFINITY; }
        if (key + 1 >= m) { s[t][1] = -INFINITY; s[t][3] = -INFINITY; }
      }
    }
    float mx[2] = {row_max[0], row_max[1]};
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      mx[0] = fmaxf(mx[0], fmaxf(s[t][0], s[t][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[t][2], s[t][3]));
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
    }
    float psum[2] = {0.f, 0.f};
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      psum[0] += fast_exp2((s[t][0] - mx[0]) * scale_log2e) + fast_exp2((s[t][1] - mx[0]) * scale_log2e);
      psum[1] += fast_exp2((s[t][2] - mx[1]) * scale_log2e) + fast_exp2((s[t][3] - mx[1]) * scale_log2e);
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      row_sum[r] = row_sum[r] * fast_exp2((row_max[r] - mx[r]) * scale_log2e) + psum[r];
      row_max[r] = mx[r];
    }
    cp_async_wait<0>();
    __syncthreads();
  }
  float lse2[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    row_sum[r] += __shfl_xor_sync(0xffffffffu, row_sum[r], 1);
    row_sum[r] += __shfl_xor_sync(0xffffffffu, row_sum[r], 2);
    lse2[r] = row_max[r] * scale_log2e + log2f(row_sum[r]);
  }
  const int r0 = q0 + warp * 16 + (lane >> 2), r1 = r0 + 8;
  if (have_lse) {
    const float* lg = lse_out + ((long long)b * heads + h) * n;
    lse2[0] = r0 < n ? lg[r0] : 0.f;
    lse2[1] = r1 < n ? lg[r1] : 0.f;
  } else if ((lane & 3) == 0) {
    float* lg = lse_out + ((long long)b * heads + h) * n;
    if (r0 < n) lg[r0] = lse2[0];
    if (r1 < n) lg[r1] = lse2[1];
  }
  const float dl0 = sDelta[warp * 16 + (lane >> 2)], dl1 = sDelta[warp * 16 + (lane >> 2) + 8];

  // ---- pass 2: dQ -----------------------------------------------------------------------------------
  ab_stage(sK[0], kg, ldk, 0, m, map_k);
  ab_stage(sV[0], vg, ldv, 0, m, map_k);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  float acc[4][4];
#pragma unroll
  for (int t = 0; t < 4; ++t) { acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f; }
  for (int j = 0; j < n_tiles; ++j) {
    const int buf = j & 1;
    if (j + 1 < n_tiles) {
      ab_stage(sK[buf ^ 1], kg, ldk, (j + 1) * 64, m, map_k);
      ab_stage(sV[buf ^ 1], vg, ldv, (j + 1) * 64, m, map_k);
      cp_async_commit();
    }
    float dp[8][4];
    ab_mma_nt(s, qf, sK[buf], lane);
    ab_mma_nt(dp, dof, sV[buf], lane);
    const int key_base = j * 64 + (lane & 3) * 2;
    uint32_t dsf[4][4];
    // dS / scale = P o (dP - delta); the scale is applied once to the accumulated dQ.  Keys past m exist only in
    // the ragged last tile: their scores are forced to -inf there (P = 0), every other tile runs unmasked.
    if (j == n_tiles - 1 && (m & 63) != 0) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int key = key_base + t * 8;
        if (key >= m) { s[t][0] = -INFINITY; s[t][2] = -INFINITY; }
        if (key + 1 >= m) { s[t][1] = -INFINITY; s[t][3] = -INFINITY; }
      }
    }
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const float p0 = fast_exp2(fmaf(s[t][0], scale_log2e, -lse2[0]));
      const float p1 = fast_exp2(fmaf(s[t][1], scale_log2e, -lse2[0]));
      const float p2 = fast_exp2(fmaf(s[t][2], scale_log2e, -lse2[1]));
      const float p3 = fast_exp2(fmaf(s[t][3], scale_log2e, -lse2[1]));
      dsf[t >> 1][(t & 1) * 2] = pack_bf16x2(p0 * (dp[t][0] - dl0), p1 * (dp[t][1] - dl0));
      dsf[t >> 1][(t & 1) * 2 + 1] = pack_bf16x2(p2 * (dp[t][2] - dl1), p3 * (dp[t][3] - dl1));
    }
    ab_mma_nn(acc, dsf, sK[buf], lane);
    cp_async_wait<0>();
    __syncthreads();
  }
  __nv_bfloat16* dqg = dq + (long long)b * n * lddq + h * hp;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int col = t * 8 + (lane & 3) * 2;
    if (col < w) {
      if (r0 < n) *reinterpret_cast<uint32_t*>(dqg + (long long)r0 * lddq + col) = pack_bf16x2(acc[t][0] * scale, acc[t][1] * scale);
      if (r1 < n) *reinterpret_cast<uint32_t*>(dqg + (long long)r1 * lddq + col) = pack_bf16x2(acc[t][2] * scale, acc[t][3] * scale);
    }
  }
}

__global__ void __launch_bounds__(kAttThreads)
attn_bwd_dkv_kernel(const __nv_bfloat16* __restrict__ q, long long ldq, const __nv_bfloat16* __restrict__ k, long long ldk,
                    const __nv_bfloat16* __restrict__ v, long long ldv, const __nv_bfloat16* __restrict__ dO, long long lddo,
                    const float* __restrict__ lse, const float* __restrict__ delta, __nv_bfloat16* __restrict__ dk,
                    long long lddk, __nv_bfloat16* __restrict__ dv, long long lddv, int n, int m, int d, int hp, float scale,
                    float scale_log2e) {
  __shared__ __align__(16) __nv_bfloat16 sK[64 * kAbPitch], sV[64 * kAbPitch], sQ[2][64 * kAbPitch], sdO[2][64 * kAbPitch];
  __shared__ __align__(8) float sLse[2][64], sDelta[2][64];
  const int b = blockIdx.z, h = blockIdx.y, k0 = blockIdx.x * 64, heads = gridDim.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w = hp >= 32 ? 32 : d;
  const __nv_bfloat16* qg = q + (long long)b * n * ldq + h * hp;
  const __nv_bfloat16* kg = k + (long long)b * m * ldk + h * hp;
  const __nv_bfloat16* vg = v + (long long)b * m * ldv + h * hp;
  const __nv_bfloat16* dog = dO + (long long)b * n * lddo + h * d;
  const float* lg = lse + ((long long)b * heads + h) * n;
  const float* dg = delta + ((long long)b * heads + h) * n;
  {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (int i = threadIdx.x; i < 64 * kAbPitch / 8; i += kAttThreads) {
      reinterpret_cast<uint4*>(sK)[i] = z; reinterpret_cast<uint4*>(sV)[i] = z;
      reinterpret_cast<uint4*>(sQ[0])[i] = z; reinterpret_cast<uint4*>(sQ[1])[i] = z;
      reinterpret_cast<uint4*>(sdO[0])[i] = z; reinterpret_cast<uint4*>(sdO[1])[i] = z;
    }
  }
  __syncthreads();
  const AbRowMap map_q = ab_row_map(w, ab_wide_ok(qg, ldq, w));
  const AbRowMap map_do = ab_row_map(d, ab_wide_ok(dog, lddo, d));
  auto stage_q = [&](int buf, int row0) {
    ab_stage(sQ[buf], qg, ldq, row0, n, map_q);
    ab_stage(sdO[buf], dog, lddo, row0, n, map_do);
    if (threadIdx.x < 64) {
      const int r = row0 + threadIdx.x;
      sLse[buf][threadIdx.x] = r < n ? -lg[r] : 0.f;     // stored negated; rows past n: Q = dO = 0, so P^T = 1 meets only zeros
      sDelta[buf][threadIdx.x] = r < n ? dg[r] : 0.f;
    }
  };
  stage_rows<kAbPitch>(sK, kg, ldk, k0, m, 64, w);
  stage_rows<kAbPitch>(sV, vg, ldv, k0, m, 64, w);
  stage_q(0, 0);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  uint32_t kf[2][4], vf[2][4];
  ab_load_a(sK, warp * 16, lane, kf);
  ab_load_a(sV, warp * 16, lane, vf);
  float dka[4][4], dva[4][4];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    dka[t][0] = dka[t][1] = dka[t][2] = dka[t][3] = 0.f;
    dva[t][0] = dva[t][1] = dva[t][2] = dva[t][3] = 0.f;
  }
  const int q_tiles = (n + 63) / 64;
  for (int j = 0; j < q_tiles; ++j) {
    const int buf = j & 1;
    if (j + 1 < q_tiles) {
      stage_q(buf ^ 1, (j + 1) * 64);
      cp_async_commit();
    }
    float st[8][4], dpt[8][4];
    ab_mma_nt(st, kf, sQ[buf], lane);       // S^T  [16 keys x 64 queries]
    ab_mma_nt(dpt, vf, sdO[buf], lane);     // dP^T
    uint32_t pf[4][4], dsf[4][4];
    // dS^T / scale = P^T o (dP^T - delta); the scale is applied once to the accumulated dK
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const int qc = t * 8 + (lane & 3) * 2;
      const float2 l = *reinterpret_cast<const float2*>(&sLse[buf][qc]);      // negated log-sum-exp
      const float2 e = *reinterpret_cast<const float2*>(&sDelta[buf][qc]);
      const float p0 = fast_exp2(fmaf(st[t][0], scale_log2e, l.x)), p1 = fast_exp2(fmaf(st[t][1], scale_log2e, l.y));
      const float p2 = fast_exp2(fmaf(st[t][2], scale_log2e, l.x)), p3 = fast_exp2(fmaf(st[t][3], scale_log2e, l.y));
      pf[t >> 1][(t & 1) * 2] = pack_bf16x2(p0, p1);
      pf[t >> 1][(t & 1) * 2 + 1] = pack_bf16x2(p2, p3);
      dsf[t >> 1][(t & 1) * 2] = pack_bf16x2(p0 * (dpt[t][0] - e.x), p1 * (dpt[t][1] - e.y));
      dsf[t >> 1][(t & 1) * 2 + 1] = pack_bf16x2(p2 * (dpt[t][2] - e.x), p3 * (dpt[t][3] - e.y));
    }
    ab_mma_nn(dva, pf, sdO[buf], lane);     // dV += P^T dO
    ab_mma_nn(dka, dsf, sQ[buf], lane);     // dK += dS^T Q
    cp_async_wait<0>();
    __syncthreads();
  }
  const int r0 = k0 + warp * 16 + (lane >> 2), r1 = r0 + 8;
  __nv_bfloat16* dkg = dk + (long long)b * m * lddk + h * hp;
  __nv_bfloat16* dvg = dv + (long long)b * m * lddv + h * hp;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int col = t * 8 + (lane & 3) * 2;
    if (col < w) {
      if (r0 < m) {
        *reinterpret_cast<uint32_t*>(dkg + (long long)r0 * lddk + col) = pack_bf16x2(dka[t][0] * scale, dka[t][1] * scale);
        *reinterpret_cast<uint32_t*>(dvg + (long long)r0 * lddv + col) = pack_bf16x2(dva[t][0], dva[t][1]);
      }
      if (r1 < m) {
        *reinterpret_cast<uint32_t*>(dkg + (long long)r1 * lddk + col) = pack_bf16x2(dka[t][2] * scale, dka[t][3] * scale);
        *reinterpret_cast<uint32_t*>(dvg + (long long)r1 * lddv + col) = pack_bf16x2(dva[t][2], dva[t][3]);
      }
    }
  }
}

}  // namespace sdb

extern "C" int sdb_attention_bwd(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                 const void* o, int64_t ldo, const void* dO, int64_t lddo, void* dq, int64_t lddq, void* dk,
                                 int64_t lddk, void* dv, int64_t lddv, float* lse_scratch, float* delta_scratch, int lse_valid,
                                 int B, int heads, int n, int m, int d, int head_pitch, float scale, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(q && k && v && o && dO && dq && dk && dv && lse_scratch && delta_scratch, "attention_bwd: null pointer");
  SDB_CHECK_ARG(B > 0 && heads > 0 && n > 0 && m > 0 && d > 0 && d % 4 == 0 && head_pitch >= d && head_pitch % 4 == 0,
                "attention_bwd: bad shape");
  if (d > 32) {
    sdb_set_error("attention_bwd: head dim %d not covered (<= 32 is built: the UNet's heads are 10..20 wide)", d);
    return SDB_EUNSUPPORTED;
  }
  SDB_CHECK_ARG(ldq % 4 == 0 && ldk % 4 == 0 && ldv % 4 == 0 && ldo % 4 == 0 && lddo % 4 == 0 && lddq % 2 == 0 &&
                lddk % 2 == 0 && lddv % 2 == 0, "attention_bwd: row strides must be multiples of 4 elements");
  SDB_CHECK_ARG(heads <= 65535 && B <= 65535, "attention_bwd: grid too large");
  cudaStream_t st = sdb_stream(stream);
  const float sl2e = scale * 1.4426950408889634f;
  attn_bwd_dq_kernel<<<dim3((n + 63) / 64, heads, B), kAttThreads, 0, st>>>(
      (const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)k, ldk, (const __nv_bfloat16*)v, ldv, (const __nv_bfloat16*)o, ldo,
      (const __nv_bfloat16*)dO, lddo, (__nv_bfloat16*)dq, lddq, lse_scratch, delta_scratch, n, m, d, head_pitch, scale, sl2e,
      lse_valid);
  SDB_LAUNCH_OK();
  attn_bwd_dkv_kernel<<<dim3((m + 63) / 64, heads, B), kAttThreads, 0, st>>>(
      (const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)k, ldk, (const __nv_bfloat16*)v, ldv, (const __nv_bfloat16*)dO, lddo,
      lse_scratch, delta_scratch, (__nv_bfloat16*)dk, lddk, (__nv_bfloat16*)dv, lddv, n, m, d, head_pitch, scale, sl2e);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

// dK / dV only (the dQ half runs on the tcgen05 kernel of attention_bwd_umma.cu); lse / delta: fp32 [B, heads, n], filled.
extern "C" int sdb_attention_bwd_dkv(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                     const void* dO, int64_t lddo, void* dk, int64_t lddk, void* dv, int64_t lddv,
                                     const float* lse, const float* delta, int B, int heads, int n, int m, int d,
                                     int head_pitch, float scale, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(q && k && v && dO && dk && dv && lse && delta, "attention_bwd_dkv: null pointer");
  SDB_CHECK_ARG(B > 0 && heads > 0 && n > 0 && m > 0 && d > 0 && d <= 32 && d % 4 == 0 && head_pitch >= d && head_pitch % 4 == 0,
                "attention_bwd_dkv: bad shape");
  SDB_CHECK_ARG(ldq % 4 == 0 && ldk % 4 == 0 && ldv % 4 == 0 && lddo % 4 == 0 && lddk % 2 == 0 && lddv % 2 == 0,
                "attention_bwd_dkv: row strides must be multiples of 4 elements");
  SDB_CHECK_ARG(heads <= 65535 && B <= 65535, "attention_bwd_dkv: grid too large");
  const float sl2e = scale * 1.4426950408889634f;
  attn_bwd_dkv_kernel<<<dim3((m + 63) / 64, heads, B), kAttThreads, 0, sdb_stream(stream)>>>(
      (const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)k, ldk, (const __nv_bfloat16*)v, ldv, (const __nv_bfloat16*)dO, lddo,
      lse, delta, (__nv_bfloat16*)dk, lddk, (__nv_bfloat16*)dv, lddv, n, m, d, head_pitch, scale, sl2e);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
