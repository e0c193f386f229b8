// Backward of the short-context attention (K4c), dQ side: cross-attention over <= 128 context tokens (CLIP's 77).
// Same shape as the forward kernel: K and V of one (batch, head) staged once in shared memory, each warp walks 16-query
// blocks.  The whole context is one tile, so per block
//   S = Q K^T,  P = softmax(S c)  (row max / sum in registers -- no log-sum-exp needed from the forward),
//   dP = dO V^T,  delta = rowsum(P o dP)  (== rowsum(dO o O)),  dS = P o (dP - delta),  dQ = scale * dS K,
// all on mma.sync fragments; the row statistics (log2-domain lse, delta) are also written out for the dK / dV kernel.
// The generic flash backward (attention_bwd.cu) spent 0.53 ms per layer here (two passes, one 64-query CTA at a time);
// the bytes (Q, dO in; dQ out) need ~0.06 ms.
#include "common.cuh"

namespace sdb {

template <int KG>   // key groups of 16
__global__ void __launch_bounds__(kCtxThreads, 2)
attn_ctx_bwd_dq_kernel(const __nv_bfloat16* __restrict__ q, long long ldq, const __nv_bfloat16* __restrict__ k, long long ldk,
                       const __nv_bfloat16* __restrict__ v, long long ldv, const __nv_bfloat16* __restrict__ dO,
                       long long lddo, __nv_bfloat16* __restrict__ dq, long long lddq, float* __restrict__ lse,
                       float* __restrict__ delta, int n, int m, int d, float scale, float scale_log2e, int q_per_cta) {
  __shared__ __align__(16) __nv_bfloat16 sK[KG * 16 * kCtxPitch];
  __shared__ __align__(16) __nv_bfloat16 sV[KG * 16 * kCtxPitch];
  const int b = blockIdx.z, h = blockIdx.y, heads = gridDim.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const __nv_bfloat16* qg = q + (long long)b * n * ldq + h * 32;
  const __nv_bfloat16* dog = dO + (long long)b * n * lddo + h * d;
  const __nv_bfloat16* kg = k + (long long)b * m * ldk + h * 32;
  const __nv_bfloat16* vg = v + (long long)b * m * ldv + h * 32;
  const int q_begin = blockIdx.x * q_per_cta;
  const int q_end = q_begin + q_per_cta < n ? q_begin + q_per_cta : n;
  const int g = lane >> 2, c2 = (lane & 3) * 2;

  {
    constexpr int kChunks = KG * 16 * 4;
    constexpr int kPer = (kChunks + kCtxThreads - 1) / kCtxThreads;
    uint4 kk[kPer], vv[kPer];
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
      const int i = threadIdx.x + u * kCtxThreads;
      const int row = i >> 2, ch = i & 3;
      kk[u] = make_uint4(0u, 0u, 0u, 0u);
      vv[u] = kk[u];
      if (i < kChunks && row < m) {
        kk[u] = __ldg(reinterpret_cast<const uint4*>(kg + (long long)row * ldk + ch * 8));
        vv[u] = __ldg(reinterpret_cast<const uint4*>(vg + (long long)row * ldv + ch * 8));
      }
    }
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
      const int i = threadIdx.x + u * kCtxThreads;
      const int row = i >> 2, ch = i & 3;
      if (i < kChunks) {
        *reinterpret_cast<uint4*>(sK + row * kCtxPitch + ch * 8) = kk[u];
        *reinterpret_cast<uint4*>(sV + row * kCtxPitch + ch * 8) = vv[u];
      }
    }
  }
  __syncthreads();

  for (int q0 = q_begin + warp * 16; q0 < q_end; q0 += (kCtxThreads / 32) * 16) {
    const int r0 = q0 + g, r1 = r0 + 8;
    // A fragments: Q (head-padded, 32 columns) and dO (dense, d columns; zero beyond d -- V's ones column must not count)
    uint32_t qf[2][4], df[2][4];
    {
      const __nv_bfloat16* p0 = qg + (long long)r0 * ldq + c2;
      const __nv_bfloat16* p1 = qg + (long long)r1 * ldq + c2;
      const __nv_bfloat16* e0 = dog + (long long)r0 * lddo + c2;
      const __nv_bfloat16* e1 = dog + (long long)r1 * lddo + c2;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) {
        qf[ks][0] = r0 < q_end ? __ldg(reinterpret_cast<const uint32_t*>(p0 + ks * 16)) : 0u;
        qf[ks][1] = r1 < q_end ? __ldg(reinterpret_cast<const uint32_t*>(p1 + ks * 16)) : 0u;
        qf[ks][2] = r0 < q_end ? __ldg(reinterpret_cast<const uint32_t*>(p0 + ks * 16 + 8)) : 0u;
        qf[ks][3] = r1 < q_end ? __ldg(reinterpret_cast<const uint32_t*>(p1 + ks * 16 + 8)) : 0u;
        const bool lo = ks * 16 + c2 < d, hi = ks * 16 + 8 + c2 < d;
        df[ks][0] = (lo && r0 < q_end) ? __ldg(reinterpret_cast<const uint32_t*>(e0 + ks * 16)) : 0u;
        df[ks][1] = (lo && r1 < q_end) ? __ldg(reinterpret_cast<const uint32_t*>(e1 + ks * 16)) : 0u;
        df[ks][2] = (hi && r0 < q_end) ? __ldg(reinterpret_cast<const uint32_t*>(e0 + ks * 16 + 8)) : 0u;
        df[ks][3] = (hi && r1 < q_end) ? __ldg(reinterpret_cast<const uint32_t*>(e1 + ks * 16 + 8)) : 0u;
      }
    }
    // ---- S = Q K^T ---------------------------------------------------------------------------------------
    float s[KG * 2][4];
#pragma unroll
    for (int t = 0; t < KG * 2; ++t) { s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f; }
#pragma unroll
    for (int np = 0; np < KG; ++np) {
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) {
        const int key = np * 16 + (lane & 7) + ((lane >> 4) & 1) * 8;
        const int col = ks * 16 + ((lane >> 3) & 1) * 8;
        uint32_t b0, b1, b2, b3;
        ldsm_x4(smem_u32(sK + key * kCtxPitch + col), b0, b1, b2, b3);
        mma_bf16(s[2 * np], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], b0, b1);
        mma_bf16(s[2 * np + 1], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], b2, b3);
      }
    }
    if ((m & 15) != 0 || m < KG * 16) {
#pragma unroll
      for (int t = 0; t < KG * 2; ++t) {
        const int key = t * 8 + c2;
        if (key >= m) { s[t][0] = -INFINITY; s[t][2] = -INFINITY; }
        if (key + 1 >= m) { s[t][1] = -INFINITY; s[t][3] = -INFINITY; }
      }
    }
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int t = 0; t < KG * 2; ++t) {
      mx[0] = fmaxf(mx[0], fmaxf(s[t][0], s[t][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[t][2], s[t][3]));
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
    }
    const float msc0 = mx[0] * scale_log2e, msc1 = mx[1] * scale_log2e;
    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
    for (int t = 0; t < KG * 2; ++t) {
      s[t][0] = fast_exp2(fmaf(s[t][0], scale_log2e, -msc0));
      s[t][1] = fast_exp2(fmaf(s[t][1], scale_log2e, -msc0));
      s[t][2] = fast_exp2(fmaf(s[t][2], scale_log2e, -msc1));
      s[t][3] = fast_exp2(fmaf(s[t][3], scale_log2e, -msc1));
      sum0 += s[t][0] + s[t][1];
      sum1 += s[t][2] + s[t][3];
    }
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1);
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    const float inv0 = 1.0f / sum0, inv1 = 1.0f / sum1;
    // ---- dP = dO V^T (V read like K above: B[k = column][n = key]) ---------------------------------------------
    float dp[KG * 2][4];
#pragma unroll
    for (int t = 0; t < KG * 2; ++t) { dp[t][0] = dp[t][1] = dp[t][2] = dp[t][3] = 0.f; }
#pragma unroll
    for (int np = 0; np < KG; ++np) {
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) {
        const int key = np * 16 + (lane & 7) + ((lane >> 4) & 1) * 8;
        const int col = ks * 16 + ((lane >> 3) & 1) * 8;
        uint32_t b0, b1, b2, b3;
        ldsm_x4(smem_u32(sV + key * kCtxPitch + col), b0, b1, b2, b3);
        mma_bf16(dp[2 * np], df[ks][0], df[ks][1], df[ks][2], df[ks][3], b0, b1);
        mma_bf16(dp[2 * np + 1], df[ks][0], df[ks][1], df[ks][2], df[ks][3], b2, b3);
      }
    }
    // ---- delta = rowsum(P o dP), dS = P o (dP - delta) -----------------------------------------------------------
    float dl0 = 0.f, dl1 = 0.f;
#pragma unroll
    for (int t = 0; t < KG * 2; ++t) {
      s[t][0] *= inv0; s[t][1] *= inv0; s[t][2] *= inv1; s[t][3] *= inv1;
      dl0 = fmaf(s[t][0], dp[t][0], fmaf(s[t][1], dp[t][1], dl0));
      dl1 = fmaf(s[t][2], dp[t][2], fmaf(s[t][3], dp[t][3], dl1));
    }
    dl0 += __shfl_xor_sync(0xffffffffu, dl0, 1);
    dl0 += __shfl_xor_sync(0xffffffffu, dl0, 2);
    dl1 += __shfl_xor_sync(0xffffffffu, dl1, 1);
    dl1 += __shfl_xor_sync(0xffffffffu, dl1, 2);
    if ((lane & 3) == 0) {
      const long long o = ((long long)b * heads + h) * n;
      if (r0 < q_end) { lse[o + r0] = msc0 + log2f(sum0); delta[o + r0] = dl0; }
      if (r1 < q_end) { lse[o + r1] = msc1 + log2f(sum1); delta[o + r1] = dl1; }
    }
    uint32_t pf[KG][4];
#pragma unroll
    for (int t = 0; t < KG * 2; ++t) {
      pf[t >> 1][(t & 1) * 2] = pack_bf16x2(s[t][0] * (dp[t][0] - dl0), s[t][1] * (dp[t][1] - dl0));
      pf[t >> 1][(t & 1) * 2 + 1] = pack_bf16x2(s[t][2] * (dp[t][2] - dl1), s[t][3] * (dp[t][3] - dl1));
    }
    // ---- dQ = scale * dS K : 16 x 32 -----------------------------------------------------------------------------
    float o[4][4];
#pragma unroll
    for (int t = 0; t < 4; ++t) { o[t][0] = o[t][1] = o[t][2] = o[t][3] = 0.f; }
#pragma unroll
    for (int kk = 0; kk < KG; ++kk) {
#pragma unroll
      for (int dpi = 0; dpi < 2; ++dpi) {
        const int key = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int col = dpi * 16 + (lane >> 4) * 8;
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(smem_u32(sK + key * kCtxPitch + col), b0, b1, b2, b3);
        mma_bf16(o[2 * dpi], pf[kk][0], pf[kk][1], pf[kk][2], pf[kk][3], b0, b1);
        mma_bf16(o[2 * dpi + 1], pf[kk][0], pf[kk][1], pf[kk][2], pf[kk][3], b2, b3);
      }
    }
    __nv_bfloat16* og = dq + (long long)b * n * lddq + h * 32;
    const bool odd = lane & 1;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const uint32_t v0 = pack_bf16x2(o[t][0] * scale, o[t][1] * scale);   // row r0, cols c, c+1
      const uint32_t v1 = pack_bf16x2(o[t][2] * scale, o[t][3] * scale);   // row r1
      const uint32_t got = __shfl_xor_sync(0xffffffffu, odd ? v0 : v1, 1);
      const int col = t * 8 + (c2 & ~3);
      const int row = odd ? r1 : r0;
      if (row < q_end) {
        const uint2 w = odd ? make_uint2(got, v1) : make_uint2(v0, got);
        *reinterpret_cast<uint2*>(og + (long long)row * lddq + col) = w;
      }
    }
  }
}

template <int KG>
static int launch_attn_ctx_bwd_dq(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                                  const void* dO, long long lddo, void* dq, long long lddq, float* lse, float* delta, int B,
                                  int heads, int n, int m, int d, float scale, cudaStream_t st) {
  int q_per_cta = 1024;
  while (q_per_cta > 128 && (long long)((n + q_per_cta - 1) / q_per_cta) * heads * B < 4LL * sdb_num_sms()) q_per_cta >>= 1;
  dim3 grid((n + q_per_cta - 1) / q_per_cta, heads, B);
  attn_ctx_bwd_dq_kernel<KG><<<grid, kCtxThreads, 0, st>>>(
      (const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)k, ldk, (const __nv_bfloat16*)v, ldv, (const __nv_bfloat16*)dO, lddo,
      (__nv_bfloat16*)dq, lddq, lse, delta, n, m, d, scale, scale * 1.4426950408889634f, q_per_cta);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

}  // namespace sdb

// dQ of the head-padded attention over a short context (m <= 128), plus the row statistics the dK / dV kernel
// (sdb_attention_bwd_dkv) consumes: lse [B, heads, n] (log2 domain) and delta [B, heads, n].  q / k / v head-padded
// (32 columns per head), dO dense [B, n, heads*d], dq head-padded (every column written).  d % 4 == 0, d <= 20.
extern "C" int sdb_attention_ctx_bwd_dq(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                        const void* dO, int64_t lddo, void* dq, int64_t lddq, float* lse, float* delta, int B,
                                        int heads, int n, int m, int d, float scale, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(q && k && v && dO && dq && lse && delta && B > 0 && heads > 0 && n > 0 && m > 0, "attention_ctx_bwd_dq: bad args");
  SDB_CHECK_ARG(d > 0 && d <= 20 && d % 4 == 0, "attention_ctx_bwd_dq: head dim (%d) must be a multiple of 4, <= 20", d);
  if (m > 128) {
    sdb_set_error("attention_ctx_bwd_dq: context length %d > 128", m);
    return SDB_EUNSUPPORTED;
  }
  SDB_CHECK_ARG(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && lddq % 4 == 0 && lddo % 2 == 0 && ldq >= heads * 32 &&
                    ldk >= heads * 32 && ldv >= heads * 32 && lddq >= heads * 32 && lddo >= heads * d,
                "attention_ctx_bwd_dq: bad row strides");
  SDB_CHECK_ARG(((uintptr_t)q & 15) == 0 && ((uintptr_t)k & 15) == 0 && ((uintptr_t)v & 15) == 0 && ((uintptr_t)dO & 3) == 0 &&
                    ((uintptr_t)dq & 7) == 0,
                "attention_ctx_bwd_dq: q / k / v must be 16-byte aligned, dq 8-byte, dO 4-byte");
  SDB_CHECK_ARG(heads <= 65535 && B <= 65535, "attention_ctx_bwd_dq: grid too large");
  cudaStream_t st = sdb_stream(stream);
  switch ((m + 15) / 16) {
    case 1: return launch_attn_ctx_bwd_dq<1>(q, ldq, k, ldk, v, ldv, dO, lddo, dq, lddq, lse, delta, B, heads, n, m, d, scale, st);
    case 2: return launch_attn_ctx_bwd_dq<2>(q, ldq, k, ldk, v, ldv, dO, lddo, dq, lddq, lse, delta, B, heads, n, m, d, scale, st);
    case 3: return launch_attn_ctx_bwd_dq<3>(q, ldq, k, ldk, v, ldv, dO, lddo, dq, lddq, lse, delta, B, heads, n, m, d, scale, st);
    case 4: return launch_attn_ctx_bwd_dq<4>(q, ldq, k, ldk, v, ldv, dO, lddo, dq, lddq, lse, delta, B, heads, n, m, d, scale, st);
    case 5: return launch_attn_ctx_bwd_dq<5>(q, ldq, k, ldk, v, ldv, dO, lddo, dq, lddq, lse, delta, B, heads, n, m, d, scale, st);
    case 6: return launch_attn_ctx_bwd_dq<6>(q, ldq, k, ldk, v, ldv, dO, lddo, dq, lddq, lse, delta, B, heads, n, m, d, scale, st);
    case 7: return launch_attn_ctx_bwd_dq<7>(q, ldq, k, ldk, v, ldv, dO, lddo, dq, lddq, lse, delta, B, heads, n, m, d, scale, st);
    default: return launch_attn_ctx_bwd_dq<8>(q, ldq, k, ldk, v, ldv, dO, lddo, dq, lddq, lse, delta, B, heads, n, m, d, scale, st);
  }
}
