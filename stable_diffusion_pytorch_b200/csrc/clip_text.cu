// CLIP text-encoder pieces that the UNet kernels do not already cover (the reference wraps
// transformers.CLIPTextModel, models/clip_model.py:56-96 -- modeling_clip.py CLIPTextEmbeddings / CLIPAttention):
//   embed_tokens_kernel          token_embedding[ids] + position_embedding  -> [B*L, C] in the compute dtype
//   attention_causal_small       causal multi-head self-attention over a short sequence (77 tokens): K and V of one
//                                (batch, head) live in shared memory as fp32, one warp per query row, exact softmax.
// The twelve layers' projections / MLPs run on the tcgen05 GEMM kernel, LayerNorm on add_layernorm_kernel.
#include "common.cuh"

namespace sdb {

template <typename TO>
__global__ void embed_tokens_kernel(const long long* __restrict__ ids, const float* __restrict__ tok,
                                    const float* __restrict__ pos, TO* __restrict__ out, int L, int Cc, int vocab,
                                    long long rows) {
  const long long row = blockIdx.x;
  if (row >= rows) return;
  long long id = ids[row];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);     // ids come from the tokenizer; clamp instead of faulting
  const float* t = tok + id * Cc;
  const float* p = pos + (row % L) * Cc;
  for (int c = threadIdx.x; c < Cc; c += blockDim.x) {
    const float v = t[c] + p[c];
    if constexpr (sizeof(TO) == 4) out[row * Cc + c] = v;
    else out[row * Cc + c] = __float2bfloat16(v);
  }
}

constexpr int kCaMaxN = 80, kCaMaxD = 64;   // K + V as fp32 fit the 48 KB static shared memory
constexpr int kCaScoreT = (kCaMaxN + 31) / 32;

template <typename T>
__device__ __forceinline__ float ca_ld(const T* p) {
  if constexpr (sizeof(T) == 4) return *p;
  else return __bfloat162float(*p);
}

template <typename T>
__global__ void __launch_bounds__(128) attention_causal_small_kernel(const T* __restrict__ q, long long ldq,
                                                                     const T* __restrict__ k, long long ldk,
                                                                     const T* __restrict__ v, long long ldv,
                                                                     T* __restrict__ out, long long ldo, int n, int d,
                                                                     float scale) {
  __shared__ float Ks[kCaMaxN][kCaMaxD + 1];
  __shared__ float Vs[kCaMaxN][kCaMaxD + 1];
  __shared__ float Qs[4][kCaMaxD];
  __shared__ float Ps[4][kCaScoreT * 32];
  const int b = blockIdx.y, h = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int e = threadIdx.x; e < n * d; e += 128) {
    const int j = e / d, c = e - j * d;
    Ks[j][c] = ca_ld(k + ((long long)b * n + j) * ldk + h * d + c);
    Vs[j][c] = ca_ld(v + ((long long)b * n + j) * ldv + h * d + c);
  }
  __syncthreads();
  for (int i = warp; i < n; i += 4) {
    for (int c = lane; c < d; c += 32) Qs[warp][c] = ca_ld(q + ((long long)b * n + i) * ldq + h * d + c) * scale;
    __syncwarp();
    float s[kCaScoreT];
    float mx = -INFINITY;
#pragma unroll
    for (int t = 0; t < kCaScoreT; ++t) {
      const int j = lane + 32 * t;
      float acc = -INFINITY;
      if (j <= i) {            // causal: key j visible to query i iff j <= i
        acc = 0.f;
        for (int c = 0; c < d; ++c) acc = fmaf(Qs[warp][c], Ks[j][c], acc);
      }
      s[t] = acc;
      mx = fmaxf(mx, acc);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
#pragma unroll
    for (int t = 0; t < kCaScoreT; ++t) {
      const int j = lane + 32 * t;
      const float pj = (j <= i) ? expf(s[t] - mx) : 0.f;
      Ps[warp][j] = pj;
      sum += pj;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    __syncwarp();
    const float inv = 1.f / sum;
    for (int c = lane; c < d; c += 32) {
      float acc = 0.f;
      for (int j = 0; j <= i; ++j) acc = fmaf(Ps[warp][j], Vs[j][c], acc);
      const float r = acc * inv;
      T* o = out + ((long long)b * n + i) * ldo + h * d + c;
      if constexpr (sizeof(T) == 4) *o = r;
      else *o = __float2bfloat16(r);
    }
    __syncwarp();
  }
}

}  // namespace sdb

extern "C" int sdb_embed_tokens(const int64_t* ids, const float* tok, const float* pos, void* out, int out_dtype,
                                int B, int L, int C, int vocab, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(ids && tok && pos && out && B > 0 && L > 0 && C > 0 && vocab > 0, "embed_tokens: bad args");
  const long long rows = (long long)B * L;
  if (out_dtype == SDB_F32)
    embed_tokens_kernel<float><<<(unsigned)rows, 128, 0, sdb_stream(stream)>>>((const long long*)ids, tok, pos, (float*)out, L, C, vocab, rows);
  else
    embed_tokens_kernel<__nv_bfloat16><<<(unsigned)rows, 128, 0, sdb_stream(stream)>>>((const long long*)ids, tok, pos, (__nv_bfloat16*)out, L, C, vocab, rows);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_attention_causal_small(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                                          int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int d,
                                          float scale, int dtype, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(q && k && v && out && B > 0 && heads > 0 && n > 0 && d > 0, "attention_causal_small: bad args");
  if (n > kCaMaxN || d > kCaMaxD) {
    sdb_set_error("attention_causal_small: n (%d) <= %d and d (%d) <= %d are built (CLIP: 77 tokens, 64-wide heads)",
                  n, kCaMaxN, d, kCaMaxD);
    return SDB_EUNSUPPORTED;
  }
  SDB_CHECK_ARG(B <= 65535, "attention_causal_small: grid too large");
  dim3 grid(heads, B);
  if (dtype == SDB_F32)
    attention_causal_small_kernel<float><<<grid, 128, 0, sdb_stream(stream)>>>(
        (const float*)q, ldq, (const float*)k, ldk, (const float*)v, ldv, (float*)out, ldo, n, d, scale);
  else
    attention_causal_small_kernel<__nv_bfloat16><<<grid, 128, 0, sdb_stream(stream)>>>(
        (const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)k, ldk, (const __nv_bfloat16*)v, ldv, (__nv_bfloat16*)out, ldo, n, d, scale);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
