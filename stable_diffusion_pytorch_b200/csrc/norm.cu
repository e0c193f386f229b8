// K2 GroupNorm (statistics + normalise/SiLU over a virtual channel concat), K5 residual-add +
// LayerNorm, and the layout / elementwise helpers.  All HBM-bound: 16-byte vector accesses,
// channels-last so a warp reads consecutive channels of consecutive pixels.
#include "common.cuh"

namespace sdb {

// ---- 8-channel vector load/store in fp32 registers --------------------------------------------
template <typename T> struct Vec8;
template <> struct Vec8<__nv_bfloat16> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
    float2 a = unpack_bf16x2(q.x), b = unpack_bf16x2(q.y), c = unpack_bf16x2(q.z), d = unpack_bf16x2(q.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
    *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                              pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
};
template <> struct Vec8<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[8]) {
    float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[8]) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
};

constexpr int kGnThreads = 256;
constexpr int kGnMaxChunks = 64;

// ---- GroupNorm statistics, stage 1: per (image, pixel chunk, channel) sum / sum of squares -----
// grid (chunks, B); thread (r, vec): vec = 8-channel vector of the virtual concat, r = pixel lane.
template <typename T>
__global__ void __launch_bounds__(kGnThreads)
gn_partial_kernel(const T* __restrict__ x1, int C1, const T* __restrict__ x2, int C2, long long hw,
                  int chunk_px, float* __restrict__ scratch /* [B, chunks, C, 2] */) {
  extern __shared__ float sm[];  // [R][2*C]
  const int C = C1 + C2, V = C >> 3;
  const int R = kGnThreads / V;
  const int b = blockIdx.y, chunk = blockIdx.x, chunks = gridDim.x;
  const int vec = threadIdx.x % V, r = threadIdx.x / V;
  float s[8], ss[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.f; ss[j] = 0.f; }
  const long long p0 = (long long)chunk * chunk_px;
  const long long p1 = p0 + chunk_px < hw ? p0 + chunk_px : hw;
  if (r < R) {
    const int c = vec * 8;
    const T* base;
    long long pitch;
    if (c < C1) { base = x1 + (long long)b * hw * C1 + c; pitch = C1; }
    else { base = x2 + (long long)b * hw * C2 + (c - C1); pitch = C2; }
    // four independent 16-byte loads in flight per thread: the kernel is a pure HBM stream
    long long px = p0 + r;
    for (; px + 3LL * R < p1; px += 4LL * R) {
      float v0[8], v1[8], v2[8], v3[8];
      Vec8<T>::load(base + px * pitch, v0);
      Vec8<T>::load(base + (px + R) * pitch, v1);
      Vec8<T>::load(base + (px + 2LL * R) * pitch, v2);
      Vec8<T>::load(base + (px + 3LL * R) * pitch, v3);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += (v0[j] + v1[j]) + (v2[j] + v3[j]);
        ss[j] = fmaf(v0[j], v0[j], ss[j]);
        ss[j] = fmaf(v1[j], v1[j], ss[j]);
        ss[j] = fmaf(v2[j], v2[j], ss[j]);
        ss[j] = fmaf(v3[j], v3[j], ss[j]);
      }
    }
    for (; px < p1; px += R) {
      float v[8];
      Vec8<T>::load(base + px * pitch, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) { s[j] += v[j]; ss[j] = fmaf(v[j], v[j], ss[j]); }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sm[(r * C + c + j) * 2] = s[j];
      sm[(r * C + c + j) * 2 + 1] = ss[j];
    }
  }
  __syncthreads();
  float* dst = scratch + ((long long)b * chunks + chunk) * C * 2;
  for (int i = threadIdx.x; i < 2 * C; i += kGnThreads) {
    float acc = 0.f;
    for (int rr = 0; rr < R; ++rr) acc += sm[rr * 2 * C + i];  // fixed order: deterministic
    dst[i] = acc;
  }
}

// ---- stage 2: per (image, group) mean / rstd ---------------------------------------------------
__global__ void __launch_bounds__(128)
gn_finalize_kernel(const float* __restrict__ scratch, int chunks, int C, int groups, long long hw,
                   float eps, float* __restrict__ stats /* [B, groups, 2] */) {
  const int g = blockIdx.x, b = blockIdx.y, cpg = C / groups;
  float s = 0.f, ss = 0.f;
  const int n = chunks * cpg;
  for (int i = threadIdx.x; i < n; i += 128) {
    const int ch = i / cpg, c = g * cpg + (i - ch * cpg);
    const float* src = scratch + (((long long)b * chunks + ch) * C + c) * 2;
    s += src[0];
    ss += src[1];
  }
  __shared__ float red[2][4];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s; red[1][threadIdx.x >> 5] = ss; }
  __syncthreads();
  if (threadIdx.x == 0) {
    s = red[0][0] + red[0][1] + red[0][2] + red[0][3];
    ss = red[1][0] + red[1][1] + red[1][2] + red[1][3];
    const float inv = 1.0f / ((float)cpg * (float)hw);
    const float mean = s * inv;
    float var = ss * inv - mean * mean;
    var = var > 0.f ? var : 0.f;
    stats[((long long)b * groups + g) * 2] = mean;
    stats[((long long)b * groups + g) * 2 + 1] = rsqrtf(var + eps);
  }
}

// ---- normalise (+SiLU), writing a dense NHWC tensor with C1+C2 channels -----------------------
// grid (pixel chunks, B); thread (r, vec) owns one 8-channel vector of the virtual concat -- its scale and
// shift live in registers -- and walks the chunk's pixels with four 16-byte loads in flight.
template <typename T, typename OutT>
__global__ void __launch_bounds__(kGnThreads)
gn_apply_kernel(const T* __restrict__ x1, int C1, const T* __restrict__ x2, int C2,
                const float* __restrict__ stats, const float* __restrict__ gamma,
                const float* __restrict__ beta, int act, OutT* __restrict__ out, long long hw,
                int groups, int chunk_px) {
  const int C = C1 + C2, V = C >> 3, cpg = C / groups;
  const int R = kGnThreads / V;
  const int b = blockIdx.y;
  const int vec = threadIdx.x % V, r = threadIdx.x / V;
  if (r >= R) return;
  const int c = vec * 8;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c + j) / cpg;
    const float mean = stats[((long long)b * groups + g) * 2];
    const float rstd = stats[((long long)b * groups + g) * 2 + 1];
    sc[j] = rstd * gamma[c + j];
    sh[j] = beta[c + j] - mean * sc[j];
  }
  const T* base;
  long long pitch;
  if (c < C1) { base = x1 + (long long)b * hw * C1 + c; pitch = C1; }
  else { base = x2 + (long long)b * hw * C2 + (c - C1); pitch = C2; }
  OutT* obase = out + (long long)b * hw * C + c;
  const long long p0 = (long long)blockIdx.x * chunk_px;
  const long long p1 = p0 + chunk_px < hw ? p0 + chunk_px : hw;
  auto finish = [&](float (&v)[8], long long px) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float y = fmaf(v[j], sc[j], sh[j]);
      v[j] = act == SDB_ACT_SILU ? silu_f(y) : y;
    }
    Vec8<OutT>::store(obase + px * C, v);
  };
  long long px = p0 + r;
  for (; px + 3LL * R < p1; px += 4LL * R) {
    float v0[8], v1[8], v2[8], v3[8];
    Vec8<T>::load(base + px * pitch, v0);
    Vec8<T>::load(base + (px + R) * pitch, v1);
    Vec8<T>::load(base + (px + 2LL * R) * pitch, v2);
    Vec8<T>::load(base + (px + 3LL * R) * pitch, v3);
    finish(v0, px);
    finish(v1, px + R);
    finish(v2, px + 2LL * R);
    finish(v3, px + 3LL * R);
  }
  for (; px < p1; px += R) {
    float v[8];
    Vec8<T>::load(base + px * pitch, v);
    finish(v, px);
  }
}

// ---- K5: out = LayerNorm(x (+ res)) over C, two-pass in registers ------------------------------
// kL lanes share a row (32 / kL rows per warp), each holding kVPL 8-channel vectors: every lane has
// kVPL (x2 with a residual) independent 16-byte loads in flight and the reductions are log2(kL) shuffles.
template <typename T, int kL, int kVPL>
__global__ void __launch_bounds__(256)
add_layernorm_kernel(const T* __restrict__ x, const T* __restrict__ res,
                     const float* __restrict__ gamma, const float* __restrict__ beta,
                     T* __restrict__ out, long long rows, int C, float eps) {
  constexpr int kRowsPerWarp = 32 / kL;
  const int lane = threadIdx.x & 31;
  const int sub = lane % kL;
  const long long row = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * kRowsPerWarp + lane / kL;
  const bool live = row < rows;
  const int V = C >> 3;
  float v[kVPL][8];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < kVPL; ++k) {
    const int vi = sub + kL * k;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[k][j] = 0.f;
    if (live && vi < V) Vec8<T>::load(x + row * C + vi * 8, v[k]);
  }
  if (res) {
#pragma unroll
    for (int k = 0; k < kVPL; ++k) {
      const int vi = sub + kL * k;
      if (live && vi < V) {
        float r[8];
        Vec8<T>::load(res + row * C + vi * 8, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[k][j] += r[j];
      }
    }
  }
#pragma unroll
  for (int k = 0; k < kVPL; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) s += v[k][j];
#pragma unroll
  for (int o = kL >> 1; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / (float)C;
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < kVPL; ++k) {
    if (sub + kL * k < V) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float d = v[k][j] - mean; ss = fmaf(d, d, ss); }
    }
  }
#pragma unroll
  for (int o = kL >> 1; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  const float rstd = rsqrtf(ss / (float)C + eps);
#pragma unroll
  for (int k = 0; k < kVPL; ++k) {
    const int vi = sub + kL * k;
    if (live && vi < V) {
      float g[8], bt[8], o[8];
      Vec8<float>::load(gamma + vi * 8, g);
      Vec8<float>::load(beta + vi * 8, bt);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fmaf((v[k][j] - mean) * rstd, g[j], bt[j]);
      Vec8<T>::store(out + row * C + vi * 8, o);
    }
  }
}

// ---- batched transpose [B, R, Cc] -> [B, Cc, R] with dtype conversion -------------------------
template <typename S, typename D>
__device__ __forceinline__ D cvt(S v);
template <> __device__ __forceinline__ float cvt<float, float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 cvt<float, __nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ float cvt<__nv_bfloat16, float>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ __nv_bfloat16 cvt<__nv_bfloat16, __nv_bfloat16>(__nv_bfloat16 v) { return v; }

template <typename S, typename D>
__global__ void transpose_kernel(const S* __restrict__ src, D* __restrict__ dst, int R, int Cc) {
  __shared__ float tile[32][33];
  const long long boff = (long long)blockIdx.z * R * Cc;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    if (r < R && c < Cc) tile[i][threadIdx.x] = cvt<S, float>(src[boff + (long long)r * Cc + c]);
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (r < R && c < Cc) dst[boff + (long long)c * R + r] = cvt<float, D>(tile[threadIdx.x][i]);
  }
}

// [B, HW, Cp] (channels-last, Cp >= C padded channels, fp32) -> [B, C, HW]: the output head runs on the tensor
// cores with its 3..8 output channels padded to 16; this drops the padding and restores NCHW.  Thread = pixel:
// 16-byte reads of its channel vector, plane-wise coalesced writes.
template <typename D>
__global__ void nhwc_slice_to_nchw_kernel(const float* __restrict__ src, D* __restrict__ dst, long long hw,
                                          int Cp, int C, long long total /* B*hw */) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / hw, px = i - b * hw;
    const float4* s = reinterpret_cast<const float4*>(src + i * Cp);
    for (int c4 = 0; c4 * 4 < C; ++c4) {
      const float4 v = __ldg(s + c4);
      const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (c4 * 4 + j < C) dst[(b * C + c4 * 4 + j) * hw + px] = cvt<float, D>(vv[j]);
    }
  }
}

template <typename T>
__global__ void upsample2x_kernel(const T* __restrict__ x, T* __restrict__ out, int H, int W, int C,
                                  long long total_vec) {
  const int V = C >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec;
       i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % V);
    long long t = i / V;
    const int wo = (int)(t % (2 * W)); t /= (2 * W);
    const int ho = (int)(t % (2 * H));
    const long long b = t / (2 * H);
    const uint4* s = reinterpret_cast<const uint4*>(x + (((b * H + (ho >> 1)) * W + (wo >> 1)) * C + v * 8));
    uint4* d = reinterpret_cast<uint4*>(out + (((b * 2 * H + ho) * 2 * W + wo) * C + v * 8));
    if constexpr (sizeof(T) == 2) { *d = __ldg(s); }
    else { d[0] = __ldg(s); d[1] = __ldg(s + 1); }
  }
}

template <typename S, typename D>
__global__ void silu_cast_kernel(const S* __restrict__ x, D* __restrict__ out, long long n, int do_silu) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float v = cvt<S, float>(x[i]);
    out[i] = cvt<float, D>(do_silu ? silu_f(v) : v);
  }
}

template <typename T>
__global__ void to_uint8_kernel(const T* __restrict__ x, uint8_t* __restrict__ out, int C, long long hw,
                                long long total) {
  // utils/prepare_dataset.py:145-150: (x+1)/2, clip to [0,1], *255, truncate to uint8
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long px = (i / C) % hw, b = i / (C * hw);
    float v = cvt<T, float>(x[(b * C + c) * hw + px]);
    v = (v + 1.0f) / 2.0f;
    v = fminf(fmaxf(v, 0.0f), 1.0f);
    out[i] = (uint8_t)(v * 255.0f);
  }
}

static inline int ew_grid(long long n, int threads) {
  long long want = (n + threads - 1) / threads;
  long long cap = (long long)sdb_num_sms() * 16;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace sdb

using namespace sdb;

extern "C" int sdb_groupnorm_chunks(int64_t hw) {
  int64_t c = (hw + 63) / 64;
  if (c < 1) c = 1;
  if (c > kGnMaxChunks) c = kGnMaxChunks;
  return (int)c;
}

extern "C" int sdb_groupnorm_stats(const void* x1, int C1, const void* x2, int C2, int dtype, int B,
                                   int64_t hw, int groups, float eps, float* scratch, float* stats,
                                   void* stream) {
  if (!x2) C2 = 0;
  const int C = C1 + C2;
  SDB_CHECK_ARG(x1 && scratch && stats && B > 0 && hw > 0, "groupnorm_stats: bad args");
  SDB_CHECK_ARG(C1 % 8 == 0 && C2 % 8 == 0, "groupnorm: C1 (%d), C2 (%d) must be multiples of 8", C1, C2);
  SDB_CHECK_ARG(groups > 0 && C % groups == 0, "groupnorm: channels (%d) must be divisible by num_groups (%d)", C, groups);
  const int V = C / 8;
  SDB_CHECK_ARG(V <= kGnThreads, "groupnorm: C (%d) too large", C);
  const int chunks = sdb_groupnorm_chunks(hw);
  const int chunk_px = (int)((hw + chunks - 1) / chunks);
  const int R = kGnThreads / V;
  const size_t smem = (size_t)R * 2 * C * sizeof(float);
  dim3 grid(chunks, B);
  if (dtype == SDB_BF16)
    gn_partial_kernel<__nv_bfloat16><<<grid, kGnThreads, smem, sdb_stream(stream)>>>(
        (const __nv_bfloat16*)x1, C1, (const __nv_bfloat16*)x2, C2, hw, chunk_px, scratch);
  else
    gn_partial_kernel<float><<<grid, kGnThreads, smem, sdb_stream(stream)>>>(
        (const float*)x1, C1, (const float*)x2, C2, hw, chunk_px, scratch);
  SDB_LAUNCH_OK();
  gn_finalize_kernel<<<dim3(groups, B), 128, 0, sdb_stream(stream)>>>(scratch, chunks, C, groups, hw, eps, stats);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

// ---- allocation-free variant of sdb_groupnorm_stats ---------------------------------------------
// The partial sums live in a file-scope device array instead of a caller buffer.  Every GPU that loads
// the library holds the 4 MiB until exit.  The host-side name `g_gn_static_scratch` is only a stand-in:
// its device address is looked up once per device with cudaGetSymbolAddress and cached here.
constexpr size_t kGnStaticScratchFloats = SDB_GN_STATIC_SCRATCH_FLOATS;
static __device__ __align__(256) float g_gn_static_scratch[kGnStaticScratchFloats];
constexpr int kGnStaticMaxDevices = 64;
static float* g_gn_static_ptr[kGnStaticMaxDevices] = {nullptr};

static int gn_static_scratch_ptr(float** out) {
  int dev = -1;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  SDB_CHECK_ARG(dev >= 0 && dev < kGnStaticMaxDevices, "groupnorm_stats_static: device ordinal %d out of range", dev);
  if (!g_gn_static_ptr[dev]) {
    float* p = nullptr;
    SDB_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_gn_static_scratch));
    g_gn_static_ptr[dev] = p;
  }
  *out = g_gn_static_ptr[dev];
  return SDB_OK;
}

extern "C" int sdb_groupnorm_stats_static(const void* x1, int C1, const void* x2, int C2, int dtype, int B,
                                          int64_t hw, int groups, float eps, float* stats, void* stream) {
  if (!x2) C2 = 0;
  SDB_CHECK_ARG(B > 0 && hw > 0 && C1 > 0 && C2 >= 0, "groupnorm_stats_static: bad args");
  const size_t need = (size_t)B * (size_t)sdb_groupnorm_chunks(hw) * (size_t)(C1 + C2) * 2;
  SDB_CHECK_ARG(need <= kGnStaticScratchFloats,
                "groupnorm_stats_static: B * chunks * C * 2 = %zu floats exceeds the static scratch of %zu floats "
                "(SDB_GN_STATIC_SCRATCH_FLOATS); use sdb_groupnorm_stats with a caller buffer",
                need, kGnStaticScratchFloats);
  float* scratch = nullptr;
  const int rc = gn_static_scratch_ptr(&scratch);
  if (rc != SDB_OK) return rc;
  // every element of scratch[0, need) is overwritten by gn_partial_kernel before gn_finalize_kernel reads it, so
  // nothing has to start at zero; what the previous call left beyond `need` is never read.
  return sdb_groupnorm_stats(x1, C1, x2, C2, dtype, B, hw, groups, eps, scratch, stats, stream);
}

extern "C" int sdb_groupnorm_apply(const void* x1, int C1, const void* x2, int C2, int dtype,
                                   const float* stats, const float* gamma, const float* beta, int act,
                                   void* out, int out_dtype, int B, int64_t hw, int groups,
                                   void* stream) {
  if (!x2) C2 = 0;
  const int C = C1 + C2;
  SDB_CHECK_ARG(x1 && stats && gamma && beta && out && B > 0 && hw > 0, "groupnorm_apply: bad args");
  SDB_CHECK_ARG(C1 % 8 == 0 && C2 % 8 == 0 && groups > 0 && C % groups == 0, "groupnorm_apply: bad channels");
  const int V = C / 8;
  SDB_CHECK_ARG(V <= kGnThreads, "groupnorm_apply: C (%d) too large", C);
  // pixel chunks: enough blocks to fill the machine a few times over, >= 4 * R pixels each
  const int R = kGnThreads / V;
  long long chunks = ((long long)sdb_num_sms() * 8 + B - 1) / B;
  const long long max_chunks = (hw + 4LL * R - 1) / (4LL * R);
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  const int chunk_px = (int)((hw + chunks - 1) / chunks);
  chunks = (hw + chunk_px - 1) / chunk_px;
  dim3 grid((unsigned)chunks, B);
  cudaStream_t st = sdb_stream(stream);
#define SDB_GN_APPLY(T, O)                                                                       \
  gn_apply_kernel<T, O><<<grid, kGnThreads, 0, st>>>((const T*)x1, C1, (const T*)x2, C2, stats, gamma, \
                                                     beta, act, (O*)out, hw, groups, chunk_px)
  if (dtype == SDB_BF16 && out_dtype == SDB_BF16) SDB_GN_APPLY(__nv_bfloat16, __nv_bfloat16);
  else if (dtype == SDB_BF16) SDB_GN_APPLY(__nv_bfloat16, float);
  else if (out_dtype == SDB_BF16) SDB_GN_APPLY(float, __nv_bfloat16);
  else SDB_GN_APPLY(float, float);
#undef SDB_GN_APPLY
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_add_layernorm(const void* x, const void* res, const float* gamma, const float* beta,
                                 void* out, int64_t rows, int C, float eps, int dtype, void* stream) {
  SDB_CHECK_ARG(x && gamma && beta && out && rows > 0, "add_layernorm: bad args");
  SDB_CHECK_ARG(C % 8 == 0 && C <= 1024, "add_layernorm: C (%d) must be a multiple of 8, <= 1024", C);
  // lanes per row: the smallest power of two >= 4 that keeps <= 5 vectors per lane
  const int V = C / 8;
  int L = 4;
  while (L < 32 && (V + L - 1) / L > 5) L *= 2;
  const int vpl = (V + L - 1) / L;
  const long long rows_per_block = 8LL * (32 / L);
  const unsigned grid = (unsigned)((rows + rows_per_block - 1) / rows_per_block);
  cudaStream_t st = sdb_stream(stream);
#define SDB_LN(T, LL, VP)                                                                          \
  add_layernorm_kernel<T, LL, VP><<<grid, 256, 0, st>>>((const T*)x, (const T*)res, gamma, beta, (T*)out, rows, C, eps)
#define SDB_LN_T(T)                                                                                \
  do {                                                                                             \
    if (L == 4) { if (vpl <= 2) SDB_LN(T, 4, 2); else if (vpl <= 4) SDB_LN(T, 4, 4); else SDB_LN(T, 4, 5); }       \
    else if (L == 8) { if (vpl <= 3) SDB_LN(T, 8, 3); else SDB_LN(T, 8, 5); }                      \
    else if (L == 16) { if (vpl <= 3) SDB_LN(T, 16, 3); else SDB_LN(T, 16, 5); }                   \
    else { if (vpl <= 3) SDB_LN(T, 32, 3); else SDB_LN(T, 32, 4); }                                \
  } while (0)
  if (dtype == SDB_BF16) SDB_LN_T(__nv_bfloat16);
  else SDB_LN_T(float);
#undef SDB_LN_T
#undef SDB_LN
  SDB_LAUNCH_OK();
  return SDB_OK;
}

// ---- row softmax of a materialised score matrix (the wide-head attention route, d > 128) ------------------------
// modules/transformer.py:101-111: sim * scale, softmax over the keys.  One warp per row; columns >= valid (the key
// padding that makes the row a multiple of 16 for the GEMMs either side) come out as exact zeros.
template <typename OutT>
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float* __restrict__ x, long long ldx, OutT* __restrict__ out, long long ldo, long long rows,
                    int cols, int valid, float scale_log2) {
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* xr = x + row * ldx;
  float mx = -INFINITY;
  for (int c = lane; c < valid; c += 32) mx = fmaxf(mx, xr[c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
  for (int c = lane; c < valid; c += 32) sum += exp2f((xr[c] - mx) * scale_log2);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float inv = 1.0f / sum;
  OutT* orow = out + row * ldo;
  for (int c = lane; c < cols; c += 32) {
    const float p = c < valid ? exp2f((xr[c] - mx) * scale_log2) * inv : 0.f;
    if constexpr (sizeof(OutT) == 2) orow[c] = __float2bfloat16(p);
    else orow[c] = p;
  }
}

extern "C" int sdb_softmax_rows(const float* x, int64_t ldx, void* out, int64_t ldo, int out_dtype, int64_t rows,
                                int cols, int valid_cols, float scale, void* stream) {
  SDB_CHECK_ARG(x && out && rows > 0 && cols > 0 && valid_cols > 0 && valid_cols <= cols && ldx >= cols && ldo >= cols,
                "softmax_rows: bad args");
  const unsigned grid = (unsigned)((rows + 7) / 8);
  const float sl2 = scale * 1.4426950408889634f;
  if (out_dtype == SDB_BF16)
    softmax_rows_kernel<__nv_bfloat16><<<grid, 256, 0, sdb_stream(stream)>>>(x, ldx, (__nv_bfloat16*)out, ldo, rows,
                                                                            cols, valid_cols, sl2);
  else
    softmax_rows_kernel<float><<<grid, 256, 0, sdb_stream(stream)>>>(x, ldx, (float*)out, ldo, rows, cols,
                                                                    valid_cols, sl2);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

static int launch_transpose(const void* src, int sdt, void* dst, int ddt, int Bn, int R, int Cc, cudaStream_t st) {
  dim3 grid((Cc + 31) / 32, (R + 31) / 32, Bn), block(32, 8);
  SDB_CHECK_ARG(grid.y <= 65535 && grid.z <= 65535, "transpose: dims too large");
  if (sdt == SDB_F32 && ddt == SDB_F32) transpose_kernel<float, float><<<grid, block, 0, st>>>((const float*)src, (float*)dst, R, Cc);
  else if (sdt == SDB_F32) transpose_kernel<float, __nv_bfloat16><<<grid, block, 0, st>>>((const float*)src, (__nv_bfloat16*)dst, R, Cc);
  else if (ddt == SDB_F32) transpose_kernel<__nv_bfloat16, float><<<grid, block, 0, st>>>((const __nv_bfloat16*)src, (float*)dst, R, Cc);
  else transpose_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, block, 0, st>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst, R, Cc);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_nchw_to_nhwc(const void* src, int src_dtype, void* dst, int dst_dtype, int B, int C,
                                int H, int W, void* stream) {
  SDB_CHECK_ARG(src && dst && B > 0 && C > 0 && H > 0 && W > 0, "nchw_to_nhwc: bad args");
  return launch_transpose(src, src_dtype, dst, dst_dtype, B, C, H * W, sdb_stream(stream));
}
extern "C" int sdb_nhwc_to_nchw(const void* src, int src_dtype, void* dst, int dst_dtype, int B, int C,
                                int H, int W, void* stream) {
  SDB_CHECK_ARG(src && dst && B > 0 && C > 0 && H > 0 && W > 0, "nhwc_to_nchw: bad args");
  return launch_transpose(src, src_dtype, dst, dst_dtype, B, H * W, C, sdb_stream(stream));
}

extern "C" int sdb_nhwc_slice_to_nchw(const float* src, int Cp, void* dst, int dst_dtype, int B, int C, int H,
                                      int W, void* stream) {
  SDB_CHECK_ARG(src && dst && B > 0 && H > 0 && W > 0 && C > 0 && C <= Cp && Cp % 4 == 0, "nhwc_slice_to_nchw: bad args");
  const long long hw = (long long)H * W, total = hw * B;
  if (dst_dtype == SDB_BF16)
    nhwc_slice_to_nchw_kernel<__nv_bfloat16><<<ew_grid(total, 256), 256, 0, sdb_stream(stream)>>>(src, (__nv_bfloat16*)dst, hw, Cp, C, total);
  else
    nhwc_slice_to_nchw_kernel<float><<<ew_grid(total, 256), 256, 0, sdb_stream(stream)>>>(src, (float*)dst, hw, Cp, C, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_upsample2x(const void* x, void* out, int dtype, int B, int H, int W, int C, void* stream) {
  SDB_CHECK_ARG(x && out && B > 0 && H > 0 && W > 0 && C % 8 == 0, "upsample2x: bad args (C %% 8)");
  long long total = (long long)B * 4 * H * W * (C / 8);
  if (dtype == SDB_BF16)
    upsample2x_kernel<__nv_bfloat16><<<ew_grid(total, 256), 256, 0, sdb_stream(stream)>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)out, H, W, C, total);
  else
    upsample2x_kernel<float><<<ew_grid(total, 256), 256, 0, sdb_stream(stream)>>>(
        (const float*)x, (float*)out, H, W, C, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

static int launch_silu_cast(const void* x, int sdt, void* out, int ddt, long long n, int do_silu, cudaStream_t st) {
  const int g = ew_grid(n, 256);
  if (sdt == SDB_F32 && ddt == SDB_F32) silu_cast_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, (float*)out, n, do_silu);
  else if (sdt == SDB_F32) silu_cast_kernel<float, __nv_bfloat16><<<g, 256, 0, st>>>((const float*)x, (__nv_bfloat16*)out, n, do_silu);
  else if (ddt == SDB_F32) silu_cast_kernel<__nv_bfloat16, float><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (float*)out, n, do_silu);
  else silu_cast_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, n, do_silu);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
extern "C" int sdb_silu(const void* x, void* out, int64_t n, int dtype, int out_dtype, void* stream) {
  SDB_CHECK_ARG(x && out && n > 0, "silu: bad args");
  return launch_silu_cast(x, dtype, out, out_dtype, n, 1, sdb_stream(stream));
}
extern "C" int sdb_cast(const void* x, int src_dtype, void* out, int dst_dtype, int64_t n, void* stream) {
  SDB_CHECK_ARG(x && out && n > 0, "cast: bad args");
  return launch_silu_cast(x, src_dtype, out, dst_dtype, n, 0, sdb_stream(stream));
}
extern "C" int sdb_to_uint8_image(const void* x_nchw, int dtype, uint8_t* out_nhwc, int B, int C, int H,
                                  int W, void* stream) {
  SDB_CHECK_ARG(x_nchw && out_nhwc && B > 0 && C > 0 && H > 0 && W > 0, "to_uint8_image: bad args");
  long long hw = (long long)H * W, total = hw * C * B;
  if (dtype == SDB_BF16)
    to_uint8_kernel<__nv_bfloat16><<<ew_grid(total, 256), 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)x_nchw, out_nhwc, C, hw, total);
  else
    to_uint8_kernel<float><<<ew_grid(total, 256), 256, 0, sdb_stream(stream)>>>((const float*)x_nchw, out_nhwc, C, hw, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
