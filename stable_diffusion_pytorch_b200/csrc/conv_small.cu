// The two thin 3x3 convolutions at the ends of the UNet / VAE decoder: conv_in (Cin = latent
// channels, models/utils.py:31-32) and the output head (Cout = latent / image channels,
// models/utils.py:197).  K = 36 or N = 3..8 cannot fill a tensor-core tile and both are HBM-bound,
// so they run on CUDA cores and also perform the NCHW <-> NHWC change of layout for free.
#include "common.cuh"

namespace sdb {

template <typename T> __device__ __forceinline__ float ld_f(const T* p);
template <> __device__ __forceinline__ float ld_f<float>(const float* p) { return __ldg(p); }
template <> __device__ __forceinline__ float ld_f<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// ---- conv_in: NCHW [B,Cin,H,W] -> NHWC [B,H,W,Cout] --------------------------------------------
// CTA = 32 pixels wide x kInRows rows x 64 couts: the 64-cout weight slab is staged once and re-used for
// every row (it used to be re-staged per 32 pixels, which cost as much as the arithmetic); the input
// window slides down one row at a time.  Thread = one pixel column x 8 couts.
constexpr int kCinMax = 8, kInPx = 32, kInCo = 64, kInRows = 8;

template <typename TI, typename TO>
__global__ void __launch_bounds__(256)
conv3x3_in_kernel(const TI* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                  TO* __restrict__ out, int src_B, int H, int W, int Cin, int Cout) {
  __shared__ float s_in[kCinMax][kInRows + 2][kInPx + 2];
  __shared__ __align__(16) float s_w[kCinMax * 9][kInCo];  // [ci*9 + tap][co]
  const int tiles_w = (W + kInPx - 1) / kInPx;
  const int wt = blockIdx.x % tiles_w;
  const int h0 = (blockIdx.x / tiles_w) * kInRows;
  const int b = blockIdx.z, co0 = blockIdx.y * kInCo, w0 = wt * kInPx;
  for (int i = threadIdx.x; i < Cin * (kInRows + 2) * (kInPx + 2); i += 256) {
    const int ci = i / ((kInRows + 2) * (kInPx + 2)), rem = i - ci * (kInRows + 2) * (kInPx + 2);
    const int dy = rem / (kInPx + 2), dx = rem - dy * (kInPx + 2);
    const int hh = h0 + dy - 1, ww = w0 + dx - 1;
    float v = 0.f;
    if (hh >= 0 && hh < H && ww >= 0 && ww < W)
      v = ld_f<TI>(x + (((long long)(b % src_B) * Cin + ci) * H + hh) * W + ww);
    s_in[ci][dy][dx] = v;
  }
  for (int i = threadIdx.x; i < Cin * 9 * kInCo; i += 256) {
    const int k = i / kInCo, co = i - k * kInCo;
    s_w[k][co] = (co0 + co < Cout) ? w[(long long)(co0 + co) * Cin * 9 + k] : 0.f;
  }
  __syncthreads();
  const int cg = threadIdx.x & 7, px = threadIdx.x >> 3;  // 8 couts per thread, 32 pixels
  const int wo = w0 + px, co = co0 + cg * 8;
  float bv[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bv[j] = (co + j < Cout) ? bias[co + j] : 0.f;
  // four rows at a time: every weight vector fetched from shared memory feeds 4 x 8 FMAs
#pragma unroll 1
  for (int r0 = 0; r0 < kInRows; r0 += 4) {
    if (h0 + r0 >= H) break;
    float acc[4][8];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[r][j] = bv[j];
    for (int ci = 0; ci < Cin; ++ci) {
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float4 w0v = *reinterpret_cast<const float4*>(&s_w[ci * 9 + t][cg * 8]);
        const float4 w1v = *reinterpret_cast<const float4*>(&s_w[ci * 9 + t][cg * 8 + 4]);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float a = s_in[ci][r0 + r + t / 3][px + t % 3];
          acc[r][0] = fmaf(a, w0v.x, acc[r][0]); acc[r][1] = fmaf(a, w0v.y, acc[r][1]);
          acc[r][2] = fmaf(a, w0v.z, acc[r][2]); acc[r][3] = fmaf(a, w0v.w, acc[r][3]);
          acc[r][4] = fmaf(a, w1v.x, acc[r][4]); acc[r][5] = fmaf(a, w1v.y, acc[r][5]);
          acc[r][6] = fmaf(a, w1v.z, acc[r][6]); acc[r][7] = fmaf(a, w1v.w, acc[r][7]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int h = h0 + r0 + r;
      if (h < H && wo < W && co < Cout) {  // Cout % 8 == 0 is checked on the host
        TO* dst = out + (((long long)b * H + h) * W + wo) * Cout + co;
        if constexpr (sizeof(TO) == 2) {
          *reinterpret_cast<uint4*>(dst) = make_uint4(pack_bf16x2(acc[r][0], acc[r][1]), pack_bf16x2(acc[r][2], acc[r][3]),
                                                      pack_bf16x2(acc[r][4], acc[r][5]), pack_bf16x2(acc[r][6], acc[r][7]));
        } else {
          reinterpret_cast<float4*>(dst)[0] = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
          reinterpret_cast<float4*>(dst)[1] = make_float4(acc[r][4], acc[r][5], acc[r][6], acc[r][7]);
        }
      }
    }
  }
}

// ---- conv_out: NHWC [B,H,W,Cin] -> NCHW [B,COUT,H,W];  CTA = 8 x 32 pixels, thread = pixel -----
constexpr int kOutTH = 8, kOutTW = 32, kOutCC = 16, kOutPitch = kOutCC + 4;

template <typename TI, typename TO, int COUT>
__global__ void __launch_bounds__(256)
conv3x3_out_kernel(const TI* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                   TO* __restrict__ out, int H, int W, int Cin) {
  __shared__ __align__(16) float s_x[(kOutTH + 2) * (kOutTW + 2) * kOutPitch];
  __shared__ __align__(16) float s_w[COUT * 9 * kOutCC];  // [co][tap][c]
  const int tiles_w = (W + kOutTW - 1) / kOutTW;
  const int w0 = (blockIdx.x % tiles_w) * kOutTW, h0 = (blockIdx.x / tiles_w) * kOutTH;
  const int b = blockIdx.z;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  float acc[COUT];
#pragma unroll
  for (int co = 0; co < COUT; ++co) acc[co] = bias[co];
  for (int c0 = 0; c0 < Cin; c0 += kOutCC) {
    const int cc = Cin - c0 < kOutCC ? Cin - c0 : kOutCC;
    __syncthreads();
    for (int i = threadIdx.x; i < (kOutTH + 2) * (kOutTW + 2) * kOutCC; i += 256) {
      const int c = i % kOutCC, p = i / kOutCC;
      const int dy = p / (kOutTW + 2), dx = p - dy * (kOutTW + 2);
      const int hh = h0 + dy - 1, ww = w0 + dx - 1;
      float v = 0.f;
      if (c < cc && hh >= 0 && hh < H && ww >= 0 && ww < W)
        v = ld_f<TI>(x + (((long long)b * H + hh) * W + ww) * Cin + c0 + c);
      s_x[p * kOutPitch + c] = v;
    }
    for (int i = threadIdx.x; i < COUT * 9 * kOutCC; i += 256) {
      const int c = i % kOutCC, t = (i / kOutCC) % 9, co = i / (kOutCC * 9);
      s_w[i] = c < cc ? w[((long long)co * Cin + c0 + c) * 9 + t] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float* xp = s_x + ((ty + t / 3) * (kOutTW + 2) + tx + t % 3) * kOutPitch;
#pragma unroll
      for (int c = 0; c < kOutCC; c += 4) {
        const float4 a = *reinterpret_cast<const float4*>(xp + c);
#pragma unroll
        for (int co = 0; co < COUT; ++co) {
          const float4 ww = *reinterpret_cast<const float4*>(s_w + (co * 9 + t) * kOutCC + c);
          acc[co] = fmaf(a.x, ww.x, acc[co]); acc[co] = fmaf(a.y, ww.y, acc[co]);
          acc[co] = fmaf(a.z, ww.z, acc[co]); acc[co] = fmaf(a.w, ww.w, acc[co]);
        }
      }
    }
  }
  const int ho = h0 + ty, wo = w0 + tx;
  if (ho < H && wo < W) {
#pragma unroll
    for (int co = 0; co < COUT; ++co) {
      TO* dst = out + (((long long)b * COUT + co) * H + ho) * W + wo;
      if constexpr (sizeof(TO) == 2) *dst = __float2bfloat16_rn(acc[co]);
      else *dst = acc[co];
    }
  }
}

// ---- 1x1 conv on a thin NCHW tensor (post_quant_conv, models/autoencoder.py:110,160) ----------
template <typename T>
__global__ void conv1x1_small_kernel(const T* __restrict__ x, const float* __restrict__ w,
                                     const float* __restrict__ bias, T* __restrict__ out, long long hw,
                                     int Cin, int Cout, long long total /* B*hw */) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / hw, px = i - b * hw;
    float xin[8];
    for (int c = 0; c < Cin; ++c) xin[c] = ld_f<T>(x + (b * Cin + c) * hw + px);
    for (int co = 0; co < Cout; ++co) {
      float a = bias[co];
      for (int c = 0; c < Cin; ++c) a = fmaf(xin[c], w[co * Cin + c], a);
      if constexpr (sizeof(T) == 2) out[(b * Cout + co) * hw + px] = __float2bfloat16_rn(a);
      else out[(b * Cout + co) * hw + px] = a;
    }
  }
}

template <typename TI, typename TO>
static int launch_out(const void* x, const float* w, const float* bias, void* out, int B, int H, int W,
                      int Cin, int Cout, cudaStream_t st) {
  dim3 grid(((W + kOutTW - 1) / kOutTW) * ((H + kOutTH - 1) / kOutTH), 1, B);
  switch (Cout) {
    case 3: conv3x3_out_kernel<TI, TO, 3><<<grid, 256, 0, st>>>((const TI*)x, w, bias, (TO*)out, H, W, Cin); break;
    case 4: conv3x3_out_kernel<TI, TO, 4><<<grid, 256, 0, st>>>((const TI*)x, w, bias, (TO*)out, H, W, Cin); break;
    case 8: conv3x3_out_kernel<TI, TO, 8><<<grid, 256, 0, st>>>((const TI*)x, w, bias, (TO*)out, H, W, Cin); break;
    default: sdb_set_error("conv3x3_out: Cout (%d) must be 3, 4 or 8", Cout); return SDB_EUNSUPPORTED;
  }
  SDB_LAUNCH_OK();
  return SDB_OK;
}

}  // namespace sdb

extern "C" int sdb_conv3x3_in(const void* x_nchw, int src_dtype, const float* w, const float* bias,
                              void* out_nhwc, int out_dtype, int B, int src_B, int H, int W, int Cin,
                              int Cout, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(x_nchw && w && bias && out_nhwc && B > 0 && H > 0 && W > 0, "conv3x3_in: bad args");
  SDB_CHECK_ARG(src_B > 0 && B % src_B == 0, "conv3x3_in: B (%d) must be a multiple of src_B (%d)", B, src_B);
  SDB_CHECK_ARG(Cin >= 1 && Cin <= kCinMax, "conv3x3_in: Cin (%d) must be <= %d", Cin, kCinMax);
  SDB_CHECK_ARG(Cout % 8 == 0, "conv3x3_in: Cout (%d) must be a multiple of 8", Cout);
  dim3 grid(((W + kInPx - 1) / kInPx) * ((H + kInRows - 1) / kInRows), (Cout + kInCo - 1) / kInCo, B);
  cudaStream_t st = sdb_stream(stream);
#define SDB_CIN(TI, TO) conv3x3_in_kernel<TI, TO><<<grid, 256, 0, st>>>((const TI*)x_nchw, w, bias, (TO*)out_nhwc, src_B, H, W, Cin, Cout)
  if (src_dtype == SDB_F32 && out_dtype == SDB_F32) SDB_CIN(float, float);
  else if (src_dtype == SDB_F32) SDB_CIN(float, __nv_bfloat16);
  else if (out_dtype == SDB_F32) SDB_CIN(__nv_bfloat16, float);
  else SDB_CIN(__nv_bfloat16, __nv_bfloat16);
#undef SDB_CIN
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_conv3x3_out(const void* x_nhwc, int src_dtype, const float* w, const float* bias,
                               void* out_nchw, int out_dtype, int B, int H, int W, int Cin, int Cout,
                               void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(x_nhwc && w && bias && out_nchw && B > 0 && H > 0 && W > 0 && Cin > 0, "conv3x3_out: bad args");
  cudaStream_t st = sdb_stream(stream);
  if (src_dtype == SDB_F32 && out_dtype == SDB_F32) return launch_out<float, float>(x_nhwc, w, bias, out_nchw, B, H, W, Cin, Cout, st);
  if (src_dtype == SDB_F32) return launch_out<float, __nv_bfloat16>(x_nhwc, w, bias, out_nchw, B, H, W, Cin, Cout, st);
  if (out_dtype == SDB_F32) return launch_out<__nv_bfloat16, float>(x_nhwc, w, bias, out_nchw, B, H, W, Cin, Cout, st);
  return launch_out<__nv_bfloat16, __nv_bfloat16>(x_nhwc, w, bias, out_nchw, B, H, W, Cin, Cout, st);
}

extern "C" int sdb_conv1x1_nchw_small(const void* x, const float* w, const float* bias, void* out,
                                      int dtype, int B, int H, int W, int Cin, int Cout, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(x && w && bias && out && B > 0 && H > 0 && W > 0, "conv1x1_small: bad args");
  SDB_CHECK_ARG(Cin >= 1 && Cin <= 8 && Cout >= 1 && Cout <= 8, "conv1x1_small: channels must be <= 8");
  const long long hw = (long long)H * W, total = hw * B;
  long long want = (total + 255) / 256;
  const int grid = (int)(want < 4096 ? want : 4096);
  if (dtype == SDB_F32)
    conv1x1_small_kernel<float><<<grid, 256, 0, sdb_stream(stream)>>>((const float*)x, w, bias, (float*)out, hw, Cin, Cout, total);
  else
    conv1x1_small_kernel<__nv_bfloat16><<<grid, 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)x, w, bias, (__nv_bfloat16*)out, hw, Cin, Cout, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
