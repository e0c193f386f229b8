// T1 (train step, train_unet.py:489-529): backward kernels of the fused forward ops.
// Round-1 scope: correct, simple kernels -- the weight/data gradients of every GEMM and convolution reuse the
// tcgen05 forward kernel (dgrad = conv / GEMM with transposed weights; wgrad = GEMM over transposed operands,
// convolutions through an explicit im2col), the rest are plain CUDA-core reductions.  The attention backward
// here covers contexts of <= 128 keys (the repo-default 8x8 / 16x16 training latents); the 4096-token flash
// backward is the next kernel to write.
#include "common.cuh"

namespace sdb {

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float block_sum(float v, float* red /* >= 32 floats */) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  float t = (int)threadIdx.x < nw ? red[threadIdx.x] : 0.f;
  if (w == 0) t = warp_sum(t);
  if (threadIdx.x == 0) red[0] = t;
  __syncthreads();
  return red[0];
}

// ---- column sums: out[g, c] += sum over the rows of group g (bias / time-embedding gradients) -----
// grid (ceil(cols/32), groups, row splits); out must be zeroed by the caller.
template <typename T>
__global__ void colsum_kernel(const T* __restrict__ x, long long ld, long long rows_per_group, int cols,
                              float* __restrict__ out) {
  __shared__ float sm[8][33];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int g = blockIdx.y;
  const long long per = (rows_per_group + gridDim.z - 1) / gridDim.z;
  const long long r0 = (long long)blockIdx.z * per, r1 = r0 + per < rows_per_group ? r0 + per : rows_per_group;
  float acc = 0.f;
  if (c < cols)
    for (long long r = r0 + (threadIdx.x >> 5); r < r1; r += 8) acc += ldf<T>(x + ((long long)g * rows_per_group + r) * ld + c);
  sm[threadIdx.x >> 5][threadIdx.x & 31] = acc;
  __syncthreads();
  if (threadIdx.x < 32 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sm[i][threadIdx.x];
    atomicAdd(out + (long long)g * cols + c, t);
  }
}

// bf16, cols % 8 == 0 (<= 2048), 16-byte aligned rows: every thread owns one 8-column vector of a row (one 16-byte
// load), the block walks 256 / (cols / 8) rows per pass -- full 128-byte lines instead of 64-byte column strips.
// grid (row splits, groups).
__global__ void __launch_bounds__(256) colsum_vec8_kernel(const __nv_bfloat16* __restrict__ x, long long ld,
                                                          long long rows_per_group, int cols, float* __restrict__ out) {
  __shared__ float sacc[2048];
  const int vecs = cols >> 3;
  const int rpp = 256 / vecs;
  const int v = threadIdx.x % vecs, rr = threadIdx.x / vecs;
  const int g = blockIdx.y;
  for (int c = threadIdx.x; c < cols; c += 256) sacc[c] = 0.f;
  __syncthreads();
  const long long per = (rows_per_group + gridDim.x - 1) / gridDim.x;
  const long long r0 = (long long)blockIdx.x * per, r1 = r0 + per < rows_per_group ? r0 + per : rows_per_group;
  if (rr < rpp) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const __nv_bfloat16* base = x + (long long)g * rows_per_group * ld + v * 8;
    for (long long r = r0 + rr; r < r1; r += rpp) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(base + r * ld));
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y;
      acc[4] += c.x; acc[5] += c.y; acc[6] += d.x; acc[7] += d.y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&sacc[v * 8 + j], acc[j]);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cols; c += 256) atomicAdd(out + (long long)g * cols + c, sacc[c]);
}

// ---- GEGLU on pair-interleaved columns (a0, a1, g0, g1, ...): forward and backward -----------------
__device__ __forceinline__ float gelu_exact(float g) { return 0.5f * g * (1.0f + erff(g * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_grad(float g) {
  const float cdf = 0.5f * (1.0f + erff(g * 0.70710678118654752440f));
  const float pdf = 0.3989422804014327f * __expf(-0.5f * g * g);
  return cdf + g * pdf;
}
__global__ void geglu_fwd_kernel(const __nv_bfloat16* __restrict__ h, __nv_bfloat16* __restrict__ out, long long M, int F) {
  const long long total = M * F;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / F;
    const int j = (int)(i - r * F);
    const __nv_bfloat16* p = h + r * 2 * F + (j >> 1) * 4 + (j & 1);
    out[i] = __float2bfloat16_rn(__bfloat162float(p[0]) * gelu_exact(__bfloat162float(p[2])));
  }
}
__global__ void geglu_bwd_kernel(const __nv_bfloat16* __restrict__ h, const __nv_bfloat16* __restrict__ dout,
                                 __nv_bfloat16* __restrict__ dh, long long M, int F) {
  const long long total = M * F;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / F;
    const int j = (int)(i - r * F);
    const long long o = r * 2 * F + (j >> 1) * 4 + (j & 1);
    const float a = __bfloat162float(h[o]), g = __bfloat162float(h[o + 2]), d = __bfloat162float(dout[i]);
    dh[o] = __float2bfloat16_rn(d * gelu_exact(g));
    dh[o + 2] = __float2bfloat16_rn(d * a * gelu_grad(g));
  }
}

// ---- SiLU backward: dx = dy * s (1 + x (1 - s)), s = sigmoid(x) ---------------------------------------
template <typename TX, typename TG>
__global__ void silu_bwd_kernel(const TX* __restrict__ x, const TG* __restrict__ dy, TG* __restrict__ dx, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float v = ldf<TX>(x + i), s = 1.0f / (1.0f + __expf(-v));
    stf<TG>(dx + i, ldf<TG>(dy + i) * s * (1.0f + v * (1.0f - s)));
  }
}

// ---- GroupNorm (+SiLU) backward over a virtual concat --------------------------------------------------
// pass A: per (image, group) s1 = sum dyh*gamma, s2 = sum dyh*gamma*xhat, and per-channel dgamma / dbeta;
// pass B: dx = rstd (dyh*gamma - (s1 + xhat s2) / N).   dyh = dy * act'(y).
template <typename T>
__device__ __forceinline__ float gn_load(const T* x1, int C1, const T* x2, int C2, long long b, long long hw,
                                         long long px, int c) {
  return c < C1 ? ldf<T>(x1 + (b * hw + px) * C1 + c) : ldf<T>(x2 + (b * hw + px) * C2 + (c - C1));
}
__device__ __forceinline__ float act_grad(float y, int act) {
  if (act != SDB_ACT_SILU) return 1.0f;
  const float s = 1.0f / (1.0f + __expf(-y));
  return s * (1.0f + y * (1.0f - s));
}
constexpr int kGnBwdMaxCpg = 512;
template <typename T>
__global__ void __launch_bounds__(256)
gn_bwd_stats_kernel(const T* __restrict__ x1, int C1, const T* __restrict__ x2, int C2, const float* __restrict__ stats,
                    const float* __restrict__ gamma, const float* __restrict__ beta, int act,
                    const T* __restrict__ dy, float* __restrict__ sums /* [B, groups, 2] */,
                    float* __restrict__ dgamma, float* __restrict__ dbeta, long long hw, int groups) {
  __shared__ float red[32];
  __shared__ float s_dg[kGnBwdMaxCpg], s_db[kGnBwdMaxCpg];
  const int C = C1 + C2, cpg = C / groups;
  const int g = blockIdx.x, b = blockIdx.y;
  const float mean = stats[((long long)b * groups + g) * 2], rstd = stats[((long long)b * groups + g) * 2 + 1];
  for (int i = threadIdx.x; i < cpg; i += blockDim.x) { s_dg[i] = 0.f; s_db[i] = 0.f; }
  __syncthreads();
  // a thread keeps ONE channel of the group and strides over pixels: dgamma / dbeta accumulate in registers
  const int lanes_c = cpg < (int)blockDim.x ? cpg : (int)blockDim.x, ppi = blockDim.x / lanes_c;
  const int cc0 = threadIdx.x % lanes_c, p0 = threadIdx.x / lanes_c;
  float s1 = 0.f, s2 = 0.f;
  if (p0 < ppi) {
    for (int cc = cc0; cc < cpg; cc += lanes_c) {
      const int c = g * cpg + cc;
      const float gam = gamma[c], bet = beta[c];
      float dg = 0.f, db = 0.f;
      for (long long px = p0; px < hw; px += ppi) {
        const float xh = (gn_load<T>(x1, C1, x2, C2, b, hw, px, c) - mean) * rstd;
        const float y = fmaf(xh, gam, bet);
        const float d = ldf<T>(dy + ((long long)b * hw + px) * C + c) * act_grad(y, act);
        s1 = fmaf(d, gam, s1);
        s2 = fmaf(d * gam, xh, s2);
        dg = fmaf(d, xh, dg);
        db += d;
      }
      atomicAdd(&s_dg[cc], dg);
      atomicAdd(&s_db[cc], db);
    }
  }
  s1 = block_sum(s1, red);
  s2 = block_sum(s2, red);
  if (threadIdx.x == 0) {
    sums[((long long)b * groups + g) * 2] = s1;
    sums[((long long)b * groups + g) * 2 + 1] = s2;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cpg; i += blockDim.x) {
    atomicAdd(dgamma + g * cpg + i, s_dg[i]);
    atomicAdd(dbeta + g * cpg + i, s_db[i]);
  }
}
template <typename T>
__global__ void gn_bwd_apply_kernel(const T* __restrict__ x1, int C1, const T* __restrict__ x2, int C2,
                                    const float* __restrict__ stats, const float* __restrict__ gamma,
                                    const float* __restrict__ beta, int act, const T* __restrict__ dy,
                                    const float* __restrict__ sums, T* __restrict__ dx1, T* __restrict__ dx2,
                                    long long hw, int groups, int B) {
  const int C = C1 + C2, cpg = C / groups;
  const long long total = (long long)B * hw * C;
  const float inv_n = 1.0f / ((float)hw * (float)cpg);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long bp = i / C, b = bp / hw, px = bp - b * hw;
    const int g = c / cpg;
    const float mean = stats[(b * groups + g) * 2], rstd = stats[(b * groups + g) * 2 + 1];
    const float xh = (gn_load<T>(x1, C1, x2, C2, b, hw, px, c) - mean) * rstd;
    const float y = fmaf(xh, gamma[c], beta[c]);
    const float d = ldf<T>(dy + i) * act_grad(y, act) * gamma[c];
    const float v = rstd * (d - (sums[(b * groups + g) * 2] + xh * sums[(b * groups + g) * 2 + 1]) * inv_n);
    if (c < C1) stf<T>(dx1 + (b * hw + px) * C1 + c, v);
    else stf<T>(dx2 + (b * hw + px) * C2 + (c - C1), v);
  }
}

// ---- LayerNorm backward: one warp per row; dgamma / dbeta through shared then global atomics ----------
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const T* __restrict__ x, const float* __restrict__ gamma, const T* __restrict__ dy,
                     T* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, long long rows,
                     int C, float eps) {
  extern __shared__ float sm[];   // [2*C]
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  for (long long row = (long long)blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += (long long)gridDim.x * wpb) {
    const T* xr = x + row * C;
    const T* dr = dy + row * C;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += ldf<T>(xr + c);
    const float mean = warp_sum(s) / (float)C;
    float ss = 0.f;
    for (int c = lane; c < C; c += 32) { const float d = ldf<T>(xr + c) - mean; ss = fmaf(d, d, ss); }
    const float rstd = rsqrtf(warp_sum(ss) / (float)C + eps);
    float m1 = 0.f, m2 = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float xh = (ldf<T>(xr + c) - mean) * rstd, g = ldf<T>(dr + c) * gamma[c];
      m1 += g;
      m2 += g * xh;
    }
    m1 = warp_sum(m1) / (float)C;
    m2 = warp_sum(m2) / (float)C;
    for (int c = lane; c < C; c += 32) {
      const float xh = (ldf<T>(xr + c) - mean) * rstd, d = ldf<T>(dr + c);
      stf<T>(dx + row * C + c, rstd * (d * gamma[c] - m1 - xh * m2));
      atomicAdd(&sm[c], d * xh);
      atomicAdd(&sm[C + c], d);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    atomicAdd(dgamma + i, sm[i]);
    atomicAdd(dbeta + i, sm[C + i]);
  }
}

// ---- resampling -----------------------------------------------------------------------------------------
// nearest x2 upsample backward: dx = sum of the 2x2 block of dy
template <typename T>
__global__ void upsample2x_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int H, int W, int C, long long total) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long t = i / C;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H);
    const long long b = t / H;
    const T* p = dy + (((b * 2 * H + 2 * h) * 2 * W + 2 * w) * C + c);
    const long long rs = 2LL * W * C;
    stf<T>(dx + i, ldf<T>(p) + ldf<T>(p + C) + ldf<T>(p + rs) + ldf<T>(p + rs + C));
  }
}
// conv-then-nearest-x0.5 backward: the gradient lands on the even pixels of the full-resolution conv output
template <typename T>
__global__ void scatter_even_kernel(const T* __restrict__ dy, T* __restrict__ out, int H, int W, int C, long long total) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long t = i / C;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H);
    const long long b = t / H;
    float v = 0.f;
    if (((h | w) & 1) == 0) v = ldf<T>(dy + (((b * (H >> 1) + (h >> 1)) * (W >> 1) + (w >> 1)) * C + c));
    stf<T>(out + i, v);
  }
}
// im2col for the 3x3 / pad 1 weight gradient: [B,H,W,C] -> [B*H*W, ldc] with columns (tap, c), zero padded
__global__ void im2col3x3_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int H, int W,
                                 int C, int ldc, long long total /* B*H*W*ldc */) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int col = (int)(i % ldc);
    long long t = i / ldc;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H);
    const long long b = t / H;
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (col < 9 * C) {
      const int tap = col / C, c = col - tap * C;
      const int hh = h + tap / 3 - 1, ww = w + tap % 3 - 1;
      if (hh >= 0 && hh < H && ww >= 0 && ww < W) v = x[((b * H + hh) * W + ww) * C + c];
    }
    out[i] = v;
  }
}
// NCHW [B,C,H,W] (any float type) -> channels-last [B,H,W,Cp] bf16 / fp32 with zero padding channels
template <typename S, typename D>
__global__ void nchw_to_nhwc_pad_kernel(const S* __restrict__ src, D* __restrict__ dst, int C, int Cp, long long hw,
                                        long long total /* B*hw*Cp */) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % Cp);
    const long long bp = i / Cp, b = bp / hw, px = bp - b * hw;
    stf<D>(dst + i, c < C ? ldf<S>(src + (b * C + c) * hw + px) : 0.f);
  }
}

// ---- loss, guidance, optimiser --------------------------------------------------------------------------
// mean squared error in fp32 (train_unet.py:517) and its gradient in one pass; *loss must be zeroed.
template <typename T>
__global__ void mse_kernel(const T* __restrict__ pred, const T* __restrict__ target, long long n, float dloss,
                           float* __restrict__ loss, T* __restrict__ dpred) {
  __shared__ float red[32];
  float acc = 0.f;
  const float inv = 1.0f / (float)n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float d = ldf<T>(pred + i) - ldf<T>(target + i);
    acc = fmaf(d, d, acc);
    if (dpred) stf<T>(dpred + i, 2.0f * d * inv * dloss);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(loss, acc * inv);
}
// eps = e0 + g (e0 - e1)  ->  de0 = (1 + g) d, de1 = -g d   (reference's operand order, latent_diffusion.py:62-65)
template <typename T>
__global__ void cfg_combine_bwd_kernel(const T* __restrict__ d, T* __restrict__ de2, long long half, float g) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < half; i += (long long)gridDim.x * blockDim.x) {
    const float v = ldf<T>(d + i);
    stf<T>(de2 + i, (1.0f + g) * v);
    stf<T>(de2 + half + i, -g * v);
  }
}
__global__ void sumsq_kernel(const float* __restrict__ x, long long n, float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    acc = fmaf(x[i], x[i], acc);
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(out, acc);
}
// fused gradient scaling (all-reduce mean), global-norm clip (torch.nn.utils.clip_grad_norm_ semantics:
// scale = min(1, max_norm / (norm + 1e-6))) and AdamW (decoupled weight decay, torch.optim.AdamW) over a flat bucket
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                             float* __restrict__ v, long long n, float lr, float b1, float b2, float eps, float wd,
                             float bc1, float bc2, float grad_scale, const float* __restrict__ gnorm_sq, float max_norm) {
  float clip = 1.0f;
  if (gnorm_sq != nullptr && max_norm > 0.f) {
    const float norm = sqrtf(*gnorm_sq) * grad_scale;
    clip = fminf(1.0f, max_norm / (norm + 1e-6f));
  }
  const float gs = grad_scale * clip;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i] * gs;
    float pi = p[i] * (1.0f - lr * wd);
    const float mi = fmaf(b1, m[i], (1.0f - b1) * gi);
    const float vi = fmaf(b2, v[i], (1.0f - b2) * gi * gi);
    m[i] = mi;
    v[i] = vi;
    pi -= lr * (mi / bc1) / (sqrtf(vi / bc2) + eps);
    p[i] = pi;
  }
}

// the same update with the per-step scalars (lr, 1 - beta1^t, 1 - beta2^t) read from device memory, so that a
// captured CUDA graph of the train step can be replayed with a new step count / learning rate
__global__ void adamw_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                 float* __restrict__ v, long long n, const float* __restrict__ hyper, float b1, float b2,
                                 float eps, float wd, float grad_scale, const float* __restrict__ gnorm_sq,
                                 float max_norm) {
  const float lr = hyper[0], bc1 = hyper[1], bc2 = hyper[2];
  float clip = 1.0f;
  if (gnorm_sq != nullptr && max_norm > 0.f) {
    const float norm = sqrtf(*gnorm_sq) * grad_scale;
    clip = fminf(1.0f, max_norm / (norm + 1e-6f));
  }
  const float gs = grad_scale * clip;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i] * gs;
    float pi = p[i] * (1.0f - lr * wd);
    const float mi = fmaf(b1, m[i], (1.0f - b1) * gi);
    const float vi = fmaf(b2, v[i], (1.0f - b2) * gi * gi);
    m[i] = mi;
    v[i] = vi;
    pi -= lr * (mi / bc1) / (sqrtf(vi / bc2) + eps);
    p[i] = pi;
  }
}

// ---- dropout (nn.Dropout in training mode: resnet2d.py:190, transformer.py:58,147) ------------------------------
// Counter-based: element i is kept iff the 16-bit slice of mix64(key + i / 4) that belongs to it is >= p * 65536,
// key = mix64(state[0], site).  No mask is stored: the backward re-runs the same kernel on the gradient with the same
// (state, site); `state` lives in device memory so a replayed CUDA graph draws fresh masks after state += 1.
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
template <typename T>
__global__ void dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, unsigned thresh, float scale,
                               const long long* __restrict__ state, unsigned long long site) {
  const unsigned long long key = mix64(mix64((unsigned long long)state[0]) ^ (site * 0xD6E8FEB86659FD93ull));
  const long long quads = (n + 3) >> 2;
  for (long long qd = (long long)blockIdx.x * blockDim.x + threadIdx.x; qd < quads; qd += (long long)gridDim.x * blockDim.x) {
    const unsigned long long r = mix64(key + (unsigned long long)qd);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = qd * 4 + j;
      if (i < n) {
        const bool keep = (unsigned)((r >> (16 * j)) & 0xffffu) >= thresh;
        stf<T>(y + i, keep ? ldf<T>(x + i) * scale : 0.f);
      }
    }
  }
}

static inline int bw_grid(long long n, int threads) {
  long long want = (n + threads - 1) / threads;
  long long cap = (long long)sdb_num_sms() * 16;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace sdb

using namespace sdb;

extern "C" int sdb_colsum(const void* x, int dtype, int64_t ld, int64_t groups, int64_t rows_per_group, int cols,
                          float* out_zeroed, void* stream) {
  SDB_CHECK_ARG(x && out_zeroed && groups > 0 && rows_per_group > 0 && cols > 0 && ld >= cols, "colsum: bad args");
  SDB_CHECK_ARG(groups <= 65535, "colsum: too many groups");
  if (dtype == SDB_BF16 && cols % 8 == 0 && cols <= 2048 && ld % 8 == 0 && ((uintptr_t)x & 15) == 0) {
    const int rpp = 256 / (cols / 8);
    long long sp = (rows_per_group + (long long)rpp * 8 - 1) / ((long long)rpp * 8);   // >= 8 passes per block
    long long cap = (4LL * sdb_num_sms() + groups - 1) / groups;
    if (sp > cap) sp = cap;
    if (sp < 1) sp = 1;
    colsum_vec8_kernel<<<dim3((unsigned)sp, (unsigned)groups), 256, 0, sdb_stream(stream)>>>(
        (const __nv_bfloat16*)x, ld, rows_per_group, cols, out_zeroed);
    SDB_LAUNCH_OK();
    return SDB_OK;
  }
  long long splits = rows_per_group / 256;
  if (splits < 1) splits = 1;
  if (splits > 64) splits = 64;
  dim3 grid((cols + 31) / 32, (unsigned)groups, (unsigned)splits);
  if (dtype == SDB_BF16) colsum_kernel<__nv_bfloat16><<<grid, 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)x, ld, rows_per_group, cols, out_zeroed);
  else colsum_kernel<float><<<grid, 256, 0, sdb_stream(stream)>>>((const float*)x, ld, rows_per_group, cols, out_zeroed);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_geglu_fwd(const void* h, void* out, int64_t M, int F, void* stream) {
  SDB_CHECK_ARG(h && out && M > 0 && F > 0 && F % 2 == 0, "geglu_fwd: bad args");
  if (geglu_fwd_vec(h, out, M, F, sdb_stream(stream))) { SDB_LAUNCH_OK(); return SDB_OK; }
  geglu_fwd_kernel<<<bw_grid(M * F, 256), 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)h, (__nv_bfloat16*)out, M, F);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
extern "C" int sdb_geglu_bwd(const void* h, const void* dout, void* dh, int64_t M, int F, void* stream) {
  SDB_CHECK_ARG(h && dout && dh && M > 0 && F > 0 && F % 2 == 0, "geglu_bwd: bad args");
  if (geglu_bwd_vec(h, dout, dh, M, F, sdb_stream(stream))) { SDB_LAUNCH_OK(); return SDB_OK; }
  geglu_bwd_kernel<<<bw_grid(M * F, 256), 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)h, (const __nv_bfloat16*)dout, (__nv_bfloat16*)dh, M, F);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_silu_bwd(const void* x, int x_dtype, const void* dy, void* dx, int g_dtype, int64_t n, void* stream) {
  SDB_CHECK_ARG(x && dy && dx && n > 0, "silu_bwd: bad args");
  const int g = bw_grid(n, 256);
  cudaStream_t st = sdb_stream(stream);
  if (x_dtype == SDB_BF16 && g_dtype == SDB_BF16) silu_bwd_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, n);
  else if (x_dtype == SDB_BF16) silu_bwd_kernel<__nv_bfloat16, float><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (const float*)dy, (float*)dx, n);
  else if (g_dtype == SDB_BF16) silu_bwd_kernel<float, __nv_bfloat16><<<g, 256, 0, st>>>((const float*)x, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, n);
  else silu_bwd_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, (const float*)dy, (float*)dx, n);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_groupnorm_bwd(const void* x1, int C1, const void* x2, int C2, int dtype, const float* stats,
                                 const float* gamma, const float* beta, int act, const void* dy, void* dx1, void* dx2,
                                 float* dgamma_acc, float* dbeta_acc, float* sums_scratch, int B, int64_t hw, int groups,
                                 void* stream) {
  if (!x2) C2 = 0;
  const int C = C1 + C2;
  SDB_CHECK_ARG(x1 && stats && gamma && beta && dy && dx1 && (C2 == 0 || dx2) && dgamma_acc && dbeta_acc && sums_scratch,
                "groupnorm_bwd: null pointer");
  SDB_CHECK_ARG(B > 0 && hw > 0 && groups > 0 && C % groups == 0 && C / groups <= kGnBwdMaxCpg,
                "groupnorm_bwd: bad shape (channels per group must be <= 512)");
  cudaStream_t st = sdb_stream(stream);
  const long long total = (long long)B * hw * C;
#define SDB_GNB(T)                                                                                                  \
  gn_bwd_stats_kernel<T><<<dim3(groups, B), 256, 0, st>>>((const T*)x1, C1, (const T*)x2, C2, stats, gamma, beta, act, \
                                                         (const T*)dy, sums_scratch, dgamma_acc, dbeta_acc, hw, groups); \
  gn_bwd_apply_kernel<T><<<bw_grid(total, 256), 256, 0, st>>>((const T*)x1, C1, (const T*)x2, C2, stats, gamma, beta, act, \
                                                            (const T*)dy, sums_scratch, (T*)dx1, (T*)dx2, hw, groups, B)
  if (dtype == SDB_BF16) { SDB_GNB(__nv_bfloat16); } else { SDB_GNB(float); }
#undef SDB_GNB
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_layernorm_bwd(const void* x, const float* gamma, const void* dy, void* dx, float* dgamma_acc,
                                 float* dbeta_acc, int64_t rows, int C, float eps, int dtype, void* stream) {
  SDB_CHECK_ARG(x && gamma && dy && dx && dgamma_acc && dbeta_acc && rows > 0 && C > 0 && C <= 4096, "layernorm_bwd: bad args");
  if (dtype == SDB_BF16 && layernorm_bwd_vec(x, gamma, dy, dx, dgamma_acc, dbeta_acc, rows, C, eps, sdb_stream(stream))) {
    SDB_LAUNCH_OK();
    return SDB_OK;
  }
  long long blocks = (rows + 7) / 8;
  const long long cap = (long long)sdb_num_sms() * 4;
  if (blocks > cap) blocks = cap;
  const size_t smem = (size_t)2 * C * sizeof(float);
  if (dtype == SDB_BF16)
    layernorm_bwd_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, smem, sdb_stream(stream)>>>(
        (const __nv_bfloat16*)x, gamma, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, dgamma_acc, dbeta_acc, rows, C, eps);
  else
    layernorm_bwd_kernel<float><<<(unsigned)blocks, 256, smem, sdb_stream(stream)>>>(
        (const float*)x, gamma, (const float*)dy, (float*)dx, dgamma_acc, dbeta_acc, rows, C, eps);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_upsample2x_bwd(const void* dy, void* dx, int dtype, int B, int H, int W, int C, void* stream) {
  SDB_CHECK_ARG(dy && dx && B > 0 && H > 0 && W > 0 && C > 0, "upsample2x_bwd: bad args");
  const long long total = (long long)B * H * W * C;
  if (dtype == SDB_BF16) upsample2x_bwd_kernel<__nv_bfloat16><<<bw_grid(total, 256), 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, H, W, C, total);
  else upsample2x_bwd_kernel<float><<<bw_grid(total, 256), 256, 0, sdb_stream(stream)>>>((const float*)dy, (float*)dx, H, W, C, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
extern "C" int sdb_scatter_even(const void* dy, void* out, int dtype, int B, int H, int W, int C, void* stream) {
  SDB_CHECK_ARG(dy && out && B > 0 && H > 0 && W > 0 && C > 0 && H % 2 == 0 && W % 2 == 0, "scatter_even: bad args");
  const long long total = (long long)B * H * W * C;
  if (dtype == SDB_BF16) scatter_even_kernel<__nv_bfloat16><<<bw_grid(total, 256), 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)dy, (__nv_bfloat16*)out, H, W, C, total);
  else scatter_even_kernel<float><<<bw_grid(total, 256), 256, 0, sdb_stream(stream)>>>((const float*)dy, (float*)out, H, W, C, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
extern "C" int sdb_im2col3x3(const void* x, void* out, int B, int H, int W, int C, int ldc, void* stream) {
  SDB_CHECK_ARG(x && out && B > 0 && H > 0 && W > 0 && C > 0 && ldc >= 9 * C, "im2col3x3: bad args");
  const long long total = (long long)B * H * W * ldc;
  im2col3x3_kernel<<<bw_grid(total, 256), 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, H, W, C, ldc, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
extern "C" int sdb_nchw_to_nhwc_pad(const void* src, int src_dtype, void* dst, int dst_dtype, int B, int C, int Cp, int H,
                                    int W, void* stream) {
  SDB_CHECK_ARG(src && dst && B > 0 && C > 0 && Cp >= C && H > 0 && W > 0, "nchw_to_nhwc_pad: bad args");
  const long long hw = (long long)H * W, total = hw * B * Cp;
  const int g = bw_grid(total, 256);
  cudaStream_t st = sdb_stream(stream);
  if (src_dtype == SDB_F32 && dst_dtype == SDB_F32) nchw_to_nhwc_pad_kernel<float, float><<<g, 256, 0, st>>>((const float*)src, (float*)dst, C, Cp, hw, total);
  else if (src_dtype == SDB_F32) nchw_to_nhwc_pad_kernel<float, __nv_bfloat16><<<g, 256, 0, st>>>((const float*)src, (__nv_bfloat16*)dst, C, Cp, hw, total);
  else if (dst_dtype == SDB_F32) nchw_to_nhwc_pad_kernel<__nv_bfloat16, float><<<g, 256, 0, st>>>((const __nv_bfloat16*)src, (float*)dst, C, Cp, hw, total);
  else nchw_to_nhwc_pad_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst, C, Cp, hw, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_mse(const void* pred, const void* target, int dtype, int64_t n, float dloss, float* loss_zeroed,
                       void* dpred_or_null, void* stream) {
  SDB_CHECK_ARG(pred && target && loss_zeroed && n > 0, "mse: bad args");
  int g = bw_grid(n, 256);
  if (g > 1024) g = 1024;
  if (dtype == SDB_BF16) mse_kernel<__nv_bfloat16><<<g, 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)pred, (const __nv_bfloat16*)target, n, dloss, loss_zeroed, (__nv_bfloat16*)dpred_or_null);
  else mse_kernel<float><<<g, 256, 0, sdb_stream(stream)>>>((const float*)pred, (const float*)target, n, dloss, loss_zeroed, (float*)dpred_or_null);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
extern "C" int sdb_cfg_combine_bwd(const void* d, void* de2, int dtype, int64_t half, float guidance, void* stream) {
  SDB_CHECK_ARG(d && de2 && half > 0, "cfg_combine_bwd: bad args");
  if (dtype == SDB_BF16) cfg_combine_bwd_kernel<__nv_bfloat16><<<bw_grid(half, 256), 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)d, (__nv_bfloat16*)de2, half, guidance);
  else cfg_combine_bwd_kernel<float><<<bw_grid(half, 256), 256, 0, sdb_stream(stream)>>>((const float*)d, (float*)de2, half, guidance);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
extern "C" int sdb_sumsq(const float* x, int64_t n, float* out_zeroed, void* stream) {
  SDB_CHECK_ARG(x && out_zeroed && n > 0, "sumsq: bad args");
  int g = bw_grid(n, 256);
  if (g > 2048) g = 2048;
  sumsq_kernel<<<g, 256, 0, sdb_stream(stream)>>>(x, n, out_zeroed);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
extern "C" int sdb_adamw(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                         float eps, float weight_decay, int step, float grad_scale, const float* gnorm_sq_or_null,
                         float max_norm, void* stream) {
  SDB_CHECK_ARG(p && g && m && v && n > 0 && step >= 1, "adamw: bad args");
  const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
  adamw_kernel<<<bw_grid(n, 256), 256, 0, sdb_stream(stream)>>>(p, g, m, v, n, lr, beta1, beta2, eps, weight_decay, bc1, bc2,
                                                              grad_scale, gnorm_sq_or_null, max_norm);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_adamw_dev(float* p, const float* g, float* m, float* v, int64_t n, const float* hyper_dev, float beta1,
                             float beta2, float eps, float weight_decay, float grad_scale,
                             const float* gnorm_sq_or_null, float max_norm, void* stream) {
  SDB_CHECK_ARG(p && g && m && v && hyper_dev && n > 0, "adamw_dev: bad args");
  adamw_dev_kernel<<<bw_grid(n, 256), 256, 0, sdb_stream(stream)>>>(p, g, m, v, n, hyper_dev, beta1, beta2, eps,
                                                                  weight_decay, grad_scale, gnorm_sq_or_null, max_norm);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_dropout(const void* x, void* y, int64_t n, float p, const int64_t* state_dev, uint64_t site, int dtype,
                           void* stream) {
  SDB_CHECK_ARG(x && y && state_dev && n > 0 && p >= 0.f && p < 1.f, "dropout: bad args (0 <= p < 1)");
  const unsigned thresh = (unsigned)(p * 65536.0f + 0.5f);
  const float scale = 65536.0f / (65536.0f - (float)thresh);      // exact inverse of the realised keep probability
  const int g = bw_grid((n + 3) / 4, 256);
  if (dtype == SDB_BF16)
    dropout_kernel<__nv_bfloat16><<<g, 256, 0, sdb_stream(stream)>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n, thresh, scale, (const long long*)state_dev, site);
  else
    dropout_kernel<float><<<g, 256, 0, sdb_stream(stream)>>>((const float*)x, (float*)y, n, thresh, scale, (const long long*)state_dev, site);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

// sdb_groupnorm_bwd on 16-byte vectors (backward_vec.cu): needs colbuf_zeroed = [B][2][C] floats; shapes outside the
// vector kernels (fp32, channel counts that are not multiples of 8) run the scalar kernels.
extern "C" int sdb_groupnorm_bwd_vec(const void* x1, int C1, const void* x2, int C2, int dtype, const float* stats,
                                     const float* gamma, const float* beta, int act, const void* dy, void* dx1, void* dx2,
                                     float* dgamma_acc, float* dbeta_acc, float* sums_scratch, float* colbuf_zeroed, int B,
                                     int64_t hw, int groups, void* stream) {
  if (!x2) C2 = 0;
  SDB_CHECK_ARG(x1 && stats && gamma && beta && dy && dx1 && (C2 == 0 || dx2) && dgamma_acc && dbeta_acc && sums_scratch &&
                    colbuf_zeroed, "groupnorm_bwd_vec: null pointer");
  SDB_CHECK_ARG(B > 0 && hw > 0 && groups > 0 && (C1 + C2) % groups == 0, "groupnorm_bwd_vec: bad shape");
  if (dtype == SDB_BF16) {
    GnVecArgs a;
    a.x1 = (const __nv_bfloat16*)x1; a.C1 = C1; a.x2 = (const __nv_bfloat16*)x2; a.C2 = C2;
    a.stats = stats; a.gamma = gamma; a.beta = beta; a.act = act; a.dy = (const __nv_bfloat16*)dy;
    a.colbuf = colbuf_zeroed; a.sums = sums_scratch; a.dx1 = (__nv_bfloat16*)dx1; a.dx2 = (__nv_bfloat16*)dx2;
    a.hw = hw; a.groups = groups;
    if (groupnorm_bwd_vec(a, dgamma_acc, dbeta_acc, B, sdb_stream(stream))) {
      SDB_LAUNCH_OK();
      return SDB_OK;
    }
  }
  return sdb_groupnorm_bwd(x1, C1, x2, C2, dtype, stats, gamma, beta, act, dy, dx1, dx2, dgamma_acc, dbeta_acc,
                           sums_scratch, B, hw, groups, stream);
}
