// 16-byte vector versions of the HBM-bound backward kernels (row T1).  The first backward kernels in backward.cu
// address tensors element by element (2-byte loads, channel strips of a GroupNorm group = 10 bytes per 320-byte
// pixel row): 2.3x .. 7x off the copy roofline at config 4.  Here every thread moves 8 bf16 channels per access and
// keeps its per-channel constants / partial sums in registers.  bf16, channel counts that are multiples of 8; the
// entry points in backward.cu fall back to the scalar kernels for anything else.
#include "common.cuh"

namespace sdb {

__device__ __forceinline__ float warp_sum_v(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}
__device__ __forceinline__ float gelu_exact_v(float g) { return 0.5f * g * (1.0f + erff(g * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_grad_v(float g) {
  const float cdf = 0.5f * (1.0f + erff(g * 0.70710678118654752440f));
  const float pdf = 0.3989422804014327f * __expf(-0.5f * g * g);
  return cdf + g * pdf;
}

// ---- GEGLU on pair-interleaved columns: one 16-byte vector of h = (a0, a1, g0, g1, a2, a3, g2, g3) -> 4 outputs ----
__global__ void geglu_fwd_vec_kernel(const uint4* __restrict__ h, uint2* __restrict__ out, long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float f[8];
    unpack8(__ldg(h + i), f);
    out[i] = make_uint2(pack_bf16x2(f[0] * gelu_exact_v(f[2]), f[1] * gelu_exact_v(f[3])),
                        pack_bf16x2(f[4] * gelu_exact_v(f[6]), f[5] * gelu_exact_v(f[7])));
  }
}
__global__ void geglu_bwd_vec_kernel(const uint4* __restrict__ h, const uint2* __restrict__ dout, uint4* __restrict__ dh,
                                     long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float f[8], r[8];
    unpack8(__ldg(h + i), f);
    const uint2 du = __ldg(dout + i);
    const float2 d01 = unpack_bf16x2(du.x), d23 = unpack_bf16x2(du.y);
    r[0] = d01.x * gelu_exact_v(f[2]); r[1] = d01.y * gelu_exact_v(f[3]);
    r[2] = d01.x * f[0] * gelu_grad_v(f[2]); r[3] = d01.y * f[1] * gelu_grad_v(f[3]);
    r[4] = d23.x * gelu_exact_v(f[6]); r[5] = d23.y * gelu_exact_v(f[7]);
    r[6] = d23.x * f[4] * gelu_grad_v(f[6]); r[7] = d23.y * f[5] * gelu_grad_v(f[7]);
    dh[i] = pack8(r);
  }
}

// ---- LayerNorm backward: one warp per row, lane owns vectors lane, lane + 32 (C <= 512); x and dy stay in registers
// for the three reductions; dgamma / dbeta accumulate per lane in registers over all the rows of the warp.
template <int KV>   // vectors per lane
__global__ void __launch_bounds__(256) layernorm_bwd_vec_kernel(const __nv_bfloat16* __restrict__ x,
                                                                const float* __restrict__ gamma,
                                                                const __nv_bfloat16* __restrict__ dy,
                                                                __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                                                                float* __restrict__ dbeta, long long rows, int C, float eps) {
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int vecs = C >> 3;
  float gam[KV][8], dg[KV][8], db[KV][8];
#pragma unroll
  for (int k = 0; k < KV; ++k) {
    const int v = lane + 32 * k;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      gam[k][j] = v < vecs ? gamma[v * 8 + j] : 0.f;
      dg[k][j] = 0.f;
      db[k][j] = 0.f;
    }
  }
  const float inv_c = 1.0f / (float)C;
  for (long long row = (long long)blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += (long long)gridDim.x * wpb) {
    float xv[KV][8], dv[KV][8];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < KV; ++k) {
      const int v = lane + 32 * k;
      if (v < vecs) {
        unpack8(__ldg(reinterpret_cast<const uint4*>(x + row * C) + v), xv[k]);
        unpack8(__ldg(reinterpret_cast<const uint4*>(dy + row * C) + v), dv[k]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) { xv[k][j] = 0.f; dv[k][j] = 0.f; }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) s += xv[k][j];
    }
    const float mean = warp_sum_v(s) * inv_c;
    float ss = 0.f;
#pragma unroll
    for (int k = 0; k < KV; ++k) {
      const bool in = lane + 32 * k < vecs;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = in ? xv[k][j] - mean : 0.f;
        xv[k][j] = d;
        ss = fmaf(d, d, ss);
      }
    }
    const float rstd = rsqrtf(warp_sum_v(ss) * inv_c + eps);
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int k = 0; k < KV; ++k) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        xv[k][j] *= rstd;                         // xhat
        const float g = dv[k][j] * gam[k][j];
        m1 += g;
        m2 = fmaf(g, xv[k][j], m2);
      }
    }
    m1 = warp_sum_v(m1) * inv_c;
    m2 = warp_sum_v(m2) * inv_c;
#pragma unroll
    for (int k = 0; k < KV; ++k) {
      const int v = lane + 32 * k;
      float r[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        r[j] = rstd * (dv[k][j] * gam[k][j] - m1 - xv[k][j] * m2);
        dg[k][j] = fmaf(dv[k][j], xv[k][j], dg[k][j]);
        db[k][j] += dv[k][j];
      }
      if (v < vecs) reinterpret_cast<uint4*>(dx + row * C)[v] = pack8(r);
    }
  }
  __shared__ float sred[2][KV * 256];
  for (int i = threadIdx.x; i < KV * 256; i += blockDim.x) { sred[0][i] = 0.f; sred[1][i] = 0.f; }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < KV; ++k) {
    const int v = lane + 32 * k;
    if (v < vecs) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(&sred[0][v * 8 + j], dg[k][j]);
        atomicAdd(&sred[1][v * 8 + j], db[k][j]);
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    atomicAdd(dgamma + c, sred[0][c]);
    atomicAdd(dbeta + c, sred[1][c]);
  }
}

// ---- GroupNorm (+SiLU) backward over a virtual concat, vectorised ------------------------------------------------
// pass A  per (image, channel): A = sum_px d, Bx = sum_px d * xhat, d = dy * act'(y)   -> colbuf[b][0|1][c]
// finalize per (image, group): s1 = sum_c gamma_c A, s2 = sum_c gamma_c Bx; dgamma_c += sum_b Bx, dbeta_c += sum_b A
// pass B  dx = rstd (d gamma - (s1 + xhat s2) / N)
// Thread = one 8-channel vector (constants of its channels in registers) striding over the pixels.
__device__ __forceinline__ float act_grad_v(float y, int act) {
  if (act != SDB_ACT_SILU) return 1.0f;
  const float s = 1.0f / (1.0f + __expf(-y));
  return s * (1.0f + y * (1.0f - s));
}
struct GnVecArgs {
  const __nv_bfloat16* x1; int C1;
  const __nv_bfloat16* x2; int C2;
  const float* stats; const float* gamma; const float* beta; int act;
  const __nv_bfloat16* dy;
  float* colbuf;            // [B][2][C], zeroed
  float* sums;              // [B][groups][2]
  __nv_bfloat16* dx1; __nv_bfloat16* dx2;
  long long hw; int groups;
};

__global__ void __launch_bounds__(256) gn_bwd_cols_kernel(const GnVecArgs a) {
  __shared__ float sacc[2][2048];
  const int C = a.C1 + a.C2, cpg = C / a.groups, vecs = C >> 3, rpp = 256 / vecs;
  const int v = threadIdx.x % vecs, rr = threadIdx.x / vecs, b = blockIdx.y;
  for (int c = threadIdx.x; c < C; c += 256) { sacc[0][c] = 0.f; sacc[1][c] = 0.f; }
  __syncthreads();
  if (rr < rpp) {
    float mean[8], rstd[8], gam[8], bet[8], A[8], Bx[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = v * 8 + j, g = c / cpg;
      mean[j] = a.stats[((long long)b * a.groups + g) * 2];
      rstd[j] = a.stats[((long long)b * a.groups + g) * 2 + 1];
      gam[j] = a.gamma[c]; bet[j] = a.beta[c];
      A[j] = 0.f; Bx[j] = 0.f;
    }
    const bool first = v * 8 < a.C1;
    const __nv_bfloat16* xb = first ? a.x1 + (long long)b * a.hw * a.C1 + v * 8 : a.x2 + (long long)b * a.hw * a.C2 + (v * 8 - a.C1);
    const long long ldx = first ? a.C1 : a.C2;
    const __nv_bfloat16* db = a.dy + (long long)b * a.hw * C + v * 8;
    const long long per = (a.hw + gridDim.x - 1) / gridDim.x;
    const long long r0 = (long long)blockIdx.x * per, r1 = r0 + per < a.hw ? r0 + per : a.hw;
    for (long long r = r0 + rr; r < r1; r += rpp) {
      float xf[8], df[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(xb + r * ldx)), xf);
      unpack8(__ldg(reinterpret_cast<const uint4*>(db + r * C)), df);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (xf[j] - mean[j]) * rstd[j];
        const float d = df[j] * act_grad_v(fmaf(xh, gam[j], bet[j]), a.act);
        A[j] += d;
        Bx[j] = fmaf(d, xh, Bx[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&sacc[0][v * 8 + j], A[j]);
      atomicAdd(&sacc[1][v * 8 + j], Bx[j]);
    }
  }
  __syncthreads();
  float* cb = a.colbuf + (long long)b * 2 * C;
  for (int c = threadIdx.x; c < C; c += 256) {
    atomicAdd(cb + c, sacc[0][c]);
    atomicAdd(cb + C + c, sacc[1][c]);
  }
}

__global__ void gn_bwd_finalize_kernel(const float* __restrict__ colbuf, const float* __restrict__ gamma,
                                       float* __restrict__ sums, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                       int C, int groups) {
  const int b = blockIdx.x, cpg = C / groups;
  const float* cb = colbuf + (long long)b * 2 * C;
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    float s1 = 0.f, s2 = 0.f;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c) {
      s1 = fmaf(gamma[c], cb[c], s1);
      s2 = fmaf(gamma[c], cb[C + c], s2);
    }
    sums[((long long)b * groups + g) * 2] = s1;
    sums[((long long)b * groups + g) * 2 + 1] = s2;
  }
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    atomicAdd(dbeta + c, cb[c]);
    atomicAdd(dgamma + c, cb[C + c]);
  }
}

__global__ void __launch_bounds__(256) gn_bwd_apply_vec_kernel(const GnVecArgs a) {
  const int C = a.C1 + a.C2, cpg = C / a.groups, vecs = C >> 3, rpp = 256 / vecs;
  const int v = threadIdx.x % vecs, rr = threadIdx.x / vecs, b = blockIdx.y;
  if (rr >= rpp) return;
  float mean[8], rstd[8], gam[8], bet[8], s1[8], s2[8];
  const float inv_n = 1.0f / ((float)a.hw * (float)cpg);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = v * 8 + j, g = c / cpg;
    mean[j] = a.stats[((long long)b * a.groups + g) * 2];
    rstd[j] = a.stats[((long long)b * a.groups + g) * 2 + 1];
    gam[j] = a.gamma[c]; bet[j] = a.beta[c];
    s1[j] = a.sums[((long long)b * a.groups + g) * 2] * inv_n;
    s2[j] = a.sums[((long long)b * a.groups + g) * 2 + 1] * inv_n;
  }
  const bool first = v * 8 < a.C1;
  const long long ldx = first ? a.C1 : a.C2;
  const long long off = first ? (long long)b * a.hw * a.C1 + v * 8 : (long long)b * a.hw * a.C2 + (v * 8 - a.C1);
  const __nv_bfloat16* xb = (first ? a.x1 : a.x2) + off;
  __nv_bfloat16* ob = (first ? a.dx1 : a.dx2) + off;
  const __nv_bfloat16* db = a.dy + (long long)b * a.hw * C + v * 8;
  const long long per = (a.hw + gridDim.x - 1) / gridDim.x;
  const long long r0 = (long long)blockIdx.x * per, r1 = r0 + per < a.hw ? r0 + per : a.hw;
  for (long long r = r0 + rr; r < r1; r += rpp) {
    float xf[8], df[8], o[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(xb + r * ldx)), xf);
    unpack8(__ldg(reinterpret_cast<const uint4*>(db + r * C)), df);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float xh = (xf[j] - mean[j]) * rstd[j];
      const float d = df[j] * act_grad_v(fmaf(xh, gam[j], bet[j]), a.act) * gam[j];
      o[j] = rstd[j] * (d - (s1[j] + xh * s2[j]));
    }
    *reinterpret_cast<uint4*>(ob + r * ldx) = pack8(o);
  }
}

// ---- host launchers (return false when the shape is outside the vector kernels) --------------------------------------
static inline int vec_grid(long long n) {
  long long want = (n + 255) / 256, cap = (long long)sdb_num_sms() * 16;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}
static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static bool geglu_fwd_vec(const void* h, void* out, long long M, int F, cudaStream_t st) {
  if (F % 4 || !al16(h) || (reinterpret_cast<uintptr_t>(out) & 7)) return false;
  const long long nvec = M * (F / 4);
  geglu_fwd_vec_kernel<<<vec_grid(nvec), 256, 0, st>>>((const uint4*)h, (uint2*)out, nvec);
  return true;
}
static bool geglu_bwd_vec(const void* h, const void* dout, void* dh, long long M, int F, cudaStream_t st) {
  if (F % 4 || !al16(h) || !al16(dh) || (reinterpret_cast<uintptr_t>(dout) & 7)) return false;
  const long long nvec = M * (F / 4);
  geglu_bwd_vec_kernel<<<vec_grid(nvec), 256, 0, st>>>((const uint4*)h, (const uint2*)dout, (uint4*)dh, nvec);
  return true;
}
static bool layernorm_bwd_vec(const void* x, const float* gamma, const void* dy, void* dx, float* dgamma, float* dbeta,
                              long long rows, int C, float eps, cudaStream_t st) {
  if (C % 8 || C > 512 || !al16(x) || !al16(dy) || !al16(dx)) return false;
  long long blocks = (rows + 7) / 8;
  const long long cap = (long long)sdb_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (C <= 256)
    layernorm_bwd_vec_kernel<1><<<(unsigned)blocks, 256, 0, st>>>((const __nv_bfloat16*)x, gamma, (const __nv_bfloat16*)dy,
                                                                 (__nv_bfloat16*)dx, dgamma, dbeta, rows, C, eps);
  else
    layernorm_bwd_vec_kernel<2><<<(unsigned)blocks, 256, 0, st>>>((const __nv_bfloat16*)x, gamma, (const __nv_bfloat16*)dy,
                                                                 (__nv_bfloat16*)dx, dgamma, dbeta, rows, C, eps);
  return true;
}
// colbuf: [B][2][C] floats, zeroed by the caller
static bool groupnorm_bwd_vec(GnVecArgs a, float* dgamma, float* dbeta, int B, cudaStream_t st) {
  const int C = a.C1 + a.C2;
  if (a.C1 % 8 || a.C2 % 8 || C > 2048 || C % a.groups || !a.colbuf) return false;
  if (!al16(a.x1) || (a.C2 && !al16(a.x2)) || !al16(a.dy) || !al16(a.dx1) || (a.C2 && !al16(a.dx2))) return false;
  const int rpp = 256 / (C / 8);
  long long sp = (a.hw + (long long)rpp * 8 - 1) / ((long long)rpp * 8);
  long long cap = (8LL * sdb_num_sms() + B - 1) / B;
  if (sp > cap) sp = cap;
  if (sp < 1) sp = 1;
  dim3 grid((unsigned)sp, (unsigned)B);
  gn_bwd_cols_kernel<<<grid, 256, 0, st>>>(a);
  gn_bwd_finalize_kernel<<<B, 256, 0, st>>>(a.colbuf, a.gamma, a.sums, dgamma, dbeta, C, a.groups);
  gn_bwd_apply_vec_kernel<<<grid, 256, 0, st>>>(a);
  return true;
}

}  // namespace sdb
