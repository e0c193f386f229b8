// K6 timestep embedding and K7 sampler kernels (HBM / latency bound elementwise work).
#include "common.cuh"

namespace sdb {

// ---------------------------------------------------------------------------------------------
// K6: out[b, j] = sin(t_b * freq_j), out[b, half + j] = cos(t_b * freq_j)
// (modules/timestep_embedding.py:101-105).  Arguments reach ~9e6, so the full-range sinf/cosf
// (Payne-Hanek reduction) are mandatory; never build this file with --use_fast_math.
// ---------------------------------------------------------------------------------------------
template <typename OutT>
__global__ void timestep_embed_kernel(const int64_t* __restrict__ t, const float* __restrict__ freq,
                                      OutT* __restrict__ out, int B, int half) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * half) return;
  int b = idx / half, j = idx - b * half;
  float arg = __fmul_rn(static_cast<float>(t[b]), freq[j]);
  float s = sinf(arg), c = cosf(arg);
  if constexpr (sizeof(OutT) == 4) {
    out[(size_t)b * 2 * half + j] = s;
    out[(size_t)b * 2 * half + half + j] = c;
  } else {
    out[(size_t)b * 2 * half + j] = __float2bfloat16_rn(s);
    out[(size_t)b * 2 * half + half + j] = __float2bfloat16_rn(c);
  }
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 + Box-Muller for the on-device noise option.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
__device__ __forceinline__ void normal4(uint64_t seed, uint32_t slot, uint64_t quad, float (&z)[4]) {
  uint4 r = philox4x32_10(make_uint4((uint32_t)quad, (uint32_t)(quad >> 32), slot, 0x5db200u),
                          make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  const float k = 2.3283064365386963e-10f;  // 2^-32
  float u0 = (r.x + 0.5f) * k, u1 = (r.y + 0.5f) * k, u2 = (r.z + 0.5f) * k, u3 = (r.w + 0.5f) * k;
  float r0 = sqrtf(-2.0f * logf(u0)), r1 = sqrtf(-2.0f * logf(u2));
  float s0, c0, s1, c1;
  sincospif(2.0f * u1, &s0, &c0);
  sincospif(2.0f * u3, &s1, &c1);
  z[0] = r0 * c0; z[1] = r0 * s0; z[2] = r1 * c1; z[3] = r1 * s1;
}

// ---------------------------------------------------------------------------------------------
// K7: guidance combine + DDPM / DDIM update, 4 elements per thread.
// Every product and sum is a separately rounded IEEE op in the reference's order
// (latent_diffusion.py:65; scheduler.py:184,196,213) so fp32 results are bit-identical to torch.
// ---------------------------------------------------------------------------------------------
template <typename T> struct Vec4;
template <> struct Vec4<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
    float4 x = *reinterpret_cast<const float4*>(p);
    v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Vec4<__nv_bfloat16> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[4]) {
    uint2 x = *reinterpret_cast<const uint2*>(p);
    float2 a = unpack_bf16x2(x.x), b = unpack_bf16x2(x.y);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[4]) {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
  }
};

template <typename T>
__global__ void __launch_bounds__(256)
cfg_sampler_step_kernel(const T* __restrict__ eps, const T* __restrict__ x_t,
                        const T* __restrict__ noise, const float* __restrict__ coef,
                        const int32_t* __restrict__ slot_dev, int slot, int kind, int cfg_mode,
                        float g, float scale_factor, uint64_t seed, T* __restrict__ x_prev,
                        T* __restrict__ x0_out, int64_t total /* B*n_per, multiple of 4 */) {
  if (slot_dev) {
    slot = *slot_dev;
    if (noise) noise += (size_t)slot * (size_t)total;  // device-slot mode: noise is a per-slot table
  }
  const float* c = coef + 8 * (size_t)slot;
  const float c0 = c[0], c1 = c[1], c2 = c[2], c3 = c[3], sigma = c[4], d0 = c[5], d1 = c[6];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < total; i += stride) {
    float e0[4], e1[4], x[4], nz[4], xp[4], x0[4];
    Vec4<T>::load(eps + i, e0);
    Vec4<T>::load(x_t + i, x);
    if (cfg_mode) Vec4<T>::load(eps + total + i, e1);
    const bool add_noise = (kind == 0) && (sigma != 0.0f);
    if (add_noise) {
      if (noise) Vec4<T>::load(noise + i, nz);
      else normal4(seed, (uint32_t)slot, (uint64_t)(i >> 2), nz);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float e = cfg_mode ? __fadd_rn(e0[j], __fmul_rn(g, __fsub_rn(e0[j], e1[j]))) : e0[j];
      float v0 = __fsub_rn(__fmul_rn(c0, x[j]), __fmul_rn(c1, e));
      x0[j] = v0;
      if (kind == 0) {
        float mean = __fadd_rn(__fmul_rn(c2, v0), __fmul_rn(c3, x[j]));
        xp[j] = add_noise ? __fadd_rn(mean, __fmul_rn(sigma, __fmul_rn(nz[j], scale_factor))) : mean;
      } else {
        xp[j] = __fadd_rn(__fmul_rn(d0, v0), __fmul_rn(d1, e));
      }
    }
    Vec4<T>::store(x_prev + i, xp);
    if (x0_out) Vec4<T>::store(x0_out + i, x0);
  }
}

// guidance combine alone (latent_diffusion.py:62-65) for stand-alone pred_noise() calls
template <typename T>
__global__ void __launch_bounds__(256)
cfg_combine_kernel(const T* __restrict__ eps, float g, T* __restrict__ out, int64_t total) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < total; i += stride) {
    float e0[4], e1[4], o[4];
    Vec4<T>::load(eps + i, e0);
    Vec4<T>::load(eps + total + i, e1);
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = __fadd_rn(e0[j], __fmul_rn(g, __fsub_rn(e0[j], e1[j])));
    Vec4<T>::store(out + i, o);
  }
}

__global__ void sampler_advance_kernel(int32_t* slot_dev, const int64_t* __restrict__ t_table,
                                       int n_slots, int64_t* __restrict__ t_out, int nt) {
  int s = *slot_dev + 1;
  if (s >= n_slots) s = 0;  // wrap so a captured graph can be replayed for the next batch
  __syncthreads();
  if (threadIdx.x == 0) *slot_dev = s;
  int64_t tv = t_table[s];
  for (int i = threadIdx.x; i < nt; i += blockDim.x) t_out[i] = tv;
}

template <typename T>
__global__ void __launch_bounds__(256)
add_noise_kernel(const T* __restrict__ x0, const T* __restrict__ noise, const int64_t* __restrict__ t,
                 const float* __restrict__ ac, T* __restrict__ out, int64_t n_per /* %4 == 0 */) {
  const int b = blockIdx.y;
  float a = ac[t[b]], sa, sb;
  if constexpr (sizeof(T) == 4) {
    sa = __fsqrt_rn(a);
    sb = __fsqrt_rn(__fsub_rn(1.0f, a));
  } else {  // scheduler.py:121-131: table cast to the sample dtype, pow(0.5) taken in that dtype
    float ab = __bfloat162float(__float2bfloat16_rn(a));
    sa = __bfloat162float(__float2bfloat16_rn(sqrtf(ab)));
    sb = __bfloat162float(__float2bfloat16_rn(sqrtf(__bfloat162float(__float2bfloat16_rn(1.0f - ab)))));
  }
  const T* xb = x0 + (size_t)b * n_per;
  const T* nb = noise + (size_t)b * n_per;
  T* ob = out + (size_t)b * n_per;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n_per;
       i += (int64_t)gridDim.x * blockDim.x * 4) {
    float x[4], z[4], o[4];
    Vec4<T>::load(xb + i, x);
    Vec4<T>::load(nb + i, z);
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = __fadd_rn(__fmul_rn(sa, x[j]), __fmul_rn(sb, z[j]));
    Vec4<T>::store(ob + i, o);
  }
}

}  // namespace sdb

extern "C" int sdb_timestep_embed(const int64_t* t, const float* freq, void* out, int B, int half,
                                  int out_dtype, void* stream) {
  SDB_CHECK_ARG(t && freq && out && B > 0 && half > 0, "timestep_embed: bad args");
  int n = B * half, th = 128;
  if (out_dtype == SDB_F32)
    sdb::timestep_embed_kernel<float><<<(n + th - 1) / th, th, 0, sdb_stream(stream)>>>(
        t, freq, (float*)out, B, half);
  else
    sdb::timestep_embed_kernel<__nv_bfloat16><<<(n + th - 1) / th, th, 0, sdb_stream(stream)>>>(
        t, freq, (__nv_bfloat16*)out, B, half);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_cfg_sampler_step(const void* eps, const void* x_t, const void* noise,
                                    const float* coef, const int32_t* slot_dev, int slot, int kind,
                                    int cfg_mode, float guidance, float scale_factor, uint64_t seed,
                                    void* x_prev, void* x0, int B, int64_t n_per, int dtype,
                                    void* stream) {
  SDB_CHECK_ARG(eps && x_t && coef && x_prev && B > 0 && n_per > 0, "sampler_step: bad args");
  SDB_CHECK_ARG(kind == 0 || kind == 1, "sampler_step: kind must be 0 (ddpm) or 1 (ddim)");
  int64_t total = (int64_t)B * n_per;
  SDB_CHECK_ARG(total % 4 == 0, "sampler_step: B*n_per (%lld) must be a multiple of 4",
                (long long)total);
  int th = 256;
  int64_t want = (total / 4 + th - 1) / th;
  int grid = (int)(want < (int64_t)sdb_num_sms() * 8 ? want : (int64_t)sdb_num_sms() * 8);
  if (dtype == SDB_F32)
    sdb::cfg_sampler_step_kernel<float><<<grid, th, 0, sdb_stream(stream)>>>(
        (const float*)eps, (const float*)x_t, (const float*)noise, coef, slot_dev, slot, kind,
        cfg_mode, guidance, scale_factor, seed, (float*)x_prev, (float*)x0, total);
  else
    sdb::cfg_sampler_step_kernel<__nv_bfloat16><<<grid, th, 0, sdb_stream(stream)>>>(
        (const __nv_bfloat16*)eps, (const __nv_bfloat16*)x_t, (const __nv_bfloat16*)noise, coef,
        slot_dev, slot, kind, cfg_mode, guidance, scale_factor, seed, (__nv_bfloat16*)x_prev,
        (__nv_bfloat16*)x0, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_cfg_combine(const void* eps, float guidance, void* out, int64_t total, int dtype,
                               void* stream) {
  SDB_CHECK_ARG(eps && out && total > 0 && total % 4 == 0, "cfg_combine: bad args (total %% 4)");
  int th = 256;
  int64_t want = (total / 4 + th - 1) / th;
  int grid = (int)(want < (int64_t)sdb_num_sms() * 8 ? want : (int64_t)sdb_num_sms() * 8);
  if (dtype == SDB_F32)
    sdb::cfg_combine_kernel<float><<<grid, th, 0, sdb_stream(stream)>>>((const float*)eps, guidance, (float*)out, total);
  else
    sdb::cfg_combine_kernel<__nv_bfloat16><<<grid, th, 0, sdb_stream(stream)>>>(
        (const __nv_bfloat16*)eps, guidance, (__nv_bfloat16*)out, total);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_sampler_advance(int32_t* slot_dev, const int64_t* t_table, int n_slots,
                                   int64_t* t_out, int nt, void* stream) {
  SDB_CHECK_ARG(slot_dev && t_table && t_out && n_slots > 0 && nt > 0, "sampler_advance: bad args");
  sdb::sampler_advance_kernel<<<1, 128, 0, sdb_stream(stream)>>>(slot_dev, t_table, n_slots, t_out,
                                                                nt);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_add_noise(const void* x0, const void* noise, const int64_t* t,
                             const float* alphas_cumprod, int n_steps, void* out, int B,
                             int64_t n_per, int dtype, void* stream) {
  SDB_CHECK_ARG(x0 && noise && t && alphas_cumprod && out && B > 0 && n_per > 0 && n_steps > 0,
                "add_noise: bad args");
  SDB_CHECK_ARG(n_per % 4 == 0, "add_noise: n_per must be a multiple of 4");
  int th = 256;
  int64_t want = (n_per / 4 + th - 1) / th;
  dim3 grid((unsigned)(want < 1024 ? want : 1024), (unsigned)B);
  if (dtype == SDB_F32)
    sdb::add_noise_kernel<float><<<grid, th, 0, sdb_stream(stream)>>>(
        (const float*)x0, (const float*)noise, t, alphas_cumprod, (float*)out, n_per);
  else
    sdb::add_noise_kernel<__nv_bfloat16><<<grid, th, 0, sdb_stream(stream)>>>(
        (const __nv_bfloat16*)x0, (const __nv_bfloat16*)noise, t, alphas_cumprod,
        (__nv_bfloat16*)out, n_per);
  SDB_LAUNCH_OK();
  return SDB_OK;
}
