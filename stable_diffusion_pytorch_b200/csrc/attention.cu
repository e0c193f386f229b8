// K4: fused (flash-style) attention  softmax(Q K^T * scale) V  without materialising `sim`
// (reference: modules/transformer.py:93-114).
//
// The UNet's head dim is 20 (SURVEY top item 7): per (q, k) pair there are only ~80 useful flops
// per exp, so the kernel is bound by the MUFU ex2 pipe and the softmax ALU work, not by tensor
// throughput.  S and P therefore stay in registers (warp-level mma.sync m16n8k16 tiles, online
// softmax on the accumulator fragments, P re-used directly as the A operand of P.V) instead of
// round-tripping through TMEM; K/V tiles are staged with cp.async double buffering.
// One CTA = 64 queries x one (batch, head); 4 warps x 16 query rows; 64-key tiles.
#include "common.cuh"

namespace sdb {

constexpr int kAttBM = 64, kAttBN = 64, kAttThreads = 128;

__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// stage `rows` x d (bf16) rows of a [*, ld] matrix into smem rows of PITCH elements; rows >= valid
// are zero-filled.  8-byte granules (d % 4 == 0; head offsets h*d keep 8-byte alignment).
template <int PITCH>
__device__ __forceinline__ void stage_rows(__nv_bfloat16* s, const __nv_bfloat16* g, long long ld,
                                           int row0, int valid_rows, int rows, int d) {
  const int cpr = d >> 2;
  for (int i = threadIdx.x; i < rows * cpr; i += kAttThreads) {
    const int r = i / cpr, c = (i - r * cpr) * 4;
    __nv_bfloat16* dst = s + r * PITCH + c;
    if (row0 + r < valid_rows) cp_async8(smem_u32(dst), g + (long long)(row0 + r) * ld + c);
    else *reinterpret_cast<uint2*>(dst) = make_uint2(0u, 0u);
  }
}

template <int DPAD>
__global__ void __launch_bounds__(kAttThreads)
attention_kernel(const __nv_bfloat16* __restrict__ q, long long ldq, const __nv_bfloat16* __restrict__ k,
                 long long ldk, const __nv_bfloat16* __restrict__ v, long long ldv,
                 __nv_bfloat16* __restrict__ out, long long ldo, int n, int m, int d, float scale_log2e) {
  constexpr int PITCH = DPAD + 8;  // +16 B: ldmatrix rows land in distinct banks
  constexpr int KS = DPAD / 16;    // k-steps of Q.K^T
  constexpr int DT = DPAD / 8;     // 8-wide d tiles of the output
  extern __shared__ __align__(16) uint8_t att_smem[];
  __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(att_smem);
  __nv_bfloat16* sK = sQ + kAttBM * PITCH;
  __nv_bfloat16* sV = sK + 2 * kAttBN * PITCH;

  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kAttBM;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const __nv_bfloat16* qg = q + (long long)b * n * ldq + h * d;
  const __nv_bfloat16* kg = k + (long long)b * m * ldk + h * d;
  const __nv_bfloat16* vg = v + (long long)b * m * ldv + h * d;

  // zero everything once: the pad columns d..DPAD are never written again
  {
    uint4* z = reinterpret_cast<uint4*>(att_smem);
    constexpr int total16 = (kAttBM + 4 * kAttBN) * PITCH * 2 / 16;
    for (int i = threadIdx.x; i < total16; i += kAttThreads) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  __syncthreads();

  const int n_tiles = (m + kAttBN - 1) / kAttBN;
  stage_rows<PITCH>(sQ, qg, ldq, q0, n, kAttBM, d);
  stage_rows<PITCH>(sK, kg, ldk, 0, m, kAttBN, d);
  stage_rows<PITCH>(sV, vg, ldv, 0, m, kAttBN, d);
  cp_async_commit();

  uint32_t qf[KS][4];
  float o[DT][4];
#pragma unroll
  for (int t = 0; t < DT; ++t) { o[t][0] = o[t][1] = o[t][2] = o[t][3] = 0.f; }
  float row_max[2] = {-INFINITY, -INFINITY}, row_sum[2] = {0.f, 0.f};

  for (int j = 0; j < n_tiles; ++j) {
    const int buf = j & 1;
    if (j + 1 < n_tiles) {
      stage_rows<PITCH>(sK + (buf ^ 1) * kAttBN * PITCH, kg, ldk, (j + 1) * kAttBN, m, kAttBN, d);
      stage_rows<PITCH>(sV + (buf ^ 1) * kAttBN * PITCH, vg, ldv, (j + 1) * kAttBN, m, kAttBN, d);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (j == 0) {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const int row = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int col = ks * 16 + (lane >> 4) * 8;
        ldsm_x4(smem_u32(sQ + row * PITCH + col), qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);
      }
    }
    const __nv_bfloat16* tK = sK + buf * kAttBN * PITCH;
    const __nv_bfloat16* tV = sV + buf * kAttBN * PITCH;

    // ---- S = Q K^T : 16 x 64 per warp -----------------------------------------------------
    float s[kAttBN / 8][4];
#pragma unroll
    for (int t = 0; t < kAttBN / 8; ++t) { s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f; }
#pragma unroll
    for (int np = 0; np < kAttBN / 16; ++np) {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const int key = np * 16 + (lane & 7) + ((lane >> 4) & 1) * 8;
        const int col = ks * 16 + ((lane >> 3) & 1) * 8;
        uint32_t b0, b1, b2, b3;
        ldsm_x4(smem_u32(tK + key * PITCH + col), b0, b1, b2, b3);
        mma_bf16(s[2 * np], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], b0, b1);
        mma_bf16(s[2 * np + 1], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], b2, b3);
      }
    }
    // ---- mask the key tail, online softmax ------------------------------------------------
    const int key_base = j * kAttBN + (lane & 3) * 2;
    if (j == n_tiles - 1 && (m & (kAttBN - 1)) != 0) {
#pragma unroll
      for (int t = 0; t < kAttBN / 8; ++t) {
        const int key = key_base + t * 8;
        if (key >= m) { s[t][0] = -INFINITY; s[t][2] = -INFINITY; }
        if (key + 1 >= m) { s[t][1] = -INFINITY; s[t][3] = -INFINITY; }
      }
    }
    float mx[2] = {row_max[0], row_max[1]};
#pragma unroll
    for (int t = 0; t < kAttBN / 8; ++t) {
      mx[0] = fmaxf(mx[0], fmaxf(s[t][0], s[t][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[t][2], s[t][3]));
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
    }
    float alpha[2], msc[2], psum[2] = {0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      alpha[r] = fast_exp2((row_max[r] - mx[r]) * scale_log2e);  // exp2(-inf) = 0 on the first tile
      msc[r] = mx[r] * scale_log2e;
      row_max[r] = mx[r];
    }
    uint32_t pf[kAttBN / 16][4];
#pragma unroll
    for (int t = 0; t < kAttBN / 8; ++t) {
      const float p0 = fast_exp2(fmaf(s[t][0], scale_log2e, -msc[0]));
      const float p1 = fast_exp2(fmaf(s[t][1], scale_log2e, -msc[0]));
      const float p2 = fast_exp2(fmaf(s[t][2], scale_log2e, -msc[1]));
      const float p3 = fast_exp2(fmaf(s[t][3], scale_log2e, -msc[1]));
      psum[0] += p0 + p1;
      psum[1] += p2 + p3;
      // C fragments of two adjacent key tiles form one A fragment of P for the P.V k-step
      pf[t >> 1][(t & 1) * 2] = pack_bf16x2(p0, p1);
      pf[t >> 1][(t & 1) * 2 + 1] = pack_bf16x2(p2, p3);
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) row_sum[r] = row_sum[r] * alpha[r] + psum[r];
#pragma unroll
    for (int t = 0; t < DT; ++t) {
      o[t][0] *= alpha[0]; o[t][1] *= alpha[0];
      o[t][2] *= alpha[1]; o[t][3] *= alpha[1];
    }
    // ---- O += P V ---------------------------------------------------------------------------
#pragma unroll
    for (int kk = 0; kk < kAttBN / 16; ++kk) {
#pragma unroll
      for (int dp = 0; dp < DT / 2; ++dp) {
        const int key = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int col = dp * 16 + (lane >> 4) * 8;
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(smem_u32(tV + key * PITCH + col), b0, b1, b2, b3);
        mma_bf16(o[2 * dp], pf[kk][0], pf[kk][1], pf[kk][2], pf[kk][3], b0, b1);
        mma_bf16(o[2 * dp + 1], pf[kk][0], pf[kk][1], pf[kk][2], pf[kk][3], b2, b3);
      }
    }
    __syncthreads();  // everyone is done with `buf` before the next prefetch overwrites it
  }

  // ---- normalise and store ---------------------------------------------------------------------
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    row_sum[r] += __shfl_xor_sync(0xffffffffu, row_sum[r], 1);
    row_sum[r] += __shfl_xor_sync(0xffffffffu, row_sum[r], 2);
  }
  const float inv0 = 1.0f / row_sum[0], inv1 = 1.0f / row_sum[1];
  const int r0 = q0 + warp * 16 + (lane >> 2), r1 = r0 + 8;
  __nv_bfloat16* og = out + (long long)b * n * ldo + h * d;
#pragma unroll
  for (int t = 0; t < DT; ++t) {
    const int col = t * 8 + (lane & 3) * 2;
    if (col < d) {
      if (r0 < n) *reinterpret_cast<uint32_t*>(og + (long long)r0 * ldo + col) = pack_bf16x2(o[t][0] * inv0, o[t][1] * inv0);
      if (r1 < n) *reinterpret_cast<uint32_t*>(og + (long long)r1 * ldo + col) = pack_bf16x2(o[t][2] * inv1, o[t][3] * inv1);
    }
  }
}

template <int DPAD>
static int launch_attention(const void* q, long long ldq, const void* k, long long ldk, const void* v,
                            long long ldv, void* out, long long ldo, int B, int heads, int n, int m, int d,
                            float scale, cudaStream_t st) {
  constexpr int PITCH = DPAD + 8;
  const int smem = (kAttBM + 4 * kAttBN) * PITCH * 2;
  static bool attr_set[64] = {false};
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (smem > 48 * 1024 && (dev < 0 || dev >= 64 || !attr_set[dev])) {
    SDB_CUDA_OK(cudaFuncSetAttribute(attention_kernel<DPAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  dim3 grid((n + kAttBM - 1) / kAttBM, heads, B);
  attention_kernel<DPAD><<<grid, kAttThreads, smem, st>>>(
      (const __nv_bfloat16*)q, ldq, (const __nv_bfloat16*)k, ldk, (const __nv_bfloat16*)v, ldv,
      (__nv_bfloat16*)out, ldo, n, m, d, scale * 1.4426950408889634f);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

}  // namespace sdb

extern "C" int sdb_attention(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                             int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                             float scale, int dtype, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(q && k && v && out && B > 0 && heads > 0 && n > 0 && m > 0, "attention: bad args");
  if (dtype == SDB_F32)   // fp32 parity mode (precise_f32.cu)
    return sdb_attention_f32_impl(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, sdb_stream(stream));
  SDB_CHECK_ARG(dtype == SDB_BF16, "attention: unknown dtype %d", dtype);
  SDB_CHECK_ARG(d > 0 && d <= 128 && d % 4 == 0, "attention: head dim (%d) must be a multiple of 4, <= 128", d);
  SDB_CHECK_ARG(ldq % 4 == 0 && ldk % 4 == 0 && ldv % 4 == 0 && ldo % 2 == 0, "attention: row strides must be multiples of 4");
  SDB_CHECK_ARG(((uintptr_t)q & 7) == 0 && ((uintptr_t)k & 7) == 0 && ((uintptr_t)v & 7) == 0 && ((uintptr_t)out & 3) == 0,
                "attention: pointers must be 8-byte aligned");
  SDB_CHECK_ARG(heads <= 65535 && B <= 65535, "attention: grid too large");
  cudaStream_t st = sdb_stream(stream);
  if (d == 128) {
    // the tensor-bound shape (VAE bottleneck): tcgen05 pipeline; SDB_ATTN_D128_UMMA=0 keeps the mma.sync kernel (A/B)
    static int umma128 = -1;
    if (umma128 < 0) { const char* e = getenv("SDB_ATTN_D128_UMMA"); umma128 = e ? atoi(e) : 1; }
    if (umma128) {
      const int rc = sdb_attention_d128_dispatch(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, scale, st);
      if (rc != SDB_EUNSUPPORTED) return rc;
    }
  }
  if (d <= 32) return launch_attention<32>(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, st);
  if (d <= 64) return launch_attention<64>(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, st);
  return launch_attention<128>(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, st);
}
