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
constexpr int kGnMaxGroupsSums = 128;   // groups the sums-fed variant keeps in shared memory

template <typename T, typename OutT>
__global__ void __launch_bounds__(kGnThreads, 4)
gn_apply_kernel(const T* __restrict__ x1, int C1, const T* __restrict__ x2, int C2,
                const float* __restrict__ stats, const float* __restrict__ gamma,
                const float* __restrict__ beta, int act, OutT* __restrict__ out, long long hw,
                int groups, int chunk_px, const void* __restrict__ cs1, const void* __restrict__ cs2,
                float eps) {
  const int C = C1 + C2, V = C >> 3, cpg = C / groups;
  const int R = kGnThreads / V;
  const int b = blockIdx.y;
  const int vec = threadIdx.x % V, r = threadIdx.x / V;
  __shared__ float s_stats[2 * kGnMaxGroupsSums];
  if (cs1) {
    // statistics from the per-channel sums the producing conv / GEMM epilogues left behind (sdb_groupnorm_apply_sums):
    // a group of the virtual concat may straddle the two tensors
    for (int g = threadIdx.x; g < groups; g += kGnThreads) {
      long long s = 0;
      unsigned long long ss = 0;              // fixed point (common.cuh): exact integer sums, whatever the order
      for (int cc = g * cpg; cc < (g + 1) * cpg; ++cc) {
        const longlong2 v = cc < C1 ? __ldg(reinterpret_cast<const longlong2*>(cs1) + (long long)b * C1 + cc)
                                    : __ldg(reinterpret_cast<const longlong2*>(cs2) + (long long)b * C2 + (cc - C1));
        s += v.x;
        ss += (unsigned long long)v.y;
      }
      const double inv = 1.0 / ((double)cpg * (double)hw);
      const double mean = (double)s * (1.0 / (double)kChanSumScale) * inv;
      double var = (double)ss * (1.0 / (double)kChanSqScale) * inv - mean * mean;
      var = var > 0.0 ? var : 0.0;
      s_stats[2 * g] = (float)mean;
      s_stats[2 * g + 1] = rsqrtf((float)var + eps);
    }
    __syncthreads();
  }
  if (r >= R) return;
  const int c = vec * 8;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c + j) / cpg;
    const float mean = cs1 ? s_stats[2 * g] : stats[((long long)b * groups + g) * 2];
    const float rstd = cs1 ? s_stats[2 * g + 1] : stats[((long long)b * groups + g) * 2 + 1];
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

// ---- K2, single pass: GroupNorm (+SiLU) with the image resident in shared memory across a thread-block cluster ----
// The three-kernel route reads the tensor twice (statistics, then normalise) and launches three times.  Here one
// cluster of S CTAs owns one image: CTA `rank` bulk-copies its slice of hw / S pixels (a contiguous run of bytes in
// NHWC; two runs for the virtual concat) into shared memory in pieces, sums every piece as it lands (per-channel
// sum / sum of squares in registers, fixed-order reduction through shared memory -> deterministic), the S CTAs
// exchange their per-group partial sums over distributed shared memory, and every CTA normalises its slice out of
// shared memory into the dense output.  HBM traffic: one read + one write of the tensor.
constexpr int kGfThreads = 512;
constexpr int kGfMaxPieces = 16;
constexpr int kGfStatLanes = 8;     // pixel lanes of the statistics pass (bounds the reduction scratch to 8 * 2C floats)

// (SiLU stays the ex2 + rcp form of silu_f: the one-MUFU 0.5 x (1 + tanh(x / 2)) variant was measured -- tanh.approx's
// 2^-11 absolute error is a 10 % relative error on the negative tail, and the VAE encoder parity went 2.4e-2 -> 4.4e-2.)
struct GnFusedParams {
  const __nv_bfloat16* x1; const __nv_bfloat16* x2;
  int C1, C2, groups, act;
  long long hw;
  int px_per_cta, piece_px, n_pieces;
  const float* gamma; const float* beta;
  __nv_bfloat16* out;
  float eps;
  uint32_t off2, off_red, off_part, off_stat, off_bar;   // shared-memory layout (bytes from the aligned base)
};

__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_addr, uint32_t rank) {
  uint32_t remote;
  float v;
  asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote));
  return v;
}

__global__ void __launch_bounds__(kGfThreads)
gn_fused_cluster_kernel(const GnFusedParams p) {
  extern __shared__ __align__(128) uint8_t gf_smem[];
  const uint32_t base = (smem_u32(gf_smem) + 127u) & ~127u;
  uint8_t* gen = gf_smem + (base - smem_u32(gf_smem));
  const int C = p.C1 + p.C2, V = C >> 3, cpg = C / p.groups;
  const int R = kGfThreads / V;                      // pixel lanes of the apply pass
  const int Rs = R < kGfStatLanes ? R : kGfStatLanes;   // ... of the statistics pass (shared-memory reads: cheap)
  const int b = blockIdx.y;
  const uint32_t rank = cluster_ctarank(), nrank = cluster_nctarank();
  const long long px0 = (long long)rank * p.px_per_cta;
  const int n_px = (int)(px0 >= p.hw ? 0 : (p.hw - px0 < p.px_per_cta ? p.hw - px0 : p.px_per_cta));
  const uint32_t s1 = base, s2 = base + p.off2;
  float* red = reinterpret_cast<float*>(gen + p.off_red);      // [R][2C]
  float* part = reinterpret_cast<float*>(gen + p.off_part);    // [groups][2], read by the peers
  float* stat = reinterpret_cast<float*>(gen + p.off_stat);    // [groups][2] mean, rstd
  const uint32_t bars = base + p.off_bar;

  if (threadIdx.x == 0) {
    for (int i = 0; i < p.n_pieces; ++i) mbar_init(bars + 8u * i, 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const __nv_bfloat16* g1 = p.x1 + ((long long)b * p.hw + px0) * p.C1;
    const __nv_bfloat16* g2 = p.x2 ? p.x2 + ((long long)b * p.hw + px0) * p.C2 : nullptr;
    for (int i = 0; i < p.n_pieces; ++i) {
      const int q0 = i * p.piece_px;
      const int q = n_px - q0 < p.piece_px ? n_px - q0 : p.piece_px;
      if (q <= 0) { mbar_arrive(bars + 8u * i); continue; }
      const uint32_t b1 = (uint32_t)q * p.C1 * 2u, b2 = (uint32_t)q * p.C2 * 2u;
      mbar_arrive_expect_tx(bars + 8u * i, b1 + b2);
      bulk_load_1d(s1 + (uint32_t)q0 * p.C1 * 2u, g1 + (long long)q0 * p.C1, b1, bars + 8u * i);
      if (b2) bulk_load_1d(s2 + (uint32_t)q0 * p.C2 * 2u, g2 + (long long)q0 * p.C2, b2, bars + 8u * i);
    }
  }
  // ---- statistics: thread (r, vec) walks the pixels r, r + R, ... of every piece as it lands ----
  const int vec = threadIdx.x % V, r = threadIdx.x / V;
  const int c = vec * 8;
  const bool live = r < R;
  const uint32_t sbase = c < p.C1 ? s1 + (uint32_t)c * 2u : s2 + (uint32_t)(c - p.C1) * 2u;
  const uint32_t pitch = (uint32_t)(c < p.C1 ? p.C1 : p.C2) * 2u;
  float s[8], ss[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.f; ss[j] = 0.f; }
  bool ok = true;
  for (int i = 0; i < p.n_pieces; ++i) {
    if (ok) ok = mbar_wait(bars + 8u * i, 0);
    const int q0 = i * p.piece_px;
    const int q1 = q0 + p.piece_px < n_px ? q0 + p.piece_px : n_px;
    if (r < Rs && ok) {
      for (int px = q0 + r; px < q1; px += Rs) {
        uint4 qv;
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(qv.x), "=r"(qv.y), "=r"(qv.z), "=r"(qv.w)
                     : "r"(sbase + (uint32_t)px * pitch));
        const float2 a = unpack_bf16x2(qv.x), bb = unpack_bf16x2(qv.y), cc = unpack_bf16x2(qv.z), d = unpack_bf16x2(qv.w);
        const float v[8] = {a.x, a.y, bb.x, bb.y, cc.x, cc.y, d.x, d.y};
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j] += v[j]; ss[j] = fmaf(v[j], v[j], ss[j]); }
      }
    }
  }
  if (r < Rs) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      red[(r * C + c + j) * 2] = s[j];
      red[(r * C + c + j) * 2 + 1] = ss[j];
    }
  }
  __syncthreads();
  // per-group partial sums of this CTA, fixed order over (pixel lane, channel)
  for (int t = threadIdx.x; t < 2 * p.groups; t += kGfThreads) {
    const int g = t >> 1, which = t & 1;
    float acc = 0.f;
    for (int rr = 0; rr < Rs; ++rr)
      for (int k = 0; k < cpg; ++k) acc += red[(rr * C + g * cpg + k) * 2 + which];
    part[t] = acc;
  }
  cluster_arrive();      // release: `part` is visible to the cluster after the matching wait
  cluster_wait();
  for (int t = threadIdx.x; t < p.groups; t += kGfThreads) {
    // all remote loads in flight together (a dependent chain of 2 * nrank DSMEM round trips costs microseconds),
    // then summed in fixed rank order: every CTA computes identical statistics
    float a[16], q[16];
#pragma unroll
    for (uint32_t k = 0; k < 16; ++k) {
      a[k] = 0.f; q[k] = 0.f;
      if (k < nrank) {
        a[k] = ld_dsmem_f32(smem_u32(part + 2 * t), k);
        q[k] = ld_dsmem_f32(smem_u32(part + 2 * t + 1), k);
      }
    }
    float sum = 0.f, sq = 0.f;
#pragma unroll
    for (uint32_t k = 0; k < 16; ++k) { sum += a[k]; sq += q[k]; }
    const float inv = 1.0f / ((float)cpg * (float)p.hw);
    const float mean = sum * inv;
    float var = sq * inv - mean * mean;
    var = var > 0.f ? var : 0.f;
    stat[2 * t] = mean;
    stat[2 * t + 1] = rsqrtf(var + p.eps);
  }
  cluster_arrive();      // this CTA has finished reading its peers' partial sums
  __syncthreads();
  // ---- normalise (+SiLU) out of shared memory ----
  if (live) {
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int g = (c + j) / cpg;
      sc[j] = stat[2 * g + 1] * __ldg(p.gamma + c + j);
      sh[j] = __ldg(p.beta + c + j) - stat[2 * g] * sc[j];
    }
    __nv_bfloat16* obase = p.out + ((long long)b * p.hw + px0) * C + c;
    for (int px = r; px < n_px; px += R) {
      uint4 qv;
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(qv.x), "=r"(qv.y), "=r"(qv.z), "=r"(qv.w)
                   : "r"(sbase + (uint32_t)px * pitch));
      const float2 a = unpack_bf16x2(qv.x), bb = unpack_bf16x2(qv.y), cc = unpack_bf16x2(qv.z), d = unpack_bf16x2(qv.w);
      float v[8] = {a.x, a.y, bb.x, bb.y, cc.x, cc.y, d.x, d.y};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float y = fmaf(v[j], sc[j], sh[j]);
        v[j] = p.act == SDB_ACT_SILU ? silu_f(y) : y;
      }
      Vec8<__nv_bfloat16>::store(obase + (long long)px * C, v);
    }
  }
  cluster_wait();        // no CTA leaves (and frees its shared memory) while a peer may still be reading `part`
}

// ---- K2, one launch for images that do NOT fit on chip: persistent two-phase kernel with a grid barrier --------
// Phase 1 streams the tensor once for the statistics (unit = (image, pixel chunk); per-unit per-group partial sums,
// fixed-order reductions -> deterministic), a device-wide barrier follows (cooperative launch: every CTA is resident),
// phase 2 streams it again -- out of the 126 MB L2 when the tensor fits there -- and writes the normalised output.

struct GnCoopParams {
  const __nv_bfloat16* x1; const __nv_bfloat16* x2;
  int C1, C2, groups, act;
  long long hw;
  int B, K, unit_px;            // K pixel chunks per image, unit_px pixels each
  const float* gamma; const float* beta;
  __nv_bfloat16* out;
  float eps;
  float* partial;               // [B][K][groups][2]
  unsigned int* bar;            // [2]: arrival counter (self-resetting), generation
};

__device__ __forceinline__ void grid_barrier(unsigned int* bar, unsigned int n_ctas) {
  __syncthreads();
  if (threadIdx.x == 0) {
    volatile unsigned int* gen_p = bar + 1;
    const unsigned int gen = *gen_p;
    __threadfence();
    if (atomicAdd(bar, 1u) == n_ctas - 1) {
      bar[0] = 0u;                       // last arrival: reset for the next launch, then open the barrier
      __threadfence();
      atomicAdd(bar + 1, 1u);
    } else {
      while (*gen_p == gen) __nanosleep(64);
    }
    __threadfence();
  }
  __syncthreads();
}

template <int kT>
__global__ void __launch_bounds__(kT)
gn_coop_kernel(const GnCoopParams p) {
  extern __shared__ float gc_sm[];       // [R][2C] reduction scratch | [groups][2] stats
  const int C = p.C1 + p.C2, V = C >> 3, cpg = C / p.groups;
  const int R = kT / V;
  const int vec = threadIdx.x % V, r = threadIdx.x / V;
  const int c = vec * 8;
  const bool live = r < R;
  float* red = gc_sm;
  size_t red_floats = (size_t)R * 2 * C;
  if (red_floats < (size_t)p.K * p.groups * 2) red_floats = (size_t)p.K * p.groups * 2;
  float* stat = gc_sm + red_floats;
  const int units = p.B * p.K;
  // ---------------- phase 1: statistics ----------------
  for (int u = blockIdx.x; u < units; u += gridDim.x) {
    const int b = u / p.K, k = u - b * p.K;
    const long long p0 = (long long)k * p.unit_px;
    const long long p1 = p0 + p.unit_px < p.hw ? p0 + p.unit_px : p.hw;
    float s[8], ss[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] = 0.f; ss[j] = 0.f; }
    if (live) {
      const __nv_bfloat16* base;
      long long pitch;
      if (c < p.C1) { base = p.x1 + (long long)b * p.hw * p.C1 + c; pitch = p.C1; }
      else { base = p.x2 + (long long)b * p.hw * p.C2 + (c - p.C1); pitch = p.C2; }
      long long px = p0 + r;
      // eight 16-byte loads in flight per thread: with two 256-thread CTAs per SM the statistics pass is bound by the
      // bytes in flight (4 loads: 30 KB per SM against the ~66 KB that the HBM rate x latency asks for)
      for (; px + 7LL * R < p1; px += 8LL * R) {
        uint4 q[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = __ldg(reinterpret_cast<const uint4*>(base + (px + (long long)i * R) * pitch));
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t w[4] = {q[i].x, q[i].y, q[i].z, q[i].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float lo = __uint_as_float(w[j] << 16), hi = __uint_as_float(w[j] & 0xffff0000u);
            s[2 * j] += lo;
            s[2 * j + 1] += hi;
            ss[2 * j] = fmaf(lo, lo, ss[2 * j]);
            ss[2 * j + 1] = fmaf(hi, hi, ss[2 * j + 1]);
          }
        }
      }
      for (; px + 3LL * R < p1; px += 4LL * R) {
        float v0[8], v1[8], v2[8], v3[8];
        Vec8<__nv_bfloat16>::load(base + px * pitch, v0);
        Vec8<__nv_bfloat16>::load(base + (px + R) * pitch, v1);
        Vec8<__nv_bfloat16>::load(base + (px + 2LL * R) * pitch, v2);
        Vec8<__nv_bfloat16>::load(base + (px + 3LL * R) * pitch, v3);
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
        Vec8<__nv_bfloat16>::load(base + px * pitch, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j] += v[j]; ss[j] = fmaf(v[j], v[j], ss[j]); }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        red[(r * C + c + j) * 2] = s[j];
        red[(r * C + c + j) * 2 + 1] = ss[j];
      }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < 2 * p.groups; t += kT) {
      const int g = t >> 1, which = t & 1;
      float acc = 0.f;
      for (int rr = 0; rr < R; ++rr)
        for (int kk = 0; kk < cpg; ++kk) acc += red[(rr * C + g * cpg + kk) * 2 + which];
      p.partial[((long long)u * p.groups) * 2 + t] = acc;
    }
    __syncthreads();
  }
  grid_barrier(p.bar, gridDim.x);
  // ---------------- phase 2: normalise (+SiLU) ----------------
  for (int u = blockIdx.x; u < units; u += gridDim.x) {
    const int b = u / p.K, k = u - b * p.K;
    // the image's K x groups x 2 partial sums: fetched by all threads at once (independent L2 loads), then summed
    // per group in fixed chunk order
    {
      const float* pp = p.partial + (long long)b * p.K * p.groups * 2;
      const int n = p.K * p.groups * 2;
      for (int t = threadIdx.x; t < n; t += kT) red[t] = __ldcg(pp + t);
    }
    __syncthreads();
    for (int t = threadIdx.x; t < p.groups; t += kT) {
      float sum = 0.f, sq = 0.f;
      for (int kk = 0; kk < p.K; ++kk) {          // fixed chunk order
        sum += red[(kk * p.groups + t) * 2];
        sq += red[(kk * p.groups + t) * 2 + 1];
      }
      const float inv = 1.0f / ((float)cpg * (float)p.hw);
      const float mean = sum * inv;
      float var = sq * inv - mean * mean;
      var = var > 0.f ? var : 0.f;
      stat[2 * t] = mean;
      stat[2 * t + 1] = rsqrtf(var + p.eps);
    }
    __syncthreads();
    if (live) {
      float sc[8], sh[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int g = (c + j) / cpg;
        sc[j] = stat[2 * g + 1] * __ldg(p.gamma + c + j);
        sh[j] = __ldg(p.beta + c + j) - stat[2 * g] * sc[j];
      }
      const __nv_bfloat16* base;
      long long pitch;
      if (c < p.C1) { base = p.x1 + (long long)b * p.hw * p.C1 + c; pitch = p.C1; }
      else { base = p.x2 + (long long)b * p.hw * p.C2 + (c - p.C1); pitch = p.C2; }
      __nv_bfloat16* obase = p.out + (long long)b * p.hw * C + c;
      const long long p0 = (long long)k * p.unit_px;
      const long long p1 = p0 + p.unit_px < p.hw ? p0 + p.unit_px : p.hw;
      auto finish = [&](float (&v)[8], long long px) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float y = fmaf(v[j], sc[j], sh[j]);
          v[j] = p.act == SDB_ACT_SILU ? silu_f(y) : y;
        }
        Vec8<__nv_bfloat16>::store(obase + px * C, v);
      };
      long long px = p0 + r;
      for (; px + 3LL * R < p1; px += 4LL * R) {
        float v0[8], v1[8], v2[8], v3[8];
        Vec8<__nv_bfloat16>::load(base + px * pitch, v0);
        Vec8<__nv_bfloat16>::load(base + (px + R) * pitch, v1);
        Vec8<__nv_bfloat16>::load(base + (px + 2LL * R) * pitch, v2);
        Vec8<__nv_bfloat16>::load(base + (px + 3LL * R) * pitch, v3);
        finish(v0, px);
        finish(v1, px + R);
        finish(v2, px + 2LL * R);
        finish(v3, px + 3LL * R);
      }
      for (; px < p1; px += R) {
        float v[8];
        Vec8<__nv_bfloat16>::load(base + px * pitch, v);
        finish(v, px);
      }
    }
    __syncthreads();
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

// Single-pass GroupNorm(+SiLU) (gn_fused_cluster_kernel): bf16 in / bf16 out, the image must fit the shared memory
// of a cluster of <= 16 CTAs.  Returns SDB_EUNSUPPORTED (without an error message being an error: the caller then
// runs sdb_groupnorm_stats + sdb_groupnorm_apply) when the shape does not qualify.
extern "C" int sdb_groupnorm_fused(const void* x1, int C1, const void* x2, int C2, const float* gamma,
                                   const float* beta, int act, void* out, int B, int64_t hw, int groups, float eps,
                                   void* stream) {
  if (!x2) C2 = 0;
  const int C = C1 + C2;
  SDB_CHECK_ARG(x1 && gamma && beta && out && B > 0 && hw > 0, "groupnorm_fused: bad args");
  SDB_CHECK_ARG(C1 % 8 == 0 && C2 % 8 == 0 && groups > 0 && C % groups == 0, "groupnorm_fused: bad channels");
  SDB_CHECK_ARG(((uintptr_t)x1 & 15) == 0 && ((uintptr_t)x2 & 15) == 0 && ((uintptr_t)out & 15) == 0,
                "groupnorm_fused: pointers must be 16-byte aligned");
  const int V = C / 8;
  if (V > kGfThreads || B > 65535) { sdb_set_error("groupnorm_fused: shape not supported"); return SDB_EUNSUPPORTED; }
  const int R = kGfThreads / V < kGfStatLanes ? kGfThreads / V : kGfStatLanes;
  const size_t fixed = (size_t)R * 2 * C * 4 + (size_t)groups * 16 + kGfMaxPieces * 8 + 512;
  const size_t row_bytes = (size_t)C * 2;
  // smallest cluster whose slice fits; prefer slices that let two CTAs share an SM (overlap of load / compute / store)
  // (SDB_GN_FUSED_MAXS caps the cluster size: measurement aid)
  static int max_s = -1;
  if (max_s < 0) { const char* e = getenv("SDB_GN_FUSED_MAXS"); max_s = e ? atoi(e) : 16; }
  int S = 0;
  for (int pass = 0; pass < 2 && !S; ++pass) {
    const size_t budget = pass == 0 ? 112 * 1024 : 225 * 1024;
    for (int cand = 1; cand <= max_s; cand *= 2) {
      const long long px = (hw + cand - 1) / cand;
      if ((size_t)px * row_bytes + fixed <= budget) { S = cand; break; }
    }
  }
  if (!S) { sdb_set_error("groupnorm_fused: image does not fit a 16-CTA cluster's shared memory"); return SDB_EUNSUPPORTED; }
  GnFusedParams p;
  memset(&p, 0, sizeof(p));
  p.x1 = (const __nv_bfloat16*)x1; p.x2 = (const __nv_bfloat16*)x2;
  p.C1 = C1; p.C2 = C2; p.groups = groups; p.act = act; p.hw = hw;
  p.px_per_cta = (int)((hw + S - 1) / S);
  p.n_pieces = p.px_per_cta >= 4 * kGfMaxPieces ? kGfMaxPieces : (p.px_per_cta >= 8 ? 4 : 1);
  p.piece_px = (p.px_per_cta + p.n_pieces - 1) / p.n_pieces;
  p.gamma = gamma; p.beta = beta; p.out = (__nv_bfloat16*)out; p.eps = eps;
  auto al = [](size_t v) { return (uint32_t)((v + 127) & ~(size_t)127); };
  p.off2 = al((size_t)p.px_per_cta * C1 * 2);
  p.off_red = p.off2 + al((size_t)p.px_per_cta * C2 * 2);
  p.off_part = p.off_red + al((size_t)R * 2 * C * 4);
  p.off_stat = p.off_part + al((size_t)groups * 8);
  p.off_bar = p.off_stat + al((size_t)groups * 8);
  const size_t smem = (size_t)p.off_bar + kGfMaxPieces * 8 + 128;
  static bool attr_set[64] = {false};
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    SDB_CUDA_OK(cudaFuncSetAttribute(gn_fused_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    SDB_CUDA_OK(cudaFuncSetAttribute(gn_fused_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)S, (unsigned)B, 1);
  cfg.blockDim = dim3(kGfThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = sdb_stream(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)S;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  SDB_CUDA_OK(cudaLaunchKernelEx(&cfg, gn_fused_cluster_kernel, p));
  SDB_LAUNCH_OK();
  return SDB_OK;
}

// debug aid: how many clusters of S CTAs with `smem` bytes each the device can hold at once
extern "C" int sdb_debug_gn_fused_clusters(int S, int smem) {
  cudaFuncSetAttribute(gn_fused_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  cudaFuncSetAttribute(gn_fused_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)S, 64, 1);
  cfg.blockDim = dim3(kGfThreads, 1, 1);
  cfg.dynamicSmemBytes = (size_t)smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)S;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = -1;
  if (cudaOccupancyMaxActiveClusters(&n, gn_fused_cluster_kernel, &cfg) != cudaSuccess) return -1;
  return n;
}

// One-launch GroupNorm(+SiLU) for any image size: gn_coop_kernel (cooperative launch, grid barrier between the
// statistics and the normalise phases).  bf16 in / bf16 out.  `partial`: fp32 scratch of sdb_groupnorm_coop_scratch(B,
// hw, groups) floats; `bar`: two zero-initialised uint32 words that the kernel leaves zeroed / advanced -- one pair per
// stream of calls (launches sharing a pair must not overlap on the GPU).
static int gn_coop_per_sm() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SDB_GN_COOP_PER_SM"); v = e ? atoi(e) : 2; if (v < 1 || v > 8) v = 2; }
  return v;
}
static int gn_coop_chunks(int B, long long hw) {
  // as many units as fit the machine in ONE wave (B * K <= CTAs resident; 2 CTAs of 256 threads per SM measured best:
  // the grid barrier gets dearer with every CTA), at least 64 pixels each
  const int sms = sdb_num_sms();
  long long k = ((long long)gn_coop_per_sm() * sms) / B;
  const long long kmax = (hw + 63) / 64;
  if (k > kmax) k = kmax;
  if (k < 1) k = 1;
  return (int)k;
}
extern "C" int64_t sdb_groupnorm_coop_scratch(int B, int64_t hw, int groups) {
  return (int64_t)B * gn_coop_chunks(B, hw) * groups * 2;
}
extern "C" int sdb_groupnorm_coop(const void* x1, int C1, const void* x2, int C2, const float* gamma, const float* beta,
                                  int act, void* out, int B, int64_t hw, int groups, float eps, float* partial,
                                  uint32_t* bar, void* stream) {
  if (!x2) C2 = 0;
  const int C = C1 + C2;
  SDB_CHECK_ARG(x1 && gamma && beta && out && partial && bar && B > 0 && hw > 0, "groupnorm_coop: bad args");
  SDB_CHECK_ARG(C1 % 8 == 0 && C2 % 8 == 0 && groups > 0 && C % groups == 0, "groupnorm_coop: bad channels");
  SDB_CHECK_ARG(((uintptr_t)x1 & 15) == 0 && ((uintptr_t)x2 & 15) == 0 && ((uintptr_t)out & 15) == 0,
                "groupnorm_coop: pointers must be 16-byte aligned");
  const int V = C / 8;
  static int threads = -1;
  if (threads < 0) {
    const char* e = getenv("SDB_GN_COOP_THREADS");
    threads = e ? atoi(e) : 256;
    if (threads != 256 && threads != 512 && threads != 1024) threads = 256;
  }
  const int kGcThreads = threads;
  SDB_CHECK_ARG(V <= kGcThreads, "groupnorm_coop: C (%d) too large", C);
  GnCoopParams p;
  memset(&p, 0, sizeof(p));
  p.x1 = (const __nv_bfloat16*)x1; p.x2 = (const __nv_bfloat16*)x2;
  p.C1 = C1; p.C2 = C2; p.groups = groups; p.act = act; p.hw = hw; p.B = B;
  p.K = gn_coop_chunks(B, hw);
  p.unit_px = (int)((hw + p.K - 1) / p.K);
  p.K = (int)((hw + p.unit_px - 1) / p.unit_px);
  p.gamma = gamma; p.beta = beta; p.out = (__nv_bfloat16*)out; p.eps = eps;
  p.partial = partial; p.bar = bar;
  const int R = kGcThreads / V;
  size_t red_floats = (size_t)R * 2 * C;
  if (red_floats < (size_t)p.K * groups * 2) red_floats = (size_t)p.K * groups * 2;
  const size_t smem = (red_floats + 2 * (size_t)groups) * sizeof(float);
  const void* fn = threads == 256 ? (const void*)gn_coop_kernel<256>
                   : threads == 512 ? (const void*)gn_coop_kernel<512> : (const void*)gn_coop_kernel<1024>;
  static int occ_cache[64][2] = {{0}};    // per device: (smem it was computed for, CTAs per SM)
  int dev = 0;
  SDB_CUDA_OK(cudaGetDevice(&dev));
  int per_sm = 0;
  if (dev >= 0 && dev < 64 && occ_cache[dev][0] == (int)smem + 1) per_sm = occ_cache[dev][1];
  if (!per_sm) {
    if (smem > 48 * 1024)
      SDB_CUDA_OK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    SDB_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kGcThreads, smem));
    if (per_sm < 1) { sdb_set_error("groupnorm_coop: kernel does not fit an SM"); return SDB_EUNSUPPORTED; }
    if (dev >= 0 && dev < 64) { occ_cache[dev][0] = (int)smem + 1; occ_cache[dev][1] = per_sm; }
  }
  if (per_sm > gn_coop_per_sm()) per_sm = gn_coop_per_sm();
  const long long units = (long long)B * p.K;
  long long grid = (long long)per_sm * sdb_num_sms();
  if (grid > units) grid = units;
  void* args[] = {(void*)&p};
  SDB_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3((unsigned)grid), dim3(kGcThreads), args, smem, sdb_stream(stream)));
  SDB_LAUNCH_OK();
  return SDB_OK;
}

static int gn_apply_launch(const void* x1, int C1, const void* x2, int C2, int dtype, const float* stats,
                           const void* cs1, const void* cs2, float eps, const float* gamma, const float* beta,
                           int act, void* out, int out_dtype, int B, int64_t hw, int groups, void* stream) {
  if (!x2) C2 = 0;
  const int C = C1 + C2;
  SDB_CHECK_ARG(x1 && (stats || cs1) && gamma && beta && out && B > 0 && hw > 0, "groupnorm_apply: bad args");
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
                                                     beta, act, (O*)out, hw, groups, chunk_px, cs1, cs2, eps)
  if (dtype == SDB_BF16 && out_dtype == SDB_BF16) SDB_GN_APPLY(__nv_bfloat16, __nv_bfloat16);
  else if (dtype == SDB_BF16) SDB_GN_APPLY(__nv_bfloat16, float);
  else if (out_dtype == SDB_BF16) SDB_GN_APPLY(float, __nv_bfloat16);
  else SDB_GN_APPLY(float, float);
#undef SDB_GN_APPLY
  SDB_LAUNCH_OK();
  return SDB_OK;
}

extern "C" int sdb_groupnorm_apply(const void* x1, int C1, const void* x2, int C2, int dtype,
                                   const float* stats, const float* gamma, const float* beta, int act,
                                   void* out, int out_dtype, int B, int64_t hw, int groups,
                                   void* stream) {
  SDB_CHECK_ARG(stats, "groupnorm_apply: stats is NULL");
  return gn_apply_launch(x1, C1, x2, C2, dtype, stats, nullptr, nullptr, 0.f, gamma, beta, act, out, out_dtype, B, hw,
                         groups, stream);
}

extern "C" int sdb_groupnorm_apply_sums(const void* x1, int C1, const void* sums1, const void* x2, int C2,
                                        const void* sums2, int dtype, const float* gamma, const float* beta,
                                        float eps, int act, void* out, int out_dtype, int B, int64_t hw, int groups,
                                        void* stream) {
  SDB_CHECK_ARG(sums1 && (!x2 || sums2), "groupnorm_apply_sums: every input tensor needs its channel sums");
  SDB_CHECK_ARG(groups > 0 && groups <= kGnMaxGroupsSums, "groupnorm_apply_sums: groups (%d) must be <= %d", groups,
                kGnMaxGroupsSums);
  SDB_CHECK_ARG(((uintptr_t)sums1 & 15) == 0 && ((uintptr_t)sums2 & 15) == 0, "groupnorm_apply_sums: sums must be 16-byte aligned");
  return gn_apply_launch(x1, C1, x2, C2, dtype, nullptr, sums1, x2 ? sums2 : nullptr, eps, gamma, beta, act, out,
                         out_dtype, B, hw, groups, stream);
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
