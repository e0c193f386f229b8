// Weight-gradient GEMM on the tensor cores (row T1: autograd of nn.Conv2d 3x3 / nn.Linear inside
// train_unet.py:359-375 `accelerator.backward(loss)`).
//
//   dW[n, tap, c] = sum over pixels  dY[pixel, n] * X[pixel + tap, c]          (conv, 9 taps, zero padding)
//   dW[n, c]      = sum over rows    dY[row, n]   * X[row, c]                  (linear)
//
// The reduction runs over the PIXEL dimension (262 144 at config 4), the output is tiny (Cout x 9 Cin).  Both
// operands are consumed exactly as they lie in HBM -- channels-last rows -- as MN-MAJOR tcgen05 operands:
// a TMA box of 64 pixels x 64 channels lands as 64 rows of 128 B (SWIZZLE_128B), which is one UMMA "MN atom" per
// 8 pixels; A = dY^T (128 output channels = two atoms), B = X^T shifted by the tap through the TMA coordinates
// (out-of-image pixels are zero-filled = the conv padding).  No transposes, no im2col.
// One CTA = (128 output channels) x (a group of up to 3 (tap, Cin-slice) accumulators side by side in TMEM, so the
// dY tile is fetched once per group) x (a chunk of the pixels): split-K over one wave of CTAs, partial sums added to the
// fp32 result with atomics (the caller zeroes it).  Warp 0 = TMA producer, warp 1 = MMA issuer, warps 4-7 = epilogue.
#include "common.cuh"
#include "umma_helpers.cuh"

namespace sdb {

constexpr int kWgThreads = 256;
constexpr int kWgMaxStages = 4;
constexpr int kWgAtomBytes = 64 * 128;   // 64 pixel rows x 64 channels bf16

struct WgradParams {
  CUtensorMap tmY, tmX;
  int linear;         // 1: operands are 2-D [rows, C]
  int taps;           // 1 | 9
  int NW;             // accumulator width (input channels per sub-tile): 160 | 128 | 64
  int atoms_b;        // 64-channel atoms fetched per sub-tile
  int subs_per_tap;   // Cin / NW
  int total_subs;     // taps * subs_per_tap
  int group;          // sub-tiles per CTA
  int n_groups, chunks, kb_total, kb_per_chunk;
  int TH, TN, tiles_h;   // conv K-block = TN images x TH rows x W pixels = 64 pixels
  int Cout, Cin;
  float* out;
  long long ld_n;     // out[n * ld_n + tap * Cin + c]
  int stage_bytes, stages;
};

// MN-major operand, SWIZZLE_128B: 64 channels (128 B) per pixel row, 8-pixel groups 1024 B apart (SBO),
// 64-channel atoms 8 KB apart (LBO)
__device__ __forceinline__ uint64_t umma_smem_desc_mn128(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3ffff) >> 4);
  d |= static_cast<uint64_t>(kWgAtomBytes >> 4) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;
  return d;
}

__global__ void __launch_bounds__(kWgThreads, 1) wgrad_umma_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = base + (uint32_t)p.stages * (uint32_t)p.stage_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kWgMaxStages + s); };
  const uint32_t tfull_bar = bar_base + 8u * (2 * kWgMaxStages);
  const uint32_t tmem_slot = bar_base + 8u * (2 * kWgMaxStages + 1);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // job = (n-tile, group, chunk)
  const int chunk = (int)blockIdx.x % p.chunks;
  const int rest = (int)blockIdx.x / p.chunks;
  const int g = rest % p.n_groups;
  const int nt = rest / p.n_groups;
  const int kb0 = chunk * p.kb_per_chunk;
  const int kb1 = min(p.kb_total, kb0 + p.kb_per_chunk);
  const int nkb = kb1 - kb0;
  const int sub0 = g * p.group;
  const int nsub = min(p.group, p.total_subs - sub0);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmY);
    prefetch_tmap(&p.tmX);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0 || warp == 2 || warp == 3) {
    // ===================== TMA producers =====================
    // A stage is 2 + nsub * atoms_b boxes (up to 11).  One thread issuing them all, with the sub-tile -> (tap, channel)
    // divisions in the loop, took about as long as the stage's MMAs; so warp 0 posts the byte count and fetches the dY
    // atoms, warps 2 and 3 share the X atoms (even and odd sub-tiles), each with its coordinates precomputed.
    const uint32_t tx = (uint32_t)(2 + nsub * p.atoms_b) * (uint32_t)kWgAtomBytes;
    const int s_first = warp == 0 ? -1 : warp - 2;      // warp 0: no X atoms
    int xc[4], xdh[4], xdw[4], xs[4], nx = 0;          // (512 TMEM columns / NW >= 64: at most 8 sub-tiles)
    for (int sidx = s_first; s_first >= 0 && sidx < nsub && nx < 4; sidx += 2) {
      const int sub = sub0 + sidx;
      const int tap = sub / p.subs_per_tap;
      xs[nx] = sidx;
      xc[nx] = (sub - tap * p.subs_per_tap) * p.NW;
      xdh[nx] = p.taps == 9 ? tap / 3 - 1 : 0;
      xdw[nx] = p.taps == 9 ? tap - (tap / 3) * 3 - 1 : 0;
      ++nx;
    }
    bool ok = true;
    int stage = 0;
    uint32_t phase = 0;
    int ig = p.linear ? 0 : kb0 / p.tiles_h, hb = p.linear ? 0 : kb0 - ig * p.tiles_h;   // (image group, row block) of kb
    for (int i = 0; i < nkb && ok && (warp == 0 || nx > 0); ++i) {
      ok = mbar_wait(empty_bar(stage), phase ^ 1u);
      if (!ok) break;
      const int kb = kb0 + i;
      int w0 = 0, h0 = 0, i0 = 0;
      if (p.linear) {
        w0 = kb * 64;
      } else {
        h0 = hb * p.TH;
        i0 = ig * p.TN;
        if (++hb == p.tiles_h) { hb = 0; ++ig; }
      }
      const uint32_t sa = base + (uint32_t)stage * (uint32_t)p.stage_bytes;
      const uint32_t sb = sa + 2u * kWgAtomBytes;
      if (elect_one()) {
        if (warp == 0) {
          mbar_arrive_expect_tx(full_bar(stage), tx);
          tma_load_4d(sa, &p.tmY, full_bar(stage), nt * 128, w0, h0, i0);
          tma_load_4d(sa + kWgAtomBytes, &p.tmY, full_bar(stage), nt * 128 + 64, w0, h0, i0);
        } else {
          for (int x = 0; x < nx; ++x)
            for (int a = 0; a < p.atoms_b; ++a)
              tma_load_4d(sb + (uint32_t)(xs[x] * p.atoms_b + a) * kWgAtomBytes, &p.tmX, full_bar(stage), xc[x] + a * 64,
                          w0 + xdw[x], h0 + xdh[x], i0);
        }
      }
      __syncwarp();
      if (++stage == p.stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = umma_idesc_bf16(128u, (uint32_t)p.NW) | (1u << 15) | (1u << 16);   // A and B MN-major
    bool ok = true;
    int stage = 0;
    uint32_t phase = 0;
    for (int i = 0; i < nkb && ok; ++i) {
      ok = mbar_wait(full_bar(stage), phase);
      if (!ok) break;
      tc_fence_after();
      const uint32_t sa = base + (uint32_t)stage * (uint32_t)p.stage_bytes;
      const uint32_t sb = sa + 2u * kWgAtomBytes;
      const uint64_t adesc = umma_smem_desc_mn128(sa);
      if (elect_one()) {
        for (int s = 0; s < nsub; ++s) {
          const uint64_t bdesc = umma_smem_desc_mn128(sb + (uint32_t)(s * p.atoms_b) * kWgAtomBytes);
          const uint32_t tmem_d = tmem_base + (uint32_t)(s * p.NW);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)   // 16 pixels = 2048 B per MMA: +128 in the (addr >> 4) field
            umma_f16(tmem_d, adesc + (uint64_t)(kk * 128), bdesc + (uint64_t)(kk * 128), idesc, (i | kk) ? 1u : 0u);
        }
        umma_commit(empty_bar(stage));
      }
      __syncwarp();
      if (++stage == p.stages) { stage = 0; phase ^= 1u; }
    }
    if (ok && elect_one()) umma_commit(tfull_bar);
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> fp32 atomics =====================
    const int q = warp - 4;   // == warp % 4: the TMEM lane quadrant this warp may read
    if (mbar_wait(tfull_bar, 0)) {
      tc_fence_after();
      const int n = nt * 128 + q * 32 + lane;
      const bool vec4 = (p.ld_n & 3) == 0 && (p.Cin & 3) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;
      for (int s = 0; s < nsub; ++s) {
        const int sub = sub0 + s;
        const int tap = sub / p.subs_per_tap;
        const int c0 = (sub - tap * p.subs_per_tap) * p.NW;
        float* orow = p.out + (long long)n * p.ld_n + (long long)tap * p.Cin + c0;
        for (int cc = 0; cc < p.NW; cc += 16) {
          uint32_t v[16];
          tmem_ld_32x16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(s * p.NW + cc), v);
          tmem_ld_wait();
          if (n < p.Cout) {
            if (vec4) {       // one 16-byte reduction per four columns (a quarter of the L2 atomic requests)
#pragma unroll
              for (int j = 0; j < 16; j += 4)
                atomicAdd(reinterpret_cast<float4*>(orow + cc + j),
                          make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                      __uint_as_float(v[j + 3])));
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) atomicAdd(orow + cc + j, __uint_as_float(v[j]));
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

static int wgrad_pick_nw(int Cin) {
  if (Cin % 160 == 0) return 160;
  if (Cin % 128 == 0) return 128;
  if (Cin % 64 == 0) return 64;
  return 0;
}

static int wgrad_launch(WgradParams& p, int kb_total, cudaStream_t st) {
  p.atoms_b = (p.NW + 63) / 64;
  p.subs_per_tap = p.Cin / p.NW;
  p.total_subs = p.taps * p.subs_per_tap;
  // accumulators per CTA: bounded by the 512 TMEM columns and by a double-buffered operand ring in 220 KB.
  // Measured (tools/bench_wgrad.py, conv 640->320): 3 accumulators / 2 stages 763 TFLOP/s, 2 / 3 stages 580,
  // 1 / 4 stages 438 -- re-fetching dY costs more than a deeper ring gains.  (SDB_WGRAD_GROUP: tuning aid)
  int group = 512 / p.NW;
  while (group > 1 && 2 * (2 + group * p.atoms_b) * kWgAtomBytes > 220 * 1024) --group;
  static int group_env = -1;
  if (group_env < 0) { const char* e = getenv("SDB_WGRAD_GROUP"); group_env = e ? atoi(e) : 0; }
  if (group_env > 0 && group_env * p.NW <= 512) group = group_env;
  if (group > p.total_subs) group = p.total_subs;
  p.group = group;
  p.n_groups = (p.total_subs + group - 1) / group;
  const int n_tiles = (p.Cout + 127) / 128;
  const int base_jobs = n_tiles * p.n_groups;
  // one CTA per SM (176 KB of operand ring): exactly one wave, rounded DOWN so no nearly-empty tail wave
  int chunks = sdb_num_sms() / base_jobs;
  if (chunks > kb_total) chunks = kb_total;
  if (chunks < 1) chunks = 1;
  p.kb_total = kb_total;
  p.kb_per_chunk = (kb_total + chunks - 1) / chunks;
  p.chunks = (kb_total + p.kb_per_chunk - 1) / p.kb_per_chunk;
  p.stage_bytes = (2 + group * p.atoms_b) * kWgAtomBytes;
  p.stages = (220 * 1024) / p.stage_bytes;
  if (p.stages > kWgMaxStages) p.stages = kWgMaxStages;
  SDB_CHECK_ARG(p.stages >= 2, "wgrad: operand ring does not fit");
  const int smem = p.stages * p.stage_bytes + 128 + 1024;
  SDB_CHECK_ARG(smem <= 227 * 1024, "wgrad: smem budget exceeded (%d B)", smem);
  static int configured = 0;
  if (configured < smem) {
    SDB_CUDA_OK(cudaFuncSetAttribute(wgrad_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = 227 * 1024;
  }
  wgrad_umma_kernel<<<n_tiles * p.n_groups * p.chunks, kWgThreads, smem, st>>>(p);
  SDB_LAUNCH_OK();
  return SDB_OK;
}

}  // namespace sdb

extern "C" int sdb_wgrad_conv3x3(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                                 void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(dy && x && dw && B > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0, "wgrad_conv3x3: bad args");
  WgradParams p{};
  p.NW = wgrad_pick_nw(Cin);
  const bool geom_ok = W <= 64 && 64 % W == 0;
  int TH = geom_ok ? 64 / W : 1;
  if (TH > H) TH = H;
  if (!(p.NW && geom_ok && H % TH == 0 && 64 % (W * TH) == 0 && Cout % 8 == 0)) {
    sdb_set_error("wgrad_conv3x3: shape not built (W = %d must divide 64, Cin = %d a multiple of 64, Cout = %d of 8)",
                  W, Cin, Cout);
    return SDB_EUNSUPPORTED;
  }
  p.linear = 0; p.taps = 9; p.TH = TH; p.TN = 64 / (W * TH); p.tiles_h = H / TH;
  p.Cout = Cout; p.Cin = Cin; p.out = dw; p.ld_n = 9LL * Cin;
  const uint32_t box[4] = {64u, (uint32_t)W, (uint32_t)TH, (uint32_t)p.TN};
  {
    const uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t str[4] = {2, (uint64_t)Cout * 2, (uint64_t)W * Cout * 2, (uint64_t)H * W * Cout * 2};
    int rc = sdb_make_tensor_map(&p.tmY, dy, 2, 4, dims, str, box, 128);
    if (rc) return rc;
  }
  {
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t str[4] = {2, (uint64_t)Cin * 2, (uint64_t)W * Cin * 2, (uint64_t)H * W * Cin * 2};
    int rc = sdb_make_tensor_map(&p.tmX, x, 2, 4, dims, str, box, 128);
    if (rc) return rc;
  }
  const int img_groups = (B + p.TN - 1) / p.TN;
  return wgrad_launch(p, img_groups * p.tiles_h, sdb_stream(stream));
}

extern "C" int sdb_wgrad_linear(const void* dy, int64_t lddy, const void* x, int64_t ldx, float* dw, int64_t lddw,
                                int M, int N, int K, void* stream) {
  using namespace sdb;
  SDB_CHECK_ARG(dy && x && dw && M > 0 && N > 0 && K > 0, "wgrad_linear: bad args");
  WgradParams p{};
  p.NW = wgrad_pick_nw(K);
  if (!(p.NW && lddy % 8 == 0 && ldx % 8 == 0 && ((uintptr_t)dy & 15) == 0 && ((uintptr_t)x & 15) == 0)) {
    sdb_set_error("wgrad_linear: shape not built (K = %d must be a multiple of 64, row strides of 8 elements)", K);
    return SDB_EUNSUPPORTED;
  }
  p.linear = 1; p.taps = 1; p.TH = 1; p.TN = 1; p.tiles_h = 1;
  p.Cout = N; p.Cin = K; p.out = dw; p.ld_n = lddw;
  const uint32_t box[4] = {64u, 64u, 1u, 1u};
  {
    const uint64_t dims[4] = {(uint64_t)N, (uint64_t)M, 1, 1};
    const uint64_t str[4] = {2, (uint64_t)lddy * 2, (uint64_t)lddy * 2 * M, (uint64_t)lddy * 2 * M};
    int rc = sdb_make_tensor_map(&p.tmY, dy, 2, 4, dims, str, box, 128);
    if (rc) return rc;
  }
  {
    const uint64_t dims[4] = {(uint64_t)K, (uint64_t)M, 1, 1};
    const uint64_t str[4] = {2, (uint64_t)ldx * 2, (uint64_t)ldx * 2 * M, (uint64_t)ldx * 2 * M};
    int rc = sdb_make_tensor_map(&p.tmX, x, 2, 4, dims, str, box, 128);
    if (rc) return rc;
  }
  return wgrad_launch(p, (M + 63) / 64, sdb_stream(stream));
}
