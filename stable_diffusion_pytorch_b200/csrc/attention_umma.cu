// K4 (tensor-core path): flash-style attention for small head dims on tcgen05 / TMEM / TMA.
// (reference: modules/transformer.py:93-114 -- softmax(Q K^T * scale) V per (batch, head))
//
// The UNet's head dim is 20 (SURVEY top item 7): per (query, key) pair there are 80 useful flops
// and one exp, so the kernel is bound by the MUFU ex2 pipe (16/clk/SM measured) and the issue
// slots of the softmax warps, never by the tensor pipe.  Design:
//   * operands arrive HEAD-PADDED from the projection GEMM: head h of Q / K / V lives at columns
//     [32h, 32h+32) with zeros beyond d for Q and K, and V carries a column of ones at index d
//     so that the P.V product also yields the softmax denominator (no per-element adds);
//   * one CTA = 128 queries of one (batch, head) (attention_umma_wg1.cu): ONE softmax warpgroup, one query row per
//     thread -> no shuffles, one TMA lane, one MMA-issuing lane; four CTAs share an SM;
//   * S = Q K^T (128 x 64 fp32) is produced in TMEM by two K=16 tcgen05.mma, read back with tcgen05.ld,
//     exponentiated, written back to TMEM as packed bf16 P and consumed directly as the A operand of the P.V
//     tcgen05.mma (B = V tile, MN-major straight from the TMA box);
//   * 2/8 of the exponentials are evaluated on the FMA pipe (Cody-Waite split + degree-3 polynomial in packed
//     f32x2 arithmetic), spread over each group of eight pairs;
//   * O (128 x 32 fp32) stays in TMEM across key tiles and is rescaled only when a row maximum grows by more than
//     2^8 (lazy rescale).
// This file holds the parameter block, the constants the helpers share and the host-side dispatch; the retired
// two-warpgroup kernel and the measured variants live under tools/experiments/.
#include <stdlib.h>

#include "common.cuh"
#include "umma_helpers.cuh"

namespace sdb {

struct FaParams {
  CUtensorMap tmQ, tmK, tmV;
  __nv_bfloat16* out;
  long long ldo;
  int n, m, d, heads;
  float scale_log2;
  float* lse;   // optional [B, heads, n]: log2-domain log-sum-exp of the scaled scores (kept for the backward)
};

}  // namespace sdb

#include "attention_umma_wg1.cu"   // the one-warpgroup-per-CTA shape of the same pipeline (uses the helpers above)
#include "attention_ctx_umma.cu"   // short context (<= 96 keys): K / V resident, query tiles streamed
#include "attention_d128_umma.cu"  // dense attention, head dim 128 (the VAE bottleneck)

static int attention_p32_impl(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                              int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                              float scale, float* lse, int* lse_written, void* stream) {
  using namespace sdb;
  if (lse_written) *lse_written = 0;
  SDB_CHECK_ARG(q && k && v && out && B > 0 && heads > 0 && n > 0 && m > 0, "attention_p32: bad args");
  SDB_CHECK_ARG(d > 0 && d <= 20 && d % 4 == 0, "attention_p32: head dim (%d) must be a multiple of 4, <= 20", d);
  SDB_CHECK_ARG(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 4 == 0,
                "attention_p32: ldq/ldk/ldv must be multiples of 8 (16-byte TMA rows), ldo of 4");
  SDB_CHECK_ARG(ldq >= heads * 32 && ldk >= heads * 32 && ldv >= heads * 32,
                "attention_p32: row strides must cover heads*32 padded columns");
  SDB_CHECK_ARG(((uintptr_t)q & 15) == 0 && ((uintptr_t)k & 15) == 0 && ((uintptr_t)v & 15) == 0 &&
                    ((uintptr_t)out & 7) == 0,
                "attention_p32: q/k/v must be 16-byte aligned, out 8-byte aligned");
  SDB_CHECK_ARG(heads <= 65535 && B <= 65535, "attention_p32: grid too large");
  // a context of <= 128 tokens (cross-attention over CLIP's 77) is one key tile: the streaming mma.sync
  // kernel beats this pipeline's per-CTA set-up by an order of magnitude there
  static int no_ctx = -1;
  if (no_ctx < 0) { const char* e = getenv("SDB_FA_NO_CTX"); no_ctx = e ? atoi(e) : 0; }
  // <= 96 keys: the tcgen05 short-context kernel (K / V resident per CTA, query tiles streamed); SDB_FA_CTX_UMMA=0
  // falls back to the mma.sync one, which also serves 97..128 keys
  static int ctx_umma = -1, ctx_tiles = 8;
  if (ctx_umma < 0) {
    const char* e = getenv("SDB_FA_CTX_UMMA");
    ctx_umma = e ? atoi(e) : 1;
    const char* t = getenv("SDB_FA_CTX_TILES");
    if (t && atoi(t) > 0) ctx_tiles = atoi(t);
  }
  if (m <= kCuMaxNS && ctx_umma && !no_ctx) {      // (no short-context kernel keeps the lse: *lse_written stays 0)
    CuParams c;
    memset(&c, 0, sizeof(c));
    const int NS = (m + 15) / 16 * 16;
    const uint32_t qbox[3] = {32, 128, 1};
    const uint32_t kvbox[3] = {32, (uint32_t)NS, 1};
    uint64_t dims[3] = {(uint64_t)heads * 32, (uint64_t)n, (uint64_t)B};
    uint64_t st[3] = {0, (uint64_t)ldq * 2, (uint64_t)n * ldq * 2};
    int rc = sdb_make_tensor_map(&c.tmQ, q, 2, 3, dims, st, qbox, 64);
    if (rc) return rc;
    dims[1] = (uint64_t)m;
    st[1] = (uint64_t)ldk * 2; st[2] = (uint64_t)m * ldk * 2;
    rc = sdb_make_tensor_map(&c.tmK, k, 2, 3, dims, st, kvbox, 64);
    if (rc) return rc;
    st[1] = (uint64_t)ldv * 2; st[2] = (uint64_t)m * ldv * 2;
    rc = sdb_make_tensor_map(&c.tmV, v, 2, 3, dims, st, kvbox, 64);
    if (rc) return rc;
    c.out = (__nv_bfloat16*)out;
    c.ldo = ldo;
    c.n = n; c.m = m; c.d = d; c.heads = heads;
    c.scale_log2 = scale * 1.4426950408889634f;
    // query tiles per CTA: up to SDB_FA_CTX_TILES (8), at least 2 (one per softmax warpgroup), fewer when the grid would
    // not fill the 2 x 148 CTA slots; the chunks of one (batch, head) are kept the same length
    const int n_tiles = (n + kCuBM - 1) / kCuBM;
    long long want = (long long)heads * B * n_tiles / 296;
    want = want < 2 ? 2 : (want > ctx_tiles ? ctx_tiles : want);
    const int chunks = (n_tiles + (int)want - 1) / (int)want;
    c.tiles_per_cta = (n_tiles + chunks - 1) / chunks;
    dim3 grid(heads, (n_tiles + c.tiles_per_cta - 1) / c.tiles_per_cta, B);
    cudaStream_t s0 = sdb_stream(stream);
    switch (NS) {
      case 16: rc = launch_attention_ctx_umma<16>(c, grid, s0); break;
      case 32: rc = launch_attention_ctx_umma<32>(c, grid, s0); break;
      case 48: rc = launch_attention_ctx_umma<48>(c, grid, s0); break;
      case 64: rc = launch_attention_ctx_umma<64>(c, grid, s0); break;
      case 80: rc = launch_attention_ctx_umma<80>(c, grid, s0); break;
      default: rc = launch_attention_ctx_umma<96>(c, grid, s0); break;
    }
    if (rc) return rc;
    SDB_LAUNCH_OK();
    return SDB_OK;
  }
  if (m <= 128 && !no_ctx)
    return sdb_attention_ctx_dispatch(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, sdb_stream(stream));
  FaParams p;
  memset(&p, 0, sizeof(p));
  // kernel shape: one warpgroup per CTA, 64-key tiles, 3 or 4 CTAs per SM (SDB_FA_WG1; default 4).  The earlier
  // two-warpgroup shape (128-key tiles, 1.20 vs 1.085 ms per layer) is retired: tools/experiments/.
  static int wg1 = -1;
  if (wg1 < 0) {
    const char* e = getenv("SDB_FA_WG1");
    wg1 = e ? atoi(e) : 4;
    if (wg1 != 3 && wg1 != 4) wg1 = 4;
  }
  const uint32_t box[3] = {32, 128, 1};
  const uint32_t box_kv[3] = {32, 64u, 1};
  {
    uint64_t dims[3] = {(uint64_t)heads * 32, (uint64_t)n, (uint64_t)B};
    uint64_t st[3] = {0, (uint64_t)ldq * 2, (uint64_t)n * ldq * 2};
    int rc = sdb_make_tensor_map(&p.tmQ, q, 2, 3, dims, st, box, 64);
    if (rc) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)heads * 32, (uint64_t)m, (uint64_t)B};
    uint64_t st[3] = {0, (uint64_t)ldk * 2, (uint64_t)m * ldk * 2};
    int rc = sdb_make_tensor_map(&p.tmK, k, 2, 3, dims, st, box_kv, 64);
    if (rc) return rc;
    st[1] = (uint64_t)ldv * 2;
    st[2] = (uint64_t)m * ldv * 2;
    rc = sdb_make_tensor_map(&p.tmV, v, 2, 3, dims, st, box_kv, 64);
    if (rc) return rc;
  }
  p.out = (__nv_bfloat16*)out;
  p.ldo = ldo;
  p.n = n; p.m = m; p.d = d; p.heads = heads;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.lse = lse;
  if (lse && lse_written) *lse_written = 1;
  // share of exponentials evaluated on the FMA pipe: kPoly pairs out of 8 (tuning aid: SDB_FA_POLY)
  static int poly = -1;
  if (poly < 0) {
    const char* e = getenv("SDB_FA_POLY");
    poly = e ? atoi(e) : 2;   // measured best
    if (poly < 0 || poly > 6) poly = 2;
  }
  if (wg1) {
    dim3 grid1((n + kF1BM - 1) / kF1BM, heads, B);
    int rc;
    static int spread = -1;
    if (spread < 0) { const char* e = getenv("SDB_FA_SPREAD"); spread = e ? atoi(e) : 1; }
    // instantiated shapes: the default and its A/B neighbours (tools/experiments/README.md has the full sweep)
    if (wg1 == 4 && spread) rc = poly == 3 ? launch_attention_wg1<3, 4, true>(p, grid1, sdb_stream(stream))
                                           : launch_attention_wg1<2, 4, true>(p, grid1, sdb_stream(stream));
    else if (wg1 == 4) rc = launch_attention_wg1<2, 4>(p, grid1, sdb_stream(stream));
    else rc = launch_attention_wg1<2, 3>(p, grid1, sdb_stream(stream));
    if (rc) return rc;
    SDB_LAUNCH_OK();
    return SDB_OK;
  }
  sdb_set_error("attention_p32: no kernel shape selected");
  return SDB_EUNSUPPORTED;
}

extern "C" int sdb_attention_p32(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                                 int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                                 float scale, void* stream) {
  return attention_p32_impl(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, nullptr, nullptr, stream);
}

extern "C" int sdb_attention_p32_lse(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                                     int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                                     float scale, float* lse, int* lse_written, void* stream) {
  SDB_CHECK_ARG(lse && lse_written, "attention_p32_lse: null lse buffer");
  return attention_p32_impl(q, ldq, k, ldk, v, ldv, out, ldo, B, heads, n, m, d, scale, lse, lse_written, stream);
}
