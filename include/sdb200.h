/* libsdb200 -- C ABI of the B200-native latent-diffusion hot path.
 *
 * The reference (lwb2099/stable_diffusion_pytorch) is pure Python: its hot path has no FFI; every
 * GPU instruction it runs is reached through torch.nn / ATen.  This header is the
 * boundary a maintainer binds instead of those ATen calls; each entry point names the reference
 * lines whose arithmetic it replaces (paths relative to the reference repo root).  INTEGRATION.md
 * shows the ctypes stubs.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes; no torch types.  All pointers are DEVICE pointers
 *    unless a parameter says "host".  The caller owns every buffer; the library allocates nothing
 *    on the hot path and never synchronises: all work is enqueued on `stream` (a cudaStream_t).
 *  - return 0 on success, a negative SDB_E* code otherwise; sdb_last_error() returns a
 *    thread-local message.  Never throws, never exits.
 *  - activations are channels-last: images NHWC [B, H, W, C], tokens [B, n, C] (identical memory),
 *    so `b c h w -> b (h w) c` (modules/transformer.py:378-381) is free.
 *  - dtype codes: SDB_F32 / SDB_BF16.  bf16 tensors feed the tcgen05 tensor-core kernels; accumu-
 *    lation, bias, statistics and the sampler arithmetic are fp32.
 *  - there is NO CPU fallback: every entry point needs an sm_100 device.
 */
#ifndef SDB200_H_
#define SDB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SDB_OK 0
#define SDB_EINVAL (-1)  /* bad argument (mirrors the reference's asserts)          */
#define SDB_ECUDA (-2)   /* a CUDA runtime / driver call failed                      */
#define SDB_EUNSUPPORTED (-3) /* shape outside what the sm_100a kernels are built for */

#define SDB_F32 0
#define SDB_BF16 1
#define SDB_U8 2   /* output only: sdb_conv3x3 image head, see SdbConvArgs.out_channels */

#define SDB_ACT_NONE 0
#define SDB_ACT_SILU 1  /* x*sigmoid(x) on the result                                          */
#define SDB_ACT_GEGLU 2 /* weight rows interleaved in pairs (a0,a1,g0,g1,a2,a3,g2,g3,..): out[:, j] = a_j*gelu_erf(g_j) */

int sdb_version(void);
const char* sdb_last_error(void);
/* 1 if any kernel hit its bounded-wait timeout since the last call (debug aid); resets the flag.
 * Synchronises the device. */
int sdb_debug_timeout_flag(void);
/* number of kernel launches enqueued by this library since process start (bench: gpu_launches) */
int64_t sdb_launch_count(void);

/* ---- K6: timestep embedding ------------------------------------------------------------------
 * modules/timestep_embedding.py:73-105 (`sinusoidal_time_proj`): out[b] = cat[sin(t_b*freq),
 * cos(t_b*freq)].  `freq` is the [half] fp32 table the caller builds on the HOST with the
 * reference's own expression (positive exponent) and uploads once.  Full-range sinf/cosf. */
int sdb_timestep_embed(const int64_t* t, const float* freq, void* out, int B, int half,
                       int out_dtype, void* stream);

/* ---- K7: sampler ----------------------------------------------------------------------------
 * Per-step coefficients live in a device table coef[step_slot][8] (fp32):
 *   [0] sqrt_recip_alpha_bar  [1] sqrt_recip_m1_alpha_bar  [2] mean_x0_coef  [3] mean_xt_coef
 *   [4] sigma = exp(0.5*log_var) (0 when the step adds no noise)
 *   [5] ddim sqrt(abar_prev)  [6] ddim sqrt(1-abar_prev)  [7] unused
 * all built on the host from models/scheduler.py:52-87 with the reference's torch ops.
 * `slot_dev` (device int32*, may be NULL) selects the row at run time so one captured CUDA graph
 * serves every step; when NULL, `slot` is used.
 *
 * sdb_cfg_sampler_step fuses models/latent_diffusion.py:62-65 (guidance combine, halves in the
 * reference's [uncond, cond] order: eps = e0 + g*(e0 - e1); cfg_mode 0 = no guidance, eps = e0
 * only) with models/scheduler.py:173-213 (DDPM: x0 = c0*x - c1*eps; mean = c2*x0 + c3*x;
 * x_prev = mean + sigma*(noise*scale_factor)) or the eta=0 DDIM update (kind 1: x_prev =
 * c5*x0 + c6*eps).  fp32 mode performs each multiply/add with a separate IEEE rounding in the
 * reference's order, so it is bit-exact against the CPU reference.
 *   eps   : [2*B or B, n_per] network output (NCHW flattened), dtype `dtype`
 *   noise : same dtype, [B, n_per] pre-drawn host noise ([n_slots, B, n_per], indexed by the device
 *           slot, when slot_dev is given), or NULL -> in-kernel Philox N(0,1) keyed by
 *           (seed, slot) (statistical parity only); ignored when sigma == 0
 *   x_prev, x0 (x0 may be NULL): [B, n_per] */
int sdb_cfg_sampler_step(const void* eps, const void* x_t, const void* noise, const float* coef,
                         const int32_t* slot_dev, int slot, int kind, int cfg_mode, float guidance,
                         float scale_factor, uint64_t seed, void* x_prev, void* x0, int B,
                         int64_t n_per, int dtype, void* stream);
/* guidance combine alone: out[i] = e0[i] + g*(e0[i] - e1[i]), eps = [e0 | e1], `total` = |e0| */
int sdb_cfg_combine(const void* eps, float guidance, void* out, int64_t total, int dtype, void* stream);
/* advances the device step slot and refreshes the int64 timestep vector the UNet reads:
 * t_out[0..nt) = t_table[slot+1] (models/latent_diffusion.py:110). */
int sdb_sampler_advance(int32_t* slot_dev, const int64_t* t_table, int n_slots, int64_t* t_out,
                        int nt, void* stream);
/* models/scheduler.py:95-139 add_noise: out = sqrt(ac[t_b])*x0 + sqrt(1-ac[t_b])*noise with the
 * table cast to `dtype` first and pow(0.5) taken in that dtype, as the reference does. */
int sdb_add_noise(const void* x0, const void* noise, const int64_t* t, const float* alphas_cumprod,
                  int n_steps, void* out, int B, int64_t n_per, int dtype, void* stream);

/* ---- layout helpers --------------------------------------------------------------------------
 * NCHW (reference layout, any dtype) <-> NHWC bf16/fp32 used inside the path. */
int sdb_nchw_to_nhwc(const void* src, int src_dtype, void* dst, int dst_dtype, int B, int C, int H,
                     int W, void* stream);
int sdb_nhwc_to_nchw(const void* src, int src_dtype, void* dst, int dst_dtype, int B, int C, int H,
                     int W, void* stream);

/* ---- K2: GroupNorm ---------------------------------------------------------------------------
 * torch.nn.GroupNorm (modules/resnet2d.py:172,188; transformer.py:329; models/utils.py:195),
 * eps 1e-5, biased variance.  Two launches: statistics, then normalise(+SiLU).
 * The input may be a VIRTUAL channel concat of two NHWC tensors (x1: C1 channels, x2: C2
 * channels, x2 may be NULL) -- replaces torch.cat in models/unet.py:308.
 *   stats : fp32 [B, groups, 2] (mean, rstd), caller-allocated
 *   scratch: fp32 [B, chunks, C, 2] partial sums, chunks = sdb_groupnorm_chunks(H*W)          */
int sdb_groupnorm_chunks(int64_t hw);
int sdb_groupnorm_stats(const void* x1, int C1, const void* x2, int C2, int dtype, int B,
                        int64_t hw, int groups, float eps, float* scratch, float* stats,
                        void* stream);
/* sdb_groupnorm_stats without a caller scratch buffer.  No counterpart in the reference: added
 * for callers that cannot allocate (CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  The partial sums go to a file-scope `static __device__`
 * array of SDB_GN_STATIC_SCRATCH_FLOATS floats (4 MiB, held by every GPU that loads the library until
 * exit); the call allocates nothing.  Limit: B * sdb_groupnorm_chunks(hw) * (C1+C2) * 2 <=
 * SDB_GN_STATIC_SCRATCH_FLOATS, SDB_EINVAL beyond it (use sdb_groupnorm_stats).  All calls on one GPU
 * share the array: calls that overlap on the GPU (different streams, no ordering between them) race;
 * calls on one stream are safe.  Nothing needs clearing between calls (every element read is written
 * first).  Same statistics, bit for bit, as sdb_groupnorm_stats. */
#define SDB_GN_STATIC_SCRATCH_FLOATS (1u << 20)
int sdb_groupnorm_stats_static(const void* x1, int C1, const void* x2, int C2, int dtype, int B,
                               int64_t hw, int groups, float eps, float* stats, void* stream);
/* GroupNorm(+SiLU) in ONE pass (same reference lines): a thread-block cluster per image keeps the image in shared
 * memory, sums it as it lands, exchanges the per-group sums over distributed shared memory and normalises out of
 * shared memory -- one HBM read and one write instead of two reads and one write in three launches.  bf16 in, bf16
 * dense NHWC out (C1+C2 channels).  SDB_EUNSUPPORTED when the image does not fit the shared memory of a cluster of
 * <= 16 CTAs (about 3 MB of bf16 per image): use sdb_groupnorm_stats + sdb_groupnorm_apply then. */
int sdb_groupnorm_fused(const void* x1, int C1, const void* x2, int C2, const float* gamma, const float* beta,
                        int act, void* out, int B, int64_t hw, int groups, float eps, void* stream);
/* GroupNorm(+SiLU) in ONE launch for any image size (same reference lines): a persistent kernel streams the tensor
 * for the statistics, crosses a device-wide barrier (cooperative launch) and streams it again -- out of L2 when the
 * tensor fits there -- to write the normalised bf16 output.  `partial`: fp32 scratch of
 * sdb_groupnorm_coop_scratch(B, hw, groups) floats; `bar`: two uint32 words, zero before the first call, reusable
 * by later calls that do not overlap this one on the GPU. */
int64_t sdb_groupnorm_coop_scratch(int B, int64_t hw, int groups);
int sdb_groupnorm_coop(const void* x1, int C1, const void* x2, int C2, const float* gamma, const float* beta, int act,
                       void* out, int B, int64_t hw, int groups, float eps, float* partial, uint32_t* bar,
                       void* stream);
/* out[b,p,c] = act((x-mean)*rstd*gamma[c]+beta[c]); act = SDB_ACT_NONE | SDB_ACT_SILU.
 * out is a dense NHWC tensor with C1+C2 channels, dtype out_dtype. */
int sdb_groupnorm_apply(const void* x1, int C1, const void* x2, int C2, int dtype,
                        const float* stats, const float* gamma, const float* beta, int act,
                        void* out, int out_dtype, int B, int64_t hw, int groups, void* stream);
/* Normalise (+SiLU) with the statistics taken from per-channel (sum, sum of squares) pairs that the producers of x1
 * (and x2) accumulated in their epilogues (SdbGemmArgs.chan_sums / SdbConvArgs.chan_sums: int64 fixed point [B][C1][2]
 * and [B][C2][2]) -- the statistics pass over the tensor disappears.  One read + one write of the tensor, one launch.
 * groups <= 128. */
int sdb_groupnorm_apply_sums(const void* x1, int C1, const void* sums1, const void* x2, int C2,
                             const void* sums2, int dtype, const float* gamma, const float* beta,
                             float eps, int act, void* out, int out_dtype, int B, int64_t hw, int groups,
                             void* stream);

/* ---- K5: residual add + LayerNorm ------------------------------------------------------------
 * modules/transformer.py:274-278 post-norm: out = LN(x (+ res)) over the last dim C, eps 1e-5. */
int sdb_add_layernorm(const void* x, const void* res, const float* gamma, const float* beta,
                      void* out, int64_t rows, int C, float eps, int dtype, void* stream);

/* ---- K3: GEMM with fused epilogue (tcgen05 + TMEM + TMA) ------------------------------------
 * out[M, N] = act( A[M, K] . W[N, K]^T + bias[N] ) + residual[M, N]
 * nn.Linear / 1x1 Conv2d call sites: transformer.py:55-58,151,176,330,347; resnet2d.py:182,202;
 * unet.py:185-187.  A may be two K segments (a: K1 columns, a2: K2 columns) -- the virtual concat
 * for the 1x1 skip convolution.  bf16 operands, fp32 accumulate. */
typedef struct SdbGemmArgs {
  const void* a;   int64_t lda;  int K1;       /* bf16 [M, K1], row stride lda elements        */
  const void* a2;  int64_t lda2; int K2;       /* optional second segment, NULL / 0            */
  const void* w;                               /* bf16 [N, K1+K2] row-major                    */
  const float* bias;                           /* fp32 [N] or NULL                             */
  const void* residual; int64_t ldr; int res_dtype; /* [M, Nout] or NULL                      */
  void* out; int64_t ldo; int out_dtype;       /* [M, Nout]; Nout = N (N/2 for GEGLU)          */
  int M; int N;
  int act;                                     /* SDB_ACT_*                                    */
  /* optional LayerNorm of every finished output row in the epilogue: out = LN(A.W^T + bias + residual) * gamma + beta
   * -- the post-norm step of modules/transformer.py:274-278 without a second pass over HBM.  NULL = off.  Needs bf16
   * out + bf16 residual, act NONE and N <= 256 in one tile (SDB_EUNSUPPORTED otherwise: run sdb_add_layernorm). */
  const float* ln_gamma; const float* ln_beta; float ln_eps;
  /* optional per-channel sums of the finished output for the GroupNorm that reads it next (resnet2d.py:171,187;
   * transformer.py:332): chan_sums, int64 [M / rows_per_image][N][2], receives += (sum, sum of squares) over the rows
   * of each image of the bf16-rounded output, in fixed point (sum: signed, 2^-30 steps; squares: unsigned, 2^-24
   * steps, i.e. < 1.1e12 per image and channel) by 64-bit integer atomics -- exact, so the totals do not depend on
   * the order the CTAs arrive in.  The caller zeroes it; sdb_groupnorm_apply_sums consumes it.  NULL = off.  Needs
   * bf16 out, act NONE, no LayerNorm, rows_per_image %% 128 == 0 dividing M (SDB_EUNSUPPORTED otherwise: nothing was
   * launched, call again without it). */
  void* chan_sums; int64_t rows_per_image;
} SdbGemmArgs;
int sdb_gemm(const SdbGemmArgs* args, void* stream);

/* ---- K1: 3x3 convolution as implicit GEMM (tcgen05 + TMEM + TMA im2col-by-shifted-box) ------
 * nn.Conv2d(k=3, padding=1) call sites: resnet2d.py:41,103,174,195; models/utils.py:32,197.
 * x NHWC bf16 [B,H,W,Cin] (already normalised/activated: padding is applied AFTER GN+SiLU as in
 * resnet2d.py:171-177, i.e. out-of-image taps contribute exact zeros via TMA OOB fill).
 * w bf16 [Cout][3][3][Cin].  Epilogue: + bias[Cout] + temb[b, Cout] (resnet2d.py:235) +
 * residual NHWC (resnet2d.py:238).
 * down2 = 1: conv followed by nearest x0.5 (resnet2d.py:122-123) == keep even rows/cols.
 * up2 = 1:   nearest x2 THEN conv (resnet2d.py:61-62); x is the low-res tensor [B,H,W,Cin],
 *            out is [B,2H,2W,Cout] (residual, if any, likewise).  Runs as four 2x2-tap "phase" convolutions
 *            (4/9 of the flops, no 4x tensor): w must then be the PHASE-COLLAPSED weights bf16
 *            [4][Cout][2][2][Cin], phase 2a+b = output pixels (2y+a, 2x+b), taps over source rows
 *            {y-1,y} (a = 0) / {y,y+1} (a = 1) and the same for columns, each the fp32 sum of the 3x3
 *            taps that fall on that source pixel.  sdb_conv3x3_f32 takes the plain 3x3 weights. */
typedef struct SdbConvArgs {
  const void* x; const void* w; const float* bias;
  const float* temb; int64_t ld_temb;          /* fp32 [B, ld_temb] or NULL                    */
  const void* residual; int res_dtype;         /* NHWC [B,Ho,Wo,Cout] or NULL                  */
  void* out; int out_dtype;
  int B, H, W, Cin, Cout;
  int down2; int up2;
  /* out_dtype = SDB_U8 (the decoder's image head, models/utils.py:197 followed by the reference's `detransform`,
   * utils/prepare_dataset.py:145-150, in the conv epilogue): out is uint8 NHWC [B,H,W,out_channels] holding
   * trunc(clip((y + 1) / 2, 0, 1) * 255) of the first out_channels <= 16 = Cout output channels (the head's Cout is
   * padded to 16 for the tensor cores).  No residual, no down2 / up2.  Ignored for other out dtypes. */
  int out_channels;
  /* optional per-channel sums of the finished output, as SdbGemmArgs.chan_sums: int64 [B][Cout][2], caller-zeroed.
   * Needs bf16 out, no down2, H x W tiling into whole 128-pixel boxes (SDB_EUNSUPPORTED otherwise, nothing launched).
   * With up2 the four phase launches add into the same buffer. */
  void* chan_sums;
} SdbConvArgs;
int sdb_conv3x3(const SdbConvArgs* args, void* stream);

/* fp32 parity mode of K3 / K1 (north_star: one UNet forward within 1e-3 relative L2 in fp32; the reference
 * run with `--dtype float32`, scripts/txt2img.py:23): same argument structs, but a / a2 / w (x / w) are fp32 and
 * the products are fp32 FMAs on the CUDA cores; residual / out dtypes as given.  No alignment or shape limits
 * beyond GEGLU's N % 4 == 0; up2 is honoured.  sdb_attention(dtype = SDB_F32) is the matching attention. */
int sdb_gemm_f32(const SdbGemmArgs* args, void* stream);
int sdb_conv3x3_f32(const SdbConvArgs* args, void* stream);

/* small-channel 3x3 convs on CUDA cores (HBM-bound): conv_in (models/utils.py:31-32, Cin = latent
 * channels; reads NCHW `src_dtype`, writes NHWC; the input holds src_B images and output image b reads
 * input image b %% src_B, which is the `torch.cat([x] * 2)` of latent_diffusion.py:53 for free) and the output head conv (models/utils.py:197,
 * Cout = latent/img channels; reads NHWC, writes NCHW).  fp32 weights OIHW as in the state_dict. */
int sdb_conv3x3_in(const void* x_nchw, int src_dtype, const float* w, const float* bias, void* out_nhwc,
                   int out_dtype, int B, int src_B, int H, int W, int Cin, int Cout, void* stream);
int sdb_conv3x3_out(const void* x_nhwc, int src_dtype, const float* w, const float* bias,
                    void* out_nchw, int out_dtype, int B, int H, int W, int Cin, int Cout,
                    void* stream);
/* 1x1 conv on a thin NCHW tensor (post_quant_conv, models/autoencoder.py:110,160); C <= 8 */
int sdb_conv1x1_nchw_small(const void* x, const float* w, const float* bias, void* out, int dtype,
                           int B, int H, int W, int Cin, int Cout, void* stream);
/* nearest x2 upsample NHWC (resnet2d.py:61), used ahead of sdb_conv3x3 when up2 is not fused */
/* fp32 channels-last [B,H,W,Cp] -> NCHW [B,C,H,W] keeping the first C <= Cp channels (output head: the
 * tensor-core conv pads Cout to 16; reference layer: models/utils.py:197). */
int sdb_nhwc_slice_to_nchw(const float* src, int Cp, void* dst, int dst_dtype, int B, int C, int H, int W,
                           void* stream);
int sdb_upsample2x(const void* x, void* out, int dtype, int B, int H, int W, int C, void* stream);

/* ---- K4: fused attention --------------------------------------------------------------------
 * modules/transformer.py:93-114: softmax(Q K^T * scale) V per (batch, head) without
 * materialising `sim`.  q [B, n, *] with head h at columns [h*d, (h+1)*d) and row stride ldq
 * elements (so a fused QKV GEMM output can be passed as three offset pointers); k, v [B, m, *].
 * out [B, n, heads*d] row stride ldo.  bf16 in/out, fp32 softmax state.  d <= 128, d % 4 == 0
 * (SDB_EUNSUPPORTED beyond: see sdb_softmax_rows for the wide-head route).
 * dtype = SDB_F32: fp32 in/out on the CUDA cores (parity mode), any d <= 128, no alignment rules. */
int sdb_attention(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                  int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                  float scale, int dtype, void* stream);

/* Row softmax of a materialised fp32 score matrix: out[r, c] = softmax_c(scale * x[r, :valid_cols])[c], exact zeros
 * for valid_cols <= c < cols (modules/transformer.py:101-111).  The wide-head route (d > 128: the VAE encoder
 * bottleneck, d_head = mid_ch, models/autoencoder.py:181-183) runs attention as sdb_gemm(Q, K) -> sdb_softmax_rows ->
 * sdb_gemm(P, V^T); out bf16 or fp32, row strides in elements. */
int sdb_softmax_rows(const float* x, int64_t ldx, void* out, int64_t ldo, int out_dtype, int64_t rows, int cols,
                     int valid_cols, float scale, void* stream);

/* Tensor-core (tcgen05 + TMEM + TMA) variant for small head dims, d <= 20 (the UNet's d_head = 20).
 * Operands are HEAD-PADDED to 32 columns by the projection GEMM (weights packed once at load):
 * head h of q / k / v occupies columns [32h, 32h+32) of rows with stride ldq / ldk / ldv;
 * q and k are zero in columns d..31, v holds 1.0 in column d (so P.V also yields the softmax
 * denominator) and anything finite beyond.  out stays dense: [B, n, heads*d], row stride ldo.
 * bf16 in/out, fp32 softmax state; any n, m (ragged tails are masked). */
int sdb_attention_p32(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                      int64_t ldv, void* out, int64_t ldo, int B, int heads, int n, int m, int d,
                      float scale, void* stream);

/* ---- elementwise helpers --------------------------------------------------------------------*/
int sdb_silu(const void* x, void* out, int64_t n, int dtype, int out_dtype, void* stream);
int sdb_cast(const void* x, int src_dtype, void* out, int dst_dtype, int64_t n, void* stream);
/* VAE head: img = clip((x+1)/2, 0, 1)*255 -> uint8 (utils/prepare_dataset.py:145-150), NCHW in,
 * NHWC uint8 out. */
int sdb_to_uint8_image(const void* x_nchw, int dtype, uint8_t* out_nhwc, int B, int C, int H, int W,
                       void* stream);

/* ---- train step T1 (train_unet.py:489-529): backward kernels ------------------------------------
 * Gradients of every GEMM / convolution re-use sdb_gemm / sdb_conv3x3 (dgrad with transposed weights,
 * wgrad over transposed operands; convolutions through sdb_im2col3x3).  `*_acc` / `*_zeroed` outputs are
 * accumulated with atomics into caller-zeroed fp32 buffers. */
/* out[g, c] += sum over the rows_per_group rows of group g of x[row, c]  (bias / time-embedding grads) */
int sdb_colsum(const void* x, int dtype, int64_t ld, int64_t groups, int64_t rows_per_group, int cols,
               float* out_zeroed, void* stream);
/* GEGLU on pair-interleaved columns (a0,a1,g0,g1,..) kept unfused in training: transformer.py:179-192 */
int sdb_geglu_fwd(const void* h, void* out, int64_t M, int F, void* stream);
int sdb_geglu_bwd(const void* h, const void* dout, void* dh, int64_t M, int F, void* stream);
int sdb_silu_bwd(const void* x, int x_dtype, const void* dy, void* dx, int g_dtype, int64_t n, void* stream);
/* nn.GroupNorm(+SiLU) backward over the virtual concat [x1, x2]; stats = [B, groups, (mean, rstd)] from
 * sdb_groupnorm_stats; sums_scratch = [B, groups, 2] floats */
int sdb_groupnorm_bwd(const void* x1, int C1, const void* x2, int C2, int dtype, const float* stats,
                      const float* gamma, const float* beta, int act, const void* dy, void* dx1, void* dx2,
                      float* dgamma_acc, float* dbeta_acc, float* sums_scratch, int B, int64_t hw, int groups,
                      void* stream);
/* sdb_groupnorm_bwd on 16-byte vectors (bf16, C1 and C2 multiples of 8): same arithmetic through per-(image, channel)
 * column sums; colbuf_zeroed = [B][2][C1 + C2] floats.  Other shapes run sdb_groupnorm_bwd. */
int sdb_groupnorm_bwd_vec(const void* x1, int C1, const void* x2, int C2, int dtype, const float* stats,
                          const float* gamma, const float* beta, int act, const void* dy, void* dx1, void* dx2,
                          float* dgamma_acc, float* dbeta_acc, float* sums_scratch, float* colbuf_zeroed, int B,
                          int64_t hw, int groups, void* stream);
int sdb_layernorm_bwd(const void* x, const float* gamma, const void* dy, void* dx, float* dgamma_acc,
                      float* dbeta_acc, int64_t rows, int C, float eps, int dtype, void* stream);
/* softmax(QK^T scale)V backward (autograd of transformer.py:93-114), any n / m, d <= 32 (SDB_EUNSUPPORTED beyond):
 * flash-style recomputation, no [n, m] tensor.  head_pitch = columns per head in q/k/v and their gradients (32 for
 * the head-padded operands of sdb_attention_p32, d for dense ones); o / dO dense [B, n, heads*d];
 * lse_scratch / delta_scratch: fp32 [B * heads * n] each.  lse_valid != 0: lse_scratch already holds the row
 * log-sum-exp (log2 domain) written by sdb_attention_p32_lse, and the recomputation pass is skipped. */
int sdb_attention_bwd(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                      const void* o, int64_t ldo, const void* dO, int64_t lddo, void* dq, int64_t lddq, void* dk,
                      int64_t lddk, void* dv, int64_t lddv, float* lse_scratch, float* delta_scratch, int lse_valid,
                      int B, int heads, int n, int m, int d, int head_pitch, float scale, void* stream);
/* sdb_attention_p32 that also keeps lse[b, h, row] = log2(sum_j exp2(s_ij * scale * log2 e)) for the backward.
 * *lse_written = 1 if the buffer was filled (the short-context kernel for m <= 128 does not keep it). */
int sdb_attention_p32_lse(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                          void* out, int64_t ldo, int B, int heads, int n, int m, int d, float scale, float* lse,
                          int* lse_written, void* stream);
/* The two halves of sdb_attention_bwd for the head-padded (tcgen05) layout when the forward kept lse:
 *   sdb_attention_bwd_dq_p32: dQ on tcgen05 / TMEM (S and dP by UMMA, dS written back to TMEM as the A operand of
 *     dS.K); needs dO re-laid head-padded (dO_pad, zero pad columns) and delta[b, h, row] = sum_c dO * O;
 *   sdb_attention_bwd_dkv: the dK / dV kernel alone (dense dO).  lse, delta: fp32 [B, heads, n]. */
int sdb_attention_bwd_dq_p32(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                             const void* dO_pad, int64_t lddo, const float* lse, const float* delta, void* dq,
                             int64_t lddq, int B, int heads, int n, int m, int d, float scale, void* stream);
/* dQ of the head-padded attention over a short context (m <= 128; autograd of transformer.py:93-114 for the CLIP
 * context), one pass: softmax statistics in registers, also written out (lse log2-domain, delta; [B,heads,n] fp32) for
 * sdb_attention_bwd_dkv.  dO dense [B,n,heads*d]; dq head-padded, every column written.  SDB_EUNSUPPORTED for m > 128. */
int sdb_attention_ctx_bwd_dq(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                             const void* dO, int64_t lddo, void* dq, int64_t lddq, float* lse, float* delta, int B,
                             int heads, int n, int m, int d, float scale, void* stream);
/* One pass ahead of the tcgen05 backward kernels: do_pad [B,n,heads,32] bf16 (pad 0), delta = rowsum(dO o O), and the
 * negated delta / lse ([B,heads,n] fp32) that sdb_attention_bwd_dkv_p32 takes. */
int sdb_attention_bwd_prep(const void* dO, int64_t lddo, const void* O, int64_t ldo, const float* lse, void* do_pad,
                           float* delta, float* neg_delta, float* neg_lse, int B, int heads, int n, int d, void* stream);
/* dK / dV on tcgen05 (transposed problem: thread = key row, lse / delta per query column through TMA boxes); dk / dv
 * head-padded; lse / delta are passed NEGATED (sdb_attention_bwd_prep writes them); needs n % 4 == 0,
 * SDB_EUNSUPPORTED otherwise. */
int sdb_attention_bwd_dkv_p32(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                              const void* dO_pad, int64_t lddo, const float* neg_lse, const float* neg_delta, void* dk,
                              int64_t lddk, void* dv, int64_t lddv, int B, int heads, int n, int m, int d, float scale,
                              void* stream);
int sdb_attention_bwd_dkv(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                          const void* dO, int64_t lddo, void* dk, int64_t lddk, void* dv, int64_t lddv, const float* lse,
                          const float* delta, int B, int heads, int n, int m, int d, int head_pitch, float scale,
                          void* stream);
int sdb_upsample2x_bwd(const void* dy, void* dx, int dtype, int B, int H, int W, int C, void* stream);
/* DownSample backward: dy [B,H/2,W/2,C] lands on the even pixels of out [B,H,W,C], zeros elsewhere */
int sdb_scatter_even(const void* dy, void* out, int dtype, int B, int H, int W, int C, void* stream);
/* Weight gradients on the tensor cores (tcgen05, MN-major operands read as they lie in HBM; split over the pixel
 * dimension, fp32 atomics into `dw`, which the caller zeroes).  autograd of nn.Conv2d 3x3 (resnet2d.py:41,103,174,195)
 * and nn.Linear / 1x1 conv.  dy, x bf16.
 *   conv:   dw[n][ky][kx][c] += sum_{b,h,w} dy[b,h,w,n] * x[b,h+ky-1,w+kx-1,c]   (zero padding), NHWC operands
 *   linear: dw[n*lddw + c]   += sum_r dy[r*lddy + n] * x[r*ldx + c]
 * SDB_EUNSUPPORTED for shapes outside the kernel (W must divide 64, Cin / K a multiple of 64): the caller then
 * uses the transpose + GEMM route. */
int sdb_wgrad_conv3x3(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                      void* stream);
int sdb_wgrad_linear(const void* dy, int64_t lddy, const void* x, int64_t ldx, float* dw, int64_t lddw, int M,
                     int N, int K, void* stream);

/* [B,H,W,C] bf16 -> [B*H*W, ldc] bf16, columns (tap, c), zero padding (3x3, pad 1) and zero tail columns */
int sdb_im2col3x3(const void* x, void* out, int B, int H, int W, int C, int ldc, void* stream);
int sdb_nchw_to_nhwc_pad(const void* src, int src_dtype, void* dst, int dst_dtype, int B, int C, int Cp, int H,
                         int W, void* stream);
/* loss = mean((pred - target)^2) in fp32 (train_unet.py:517), dpred = 2 (pred - target) / n * dloss */
int sdb_mse(const void* pred, const void* target, int dtype, int64_t n, float dloss, float* loss_zeroed,
            void* dpred_or_null, void* stream);
/* eps = e0 + g (e0 - e1) (latent_diffusion.py:62-65): de2 = [(1+g) d | -g d] */
int sdb_cfg_combine_bwd(const void* d, void* de2, int dtype, int64_t half, float guidance, void* stream);
int sdb_sumsq(const float* x, int64_t n, float* out_zeroed, void* stream);
/* flat-bucket AdamW (torch.optim.AdamW semantics) with the gradient pre-scale (all-reduce mean) and
 * clip_grad_norm_(max_norm) folded in; gnorm_sq = device pointer to sum of squares of the UNSCALED flat grad */
int sdb_adamw(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
              float eps, float weight_decay, int step, float grad_scale, const float* gnorm_sq_or_null,
              float max_norm, void* stream);
/* sdb_adamw with the per-step scalars in device memory: hyper_dev = {lr, 1 - beta1^step, 1 - beta2^step}; lets a
 * captured CUDA graph of the whole train step be replayed with a new step count / learning rate. */
int sdb_adamw_dev(float* p, const float* g, float* m, float* v, int64_t n, const float* hyper_dev, float beta1,
                  float beta2, float eps, float weight_decay, float grad_scale, const float* gnorm_sq_or_null,
                  float max_norm, void* stream);

/* nn.Dropout in training mode (resnet2d.py:190; transformer.py:58,147): y = keep ? x / (1 - p) : 0 with a counter-based
 * mask keyed by (state_dev[0], site).  The same call on the gradient IS the backward (no mask is stored).  state_dev:
 * one int64 in device memory, advanced by the caller once per step (so CUDA-graph replays draw fresh masks); site: a
 * distinct number per dropout layer call within a step.  "Not covered by any oracle" (SURVEY 8c): statistical tests. */
int sdb_dropout(const void* x, void* y, int64_t n, float p, const int64_t* state_dev, uint64_t site, int dtype,
                void* stream);

/* ---- CLIP text encoder pieces (models/clip_model.py:56-96 wraps transformers.CLIPTextModel) ----
 * out[b*L + l, :] = token_embedding[ids[b, l]] + position_embedding[l]  (CLIPTextEmbeddings); ids int64. */
int sdb_embed_tokens(const int64_t* ids, const float* tok, const float* pos, void* out, int out_dtype, int B, int L,
                     int C, int vocab, void* stream);
/* causal softmax(QK^T scale)V over a short sequence (CLIPAttention with is_causal): n <= 80, d <= 64; head h at
 * columns [h*d, (h+1)*d), row strides in elements; dtype = SDB_BF16 | SDB_F32 for q, k, v and out. */
int sdb_attention_causal_small(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                               void* out, int64_t ldo, int B, int heads, int n, int d, float scale, int dtype,
                               void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SDB200_H_ */
