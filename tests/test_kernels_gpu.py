"""GPU: every libsdb200 kernel through the C ABI against a plain PyTorch fp32 reference of the same
op on the same (bf16-rounded) inputs."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

ops = None


@pytest.fixture(scope="module", autouse=True)
def _load():
    global ops
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from stable_diffusion_pytorch_b200 import ops as _ops
    ops = _ops
    # the torch references must be true fp32 (cuDNN / cuBLAS default to TF32 for fp32 inputs)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    assert ops.debug_timeout_flag() == 0, "a kernel hit its bounded mbarrier wait"


def rel_l2(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def rnd(*shape, seed=0, scale=1.0, dtype=torch.bfloat16):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to("cuda", dtype)


# ------------------------------------------------------------------ GEMM (tcgen05)
GEMM_CASES = [
    # M, N, K, bias, residual, act
    (128, 64, 64, False, False, 0),      # one tile, SW128, one k-block
    (128, 160, 160, True, False, 0),     # SW64 path (K % 64 != 0), N = 160
    (300, 160, 320, True, True, 0),      # M tail, bias + residual
    (4096, 480, 160, False, False, 0),   # fused QKV shape, many tiles (persistent loop, 2 acc stages)
    (1024, 1280, 160, True, False, 2),   # GEGLU epilogue
    (512, 160, 640, True, True, 0),      # FFN out
    (77 * 3, 320, 768, False, False, 0), # context K/V
    (32, 512, 160, True, False, 1),      # time MLP with SiLU epilogue, M < 128
    (2048, 96, 96, True, False, 0),      # small N
    (1000, 3040, 512, True, False, 0),   # batched temb projections, wide N
]


@pytest.mark.parametrize("M,N,K,has_bias,has_res,act", GEMM_CASES)
def test_gemm(M, N, K, has_bias, has_res, act):
    a = rnd(M, K, seed=1)
    w = rnd(N, K, seed=2, scale=1 / math.sqrt(K))
    bias = rnd(N, seed=3, dtype=torch.float32) if has_bias else None
    n_out = N // 2 if act == 2 else N
    res = rnd(M, n_out, seed=4) if has_res else None
    out = ops.gemm(a, w, bias, res, act)
    ref = a.float() @ w.float().t()
    if has_bias:
        ref = ref + bias
    if act == 1:
        ref = F.silu(ref)
    if act == 2:
        r4 = ref.view(M, N // 4, 4)      # columns interleaved in pairs: (a0, a1, g0, g1, ...)
        ref = (r4[..., 0:2] * F.gelu(r4[..., 2:4])).reshape(M, N // 2)
    if has_res:
        ref = ref + res.float()
    assert out.shape == ref.shape
    assert rel_l2(out.float(), ref) < 6e-3


def test_gemm_two_segments_fp32_out():
    M, K1, K2, N = 640, 320, 160, 320
    a1, a2 = rnd(M, K1, seed=1), rnd(M, K2, seed=2)
    w = rnd(N, K1 + K2, seed=3, scale=0.05)
    bias = rnd(N, seed=4, dtype=torch.float32)
    out = ops.gemm(a1, w, bias, a2=a2, out_dtype=torch.float32)
    ref = torch.cat([a1, a2], 1).float() @ w.float().t() + bias
    assert rel_l2(out, ref) < 1e-5


def test_gemm_strided_views():
    # q/k/v as column slices of a fused projection: lda != K
    M, K = 256, 160
    big = rnd(M, 3 * K, seed=5)
    w = rnd(64, K, seed=6, scale=0.1)
    out = ops.gemm(big[:, K:2 * K], w, out_dtype=torch.float32)
    ref = big[:, K:2 * K].float() @ w.float().t()
    assert rel_l2(out, ref) < 1e-5


@pytest.mark.parametrize("M,N,K", [(4096, 160, 160), (1000, 160, 640), (300, 256, 640), (128, 64, 64),
                                   (777, 320, 160)])
def test_gemm_layernorm_epilogue(M, N, K):
    """out = LN(a @ w^T + bias + residual) (post-norm, transformer.py:274-278) in the GEMM epilogue when the row
    fits one tile (N <= 256); N = 320 runs the GEMM and the LayerNorm kernel -- same result either way."""
    a = rnd(M, K, seed=1)
    w = rnd(N, K, seed=2, scale=1 / math.sqrt(K))
    bias = rnd(N, seed=3, dtype=torch.float32)
    res = rnd(M, N, seed=4)
    gamma = 1 + 0.1 * rnd(N, seed=5, dtype=torch.float32)
    beta = 0.1 * rnd(N, seed=6, dtype=torch.float32)
    n0 = ops.launch_count()
    out = ops.gemm(a, w, bias, residual=res, ln=(gamma, beta))
    assert ops.launch_count() - n0 == (1 if N <= 256 else 2)
    h = a.float() @ w.float().t() + bias + res.float()
    ref = F.layer_norm(h, (N,), gamma, beta, 1e-5)
    assert out.dtype == torch.bfloat16 and rel_l2(out.float(), ref) < 5e-3
    old = ops._LN_FUSED
    try:
        ops._LN_FUSED = False
        two = ops.gemm(a, w, bias, residual=res, ln=(gamma, beta))
    finally:
        ops._LN_FUSED = old
    assert rel_l2(out.float(), two.float()) < 5e-3


# ------------------------------------------------------------------ conv3x3 (tcgen05 implicit GEMM)
CONV_CASES = [
    # B, H, W, Cin, Cout, temb, residual, down2
    (1, 16, 16, 64, 64, False, False, False),
    (2, 32, 32, 160, 160, True, True, False),
    (3, 8, 8, 64, 128, True, False, False),     # 2 images per tile + batch tail
    (2, 16, 16, 320, 160, False, True, False),
    (2, 32, 32, 160, 160, False, False, True),  # DownSample: conv then nearest x0.5
    (1, 64, 64, 480, 320, True, True, False),   # virtual-concat sized K, 2 N tiles
    (1, 4, 4, 32, 32, False, False, False),     # tiny map
    (2, 16, 16, 96, 96, False, False, False),
    (1, 128, 128, 64, 64, False, False, False), # W = 128: one image row per tile -> the halo loader (one 3 x 130 pixel
                                                # block per channel block, nine row-shifted views), CTA pairs
    (2, 8, 256, 128, 128, True, True, False),   # halo: two tiles per image row, two channel blocks, residual + temb
    (1, 5, 128, 192, 96, True, False, False),   # halo: three channel blocks, odd tile count (a ghost tile in the pair walk)
    (1, 2, 128, 64, 64, False, True, False),    # halo without CTA pairs (too few tiles)
    (1, 16, 128, 256, 512, False, False, False),  # halo at BN = 256: two weight-ring stages fewer, two N tiles
    (1, 6, 128, 64, 64, False, False, True),    # W = 128 with DownSample (direct epilogue)
    (2, 12, 20, 64, 64, True, True, False),     # W without a power-of-two factor >= 8: clipped 8-wide boxes
    (1, 10, 6, 32, 48, False, True, False),     # W < 8
    (3, 6, 12, 64, 32, True, False, True),      # ragged + DownSample
    (1, 24, 40, 160, 160, True, True, False),
]


@pytest.mark.parametrize("B,H,W,Cin,Cout,has_temb,has_res,down2", CONV_CASES)
def test_conv3x3(B, H, W, Cin, Cout, has_temb, has_res, down2):
    x = rnd(B, H, W, Cin, seed=1)
    w = rnd(Cout, 3, 3, Cin, seed=2, scale=1 / math.sqrt(9 * Cin))
    bias = rnd(Cout, seed=3, dtype=torch.float32)
    temb = rnd(B, Cout, seed=4, dtype=torch.float32) if has_temb else None
    Ho, Wo = (H // 2, W // 2) if down2 else (H, W)
    res = rnd(B, Ho, Wo, Cout, seed=5) if has_res else None
    out = ops.conv3x3(x, w, bias, temb, res, down2=down2, out_dtype=torch.float32)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), bias, padding=1)
    if down2:
        ref = F.interpolate(ref, scale_factor=0.5, mode="nearest")
    if has_temb:
        ref = ref + temb[:, :, None, None]
    ref = ref.permute(0, 2, 3, 1)
    if has_res:
        ref = ref + res.float()
    assert rel_l2(out, ref) < 1e-5
    # bf16 output: the staged epilogue (swizzled smem tile + TMA store, TMA-loaded residual) when the
    # output keeps the input resolution, the direct one otherwise
    outb = ops.conv3x3(x, w, bias, temb, res, down2=down2, out_dtype=torch.bfloat16)
    assert outb.dtype == torch.bfloat16 and outb.shape == ref.shape
    assert rel_l2(outb.float(), ref) < 4e-3


@pytest.fixture
def sums_any_size(monkeypatch):
    """The product only asks for epilogue sums on tensors beyond the L2 (SDB_GN_SUMS_MIN_MB); the tests take them all."""
    monkeypatch.setattr(ops, "_GN_SUMS_MIN_BYTES", 0)


SUMS_CASES = [
    # B, H, W, Cin, Cout, kind
    (2, 16, 16, 32, 64, "plain"),
    (2, 64, 64, 160, 160, "temb"),     # the UNet's level-0 ResBlock conv1 (BN = 160: 80 column pairs, tail slab)
    (3, 32, 32, 64, 320, "res"),       # two n-tiles, residual
    (2, 16, 16, 128, 128, "up2"),      # four phase launches add into one buffer
    (1, 8, 16, 32, 80, "plain"),       # one tile per image, 16-column tail slab
    (2, 64, 64, 128, 256, "plain"),    # BN = 256, CTA pairs, several tiles per CTA and image
    (40, 16, 16, 64, 64, "res"),       # many images: the per-thread partials are flushed at every image change
]


@pytest.mark.parametrize("B,H,W,Cin,Cout,kind", SUMS_CASES)
def test_conv3x3_channel_sums(B, H, W, Cin, Cout, kind, sums_any_size):
    """GroupNorm statistics out of the producing conv's epilogue (resnet2d.py:171-177,187-198): inside a
    gn_sums_scope the conv also leaves per-(image, channel) sum / sum of squares of its bf16 output; group_norm
    then normalises without its statistics pass.  Checked: the conv output is unchanged, the sums match torch's, and
    the normalised result matches the two-pass route."""
    from stable_diffusion_pytorch_b200 import _pack
    up2 = kind == "up2"
    x = rnd(B, H, W, Cin, seed=1)
    bias = rnd(Cout, seed=3, dtype=torch.float32)
    temb = rnd(B, Cout, seed=4, dtype=torch.float32) if kind == "temb" else None
    Ho, Wo = (2 * H, 2 * W) if up2 else (H, W)
    res = rnd(B, Ho, Wo, Cout, seed=5) if kind == "res" else None
    if up2:
        w = _pack.conv3x3_up_w(rnd(Cout, Cin, 3, 3, seed=2, scale=1 / math.sqrt(9 * Cin), dtype=torch.float32))
    else:
        w = rnd(Cout, 3, 3, Cin, seed=2, scale=1 / math.sqrt(9 * Cin))
    plain = ops.conv3x3(x, w, bias, temb, res, up2=up2)
    gamma = rnd(Cout, seed=6, dtype=torch.float32) + 1.0
    beta = rnd(Cout, seed=7, dtype=torch.float32)
    groups = 16
    want = ops.group_norm(plain, gamma, beta, groups, ops.ACT_SILU)
    with ops.gn_sums_scope(x.device):
        out = ops.conv3x3(x, w, bias, temb, res, up2=up2, sums=True)
        sums = ops._sums_of(out)
        assert sums is not None, "the epilogue did not take the channel sums for this shape"
        got = ops.group_norm(out, gamma, beta, groups, ops.ACT_SILU)
        sums = sums.clone().view(B, Cout, 2)
    with ops.gn_sums_scope(x.device):         # 64-bit fixed-point atomics: bit-identical from run to run
        out2 = ops.conv3x3(x, w, bias, temb, res, up2=up2, sums=True)
        assert torch.equal(ops._sums_of(out2).view(B, Cout, 2), sums)
        assert torch.equal(ops.group_norm(out2, gamma, beta, groups, ops.ACT_SILU), got)
    assert torch.equal(out, plain)
    o = out.double().view(B, Ho * Wo, Cout)
    ref_s, ref_q = o.sum(1), (o * o).sum(1)
    got_s, got_q = sums[..., 0].double() / ops.SUM_SCALE, sums[..., 1].double() / ops.SQ_SCALE
    assert float((got_s - ref_s).abs().max()) < 1e-3 * float(ref_q.max().sqrt() * math.sqrt(Ho * Wo)) + 1e-3
    assert rel_l2(got_q, ref_q) < 1e-5
    assert rel_l2(got.float(), want.float()) < 2e-3
    assert ops._sums_of(out) is None          # outside the scope the tag is void


def test_channel_sums_fallbacks_and_concat(sums_any_size):
    """Shapes the epilogue cannot cover keep the usual route; a GroupNorm over the virtual concat of two tagged
    tensors (the UNet's up path) whose group straddles the boundary; a GEMM producer (SpatialTransformer.proj_out)."""
    gamma = rnd(112, seed=6, dtype=torch.float32) + 1.0
    beta = rnd(112, seed=7, dtype=torch.float32)
    x = rnd(2, 16, 16, 32, seed=1)
    wa = rnd(80, 3, 3, 32, seed=2, scale=0.06)
    wb = rnd(32, 3, 3, 32, seed=3, scale=0.06)
    a0, b0 = ops.conv3x3(x, wa, None), ops.conv3x3(x, wb, None)
    want = ops.group_norm(a0, gamma, beta, 16, ops.ACT_SILU, x2=b0)            # 7 channels per group: group 11 straddles
    with ops.gn_sums_scope(x.device):
        a1, b1 = ops.conv3x3(x, wa, None, sums=True), ops.conv3x3(x, wb, None, sums=True)
        assert ops._sums_of(a1) is not None and ops._sums_of(b1) is not None
        got = ops.group_norm(a1, gamma, beta, 16, ops.ACT_SILU, x2=b1)
        half = ops.group_norm(a1, gamma, beta, 16, ops.ACT_SILU, x2=b0)        # one side untagged: two-pass route
        # 8 x 8 images: two images per 128-pixel tile -> not taken, no tag, same result
        xs = rnd(4, 8, 8, 32, seed=4)
        small = ops.conv3x3(xs, wb, None, sums=True)
        assert ops._sums_of(small) is None
        assert torch.equal(small, ops.conv3x3(xs, wb, None))
        # GEMM producer with residual, rows_per_image = 256
        t = rnd(512, 64, seed=8)
        wg = rnd(96, 64, seed=9, scale=0.12)
        r = rnd(512, 96, seed=10)
        o = ops.gemm(t, wg, None, residual=r, sums_rows=256)
        sg = ops._sums_of(o)
        assert sg is not None
        sg = sg.clone().view(2, 96, 2)
    assert rel_l2(got.float(), want.float()) < 2e-3
    assert torch.equal(half, want)
    assert torch.equal(o, ops.gemm(t, wg, None, residual=r))
    of = o.double().view(2, 256, 96)
    assert rel_l2(sg[..., 1].double() / ops.SQ_SCALE, (of * of).sum(1)) < 1e-5
    assert float((sg[..., 0].double() / ops.SUM_SCALE - of.sum(1)).abs().max()) < 5e-2


UP_CASES = [
    # B, H, W, Cin, Cout, residual   (low-resolution input H x W -> 2H x 2W)
    (1, 8, 8, 64, 64, False),
    (2, 32, 32, 320, 320, False),     # the UNet's UpSample (config 2 runs it at B = 32)
    (2, 16, 16, 160, 96, True),       # K tail (160 = 2.5 k-blocks), residual at the output resolution
    (3, 6, 10, 64, 32, False),        # ragged: clipped boxes, W without a power-of-two factor >= 8
    (1, 64, 64, 512, 512, False),     # VAE decoder shape: BN = 256, two N tiles
    (1, 6, 128, 64, 64, True),        # W = 128: the phases run on the halo loader (2 x 129 pixel blocks)
    (2, 4, 256, 128, 64, False),
    (1, 4, 4, 32, 48, True),
]


@pytest.mark.parametrize("B,H,W,Cin,Cout,has_res", UP_CASES)
def test_conv3x3_up2_phase_collapsed(B, H, W, Cin, Cout, has_res):
    """UpSample (resnet2d.py:46-63: nearest x2 THEN conv3x3) as four 2x2-tap phase convolutions over the
    low-resolution tensor, against F.interpolate + F.conv2d on the same bf16-rounded inputs.  The collapsed weights
    are sums of up to four bf16 taps rounded once more to bf16: rel-L2 stays at the bf16 level."""
    from stable_diffusion_pytorch_b200 import _pack
    x = rnd(B, H, W, Cin, seed=1)
    w_oihw = rnd(Cout, Cin, 3, 3, seed=2, scale=1 / math.sqrt(9 * Cin), dtype=torch.float32)
    bias = rnd(Cout, seed=3, dtype=torch.float32)
    res = rnd(B, 2 * H, 2 * W, Cout, seed=5) if has_res else None
    wp = _pack.conv3x3_up_w(w_oihw)
    assert tuple(wp.shape) == (4, Cout, 2, 2, Cin) and wp.dtype == torch.bfloat16
    ref = F.conv2d(F.interpolate(x.float().permute(0, 3, 1, 2), scale_factor=2, mode="nearest"), w_oihw, bias,
                   padding=1).permute(0, 2, 3, 1)
    # exact-arithmetic check of the collapse itself: the same phases evaluated by torch in fp32
    if has_res:
        ref = ref + res.float()
    for odt, tol in ((torch.float32, 6e-3), (torch.bfloat16, 8e-3)):
        out = ops.conv3x3(x, wp, bias, None, res, up2=True, out_dtype=odt)
        assert out.dtype == odt and tuple(out.shape) == (B, 2 * H, 2 * W, Cout)
        e = rel_l2(out.float(), ref)
        assert e < tol, (odt, e)
    # and against the un-fused route of the same library (upsample2x kernel + 3x3 conv): same answer to bf16 noise
    w9 = w_oihw.permute(0, 2, 3, 1).to(torch.bfloat16).contiguous()
    unf = ops.conv3x3(ops.upsample2x(x), w9, bias, None, res, out_dtype=torch.float32)
    assert rel_l2(ops.conv3x3(x, wp, bias, None, res, up2=True, out_dtype=torch.float32), unf) < 8e-3


def test_conv_small_in_out():
    B, H, W = 2, 24, 40
    x = rnd(B, 4, H, W, seed=1, dtype=torch.float32)
    w = rnd(160, 4, 3, 3, seed=2, scale=0.2, dtype=torch.float32)
    b = rnd(160, seed=3, dtype=torch.float32)
    out = ops.conv3x3_in(x, w, b, out_dtype=torch.float32)
    ref = F.conv2d(x, w, b, padding=1).permute(0, 2, 3, 1)
    assert rel_l2(out, ref) < 1e-5
    xh = rnd(B, H, W, 160, seed=4)
    for cout in (3, 4, 8):
        w2 = rnd(cout, 160, 3, 3, seed=5, scale=0.03, dtype=torch.float32)
        b2 = rnd(cout, seed=6, dtype=torch.float32)
        out2 = ops.conv3x3_out(xh, w2, b2, out_dtype=torch.float32)
        ref2 = F.conv2d(xh.float().permute(0, 3, 1, 2), w2, b2, padding=1)
        assert rel_l2(out2, ref2) < 1e-5
    w3 = rnd(4, 4, 1, 1, seed=7, dtype=torch.float32)
    b3 = rnd(4, seed=8, dtype=torch.float32)
    assert rel_l2(ops.conv1x1_nchw_small(x, w3, b3), F.conv2d(x, w3, b3)) < 1e-6


# ------------------------------------------------------------------ norms / layout
@pytest.mark.parametrize("B,H,W,C1,C2,groups,act", [
    (2, 16, 16, 160, 0, 32, 1), (2, 32, 32, 320, 160, 32, 1), (3, 8, 8, 320, 0, 2, 1),
    (1, 64, 64, 160, 0, 32, 0), (2, 16, 16, 64, 32, 16, 1), (1, 128, 128, 64, 0, 32, 1)])
def test_groupnorm(B, H, W, C1, C2, groups, act):
    x1 = rnd(B, H, W, C1, seed=1) + 0.3
    x2 = rnd(B, H, W, C2, seed=2, scale=2.0) if C2 else None
    C = C1 + C2
    gamma = 1 + 0.1 * rnd(C, seed=3, dtype=torch.float32)
    beta = 0.1 * rnd(C, seed=4, dtype=torch.float32)
    out = ops.group_norm(x1, gamma, beta, groups, act, x2=x2, out_dtype=torch.float32)
    xc = x1 if x2 is None else torch.cat([x1, x2], -1)
    ref = F.group_norm(xc.float().permute(0, 3, 1, 2), groups, gamma, beta, 1e-5)
    if act:
        ref = F.silu(ref)
    assert rel_l2(out, ref.permute(0, 2, 3, 1)) < 2e-5


GN_FUSED_CASES = [
    # B, H, W, C1, C2, groups, act       (bf16 in / bf16 out: the single-pass cluster kernel)
    (2, 64, 64, 160, 0, 32, 1),       # UNet level 0: 1.3 MB per image -> 16-CTA cluster
    (3, 32, 32, 320, 320, 32, 1),     # virtual concat, 640 channels
    (2, 32, 32, 320, 160, 32, 1),     # 480 channels: 15 per group, concat boundary inside a group's neighbourhood
    (2, 32, 32, 320, 0, 2, 1),        # the bottleneck's groups = 2
    (5, 8, 8, 64, 0, 32, 0),          # tiny image: one CTA, one piece, no SiLU
    (1, 12, 20, 96, 32, 16, 1),       # pixels not divisible by the cluster size
    (2, 64, 64, 160, 160, 32, 0),     # 2.6 MB per image: 16 CTAs with 164 KB slices
    (1, 128, 128, 256, 0, 32, 1),     # 8 MB per image: does not fit -> falls back to the three-kernel route
]


@pytest.mark.parametrize("route", ["coop", "cluster"])
@pytest.mark.parametrize("B,H,W,C1,C2,groups,act", GN_FUSED_CASES)
def test_groupnorm_one_launch_routes(B, H, W, C1, C2, groups, act, route):
    """bf16 -> bf16 GroupNorm(+SiLU) in one launch: "coop" = persistent two-phase kernel with a grid barrier (the
    default), "cluster" = image resident in the shared memory of a thread-block cluster (falls back when it does not
    fit).  Both against torch's fp32 GroupNorm and against the three-kernel route; both deterministic."""
    x1 = rnd(B, H, W, C1, seed=1) + 0.3
    x2 = rnd(B, H, W, C2, seed=2, scale=2.0) if C2 else None
    C = C1 + C2
    gamma = 1 + 0.1 * rnd(C, seed=3, dtype=torch.float32)
    beta = 0.1 * rnd(C, seed=4, dtype=torch.float32)
    xc = x1 if x2 is None else torch.cat([x1, x2], -1)
    ref = F.group_norm(xc.float().permute(0, 3, 1, 2), groups, gamma, beta, 1e-5)
    if act:
        ref = F.silu(ref)
    ref = ref.permute(0, 2, 3, 1)
    old = ops._GN_FUSED
    try:
        ops._GN_FUSED = route
        n0 = ops.launch_count()
        out = ops.group_norm(x1, gamma, beta, groups, act, x2=x2)
        launches = ops.launch_count() - n0
        again = ops.group_norm(x1, gamma, beta, groups, act, x2=x2)
        ops._GN_FUSED = False
        three = ops.group_norm(x1, gamma, beta, groups, act, x2=x2)
    finally:
        ops._GN_FUSED = old
    fits_cluster = H * W * C * 2 <= 3 * 2 ** 20
    assert launches == (1 if route == "coop" or fits_cluster else 3)
    assert out.dtype == torch.bfloat16 and tuple(out.shape) == (B, H, W, C)
    assert rel_l2(out.float(), ref) < 4e-3          # bf16 rounding of the output
    assert rel_l2(out.float(), three.float()) < 3e-3
    assert torch.equal(out, again)                  # deterministic


@pytest.mark.parametrize("rows,C,dtype", [(1000, 160, torch.bfloat16), (513, 320, torch.float32),
                                           (64, 64, torch.bfloat16), (100, 1024, torch.float32)])
def test_add_layernorm(rows, C, dtype):
    x, r = rnd(rows, C, seed=1, dtype=dtype), rnd(rows, C, seed=2, dtype=dtype)
    gamma = 1 + 0.1 * rnd(C, seed=3, dtype=torch.float32)
    beta = 0.1 * rnd(C, seed=4, dtype=torch.float32)
    out = ops.add_layernorm(x, r, gamma, beta)
    ref = F.layer_norm(x.float() + r.float(), (C,), gamma, beta, 1e-5)
    assert rel_l2(out.float(), ref) < (5e-3 if dtype == torch.bfloat16 else 1e-5)


def test_layout_and_elementwise():
    x = rnd(2, 5, 7, 9, seed=1, dtype=torch.float32)
    y = ops.nchw_to_nhwc(x, torch.bfloat16)
    assert torch.equal(y, x.permute(0, 2, 3, 1).to(torch.bfloat16))
    z = ops.nhwc_to_nchw(y, torch.float32)
    assert torch.equal(z, x.to(torch.bfloat16).float())
    u = ops.upsample2x(rnd(2, 3, 5, 16, seed=2))
    assert torch.equal(u, F.interpolate(rnd(2, 3, 5, 16, seed=2).permute(0, 3, 1, 2).float(), scale_factor=2,
                                        mode="nearest").permute(0, 2, 3, 1).to(torch.bfloat16))
    s = ops.silu(x)
    assert rel_l2(s, F.silu(x)) < 1e-6
    img = ops.to_uint8_image(rnd(2, 3, 8, 8, seed=3, dtype=torch.float32))
    ref = ((rnd(2, 3, 8, 8, seed=3, dtype=torch.float32) + 1) / 2).clamp(0, 1).mul(255).to(torch.uint8)
    assert (img.int() - ref.permute(0, 2, 3, 1).int()).abs().max() <= 1


# ------------------------------------------------------------------ attention
@pytest.mark.parametrize("B,heads,n,m,d", [
    (2, 8, 256, 256, 20), (1, 8, 1024, 1024, 20), (2, 8, 300, 77, 20), (2, 4, 64, 64, 16),
    (1, 1, 256, 256, 64), (1, 1, 200, 200, 128), (2, 2, 128, 77, 32),
    # head dim 128: the tcgen05 kernel of the VAE bottleneck (ragged rows / keys, several heads, the config-3 shape)
    (2, 2, 1024, 1024, 128), (1, 1, 130, 77, 128), (1, 3, 64, 640, 128), (2, 1, 4096, 4096, 128)])
def test_attention(B, heads, n, m, d):
    dm = heads * d
    qkv = rnd(B, n, 3 * dm, seed=1)
    q = qkv[:, :, :dm]
    if m == n:
        k, v = qkv[:, :, dm:2 * dm], qkv[:, :, 2 * dm:]
    else:
        kv = rnd(B, m, 2 * dm, seed=2)
        k, v = kv[:, :, :dm], kv[:, :, dm:]
    scale = 1 / math.sqrt(d)
    out = ops.attention(q, k, v, heads, d, scale)
    qf = q.float().view(B, n, heads, d).transpose(1, 2)
    kf = k.float().view(B, m, heads, d).transpose(1, 2)
    vf = v.float().view(B, m, heads, d).transpose(1, 2)
    ref = torch.softmax(qf @ kf.transpose(-1, -2) * scale, -1) @ vf
    ref = ref.transpose(1, 2).reshape(B, n, dm)
    assert rel_l2(out.float(), ref) < 1e-2


@pytest.mark.parametrize("B,heads,n,m,d", [
    (1, 1, 256, 128, 20), (1, 2, 256, 256, 20), (2, 8, 1024, 1024, 20), (2, 8, 300, 77, 20),
    (1, 8, 4096, 4096, 20), (1, 3, 128, 640, 16), (2, 4, 1000, 333, 8),
    # short contexts: <= 96 keys take the tcgen05 short-context kernel (one template per 16 keys; odd tile counts,
    # ragged last tiles, chunked query ranges), 97..128 the streaming mma.sync kernel
    (1, 8, 4096, 77, 20), (2, 2, 50, 5, 20), (1, 2, 1000, 100, 8), (1, 1, 17, 16, 4), (3, 8, 1024, 77, 20),
    (1, 2, 2500, 96, 20), (2, 3, 700, 48, 12), (1, 1, 129, 33, 16), (1, 2, 1300, 64, 20), (1, 1, 128, 1, 8),
    (40, 8, 1200, 77, 20)])
def test_attention_p32(B, heads, n, m, d):
    """tcgen05 attention on head-padded operands vs fp32 softmax attention of the same bf16 inputs."""
    g = torch.Generator(device="cpu").manual_seed(n + m)

    def padded(rows, ones_col):
        t = torch.zeros(B, rows, heads, 32)
        t[..., :d] = torch.randn(B, rows, heads, d, generator=g)
        if ones_col:
            t[..., d] = 1.0
            t[..., d + 1:] = 3.0      # finite garbage beyond the ones column must not matter
        return t.reshape(B, rows, heads * 32).to("cuda", torch.bfloat16)

    q, k, v = padded(n, False), padded(m, False), padded(m, True)
    scale = 1 / math.sqrt(d)
    out = ops.attention_p32(q, k, v, heads, d, scale)
    qf = q.float().view(B, n, heads, 32)[..., :d].transpose(1, 2)
    kf = k.float().view(B, m, heads, 32)[..., :d].transpose(1, 2)
    vf = v.float().view(B, m, heads, 32)[..., :d].transpose(1, 2)
    ref = torch.softmax(qf @ kf.transpose(-1, -2) * scale, -1) @ vf
    ref = ref.transpose(1, 2).reshape(B, n, heads * d)
    assert out.shape == ref.shape
    assert rel_l2(out.float(), ref) < 1e-2


@pytest.mark.parametrize("growth", [0.3, 2.0, 12.0])
def test_attention_p32_growing_scores(growth):
    """Scores that keep growing along the keys: exercises the lazy rescale of the TMEM accumulator
    (reference maximum lagging by > 2^8) and, for the largest growth, the redo path (> 2^60 per tile)."""
    B, heads, n, m, d = 1, 2, 384, 512, 20
    g = torch.Generator(device="cpu").manual_seed(7)
    q = torch.zeros(B, n, heads, 32)
    k = torch.zeros(B, m, heads, 32)
    v = torch.zeros(B, m, heads, 32)
    q[..., :d] = 0.05 * torch.randn(B, n, heads, d, generator=g)
    k[..., :d] = 0.05 * torch.randn(B, m, heads, d, generator=g)
    q[..., 0] = 4.0
    k[..., 0] = growth * torch.arange(m).float()[None, :, None] / 4.0   # log2-score slope ~ 0.32 * growth per key
    v[..., :d] = torch.randn(B, m, heads, d, generator=g)
    v[..., d] = 1.0
    q, k, v = (t.reshape(B, -1, heads * 32).to("cuda", torch.bfloat16) for t in (q, k, v))
    scale = 1 / math.sqrt(d)
    out = ops.attention_p32(q, k, v, heads, d, scale)
    qf = q.float().view(B, n, heads, 32)[..., :d].transpose(1, 2)
    kf = k.float().view(B, m, heads, 32)[..., :d].transpose(1, 2)
    vf = v.float().view(B, m, heads, 32)[..., :d].transpose(1, 2)
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * scale, -1) @ vf).transpose(1, 2).reshape(B, n, heads * d)
    assert torch.isfinite(out.float()).all()
    assert rel_l2(out.float(), ref) < 1e-2


@pytest.mark.parametrize("growth", [0.3, 2.0, 12.0])
def test_attention_d128_growing_scores(growth):
    """Head dim 128 (tcgen05 kernel): scores that keep growing along the keys exercise the lazy rescale of the
    128-column accumulator and of the register row sum."""
    B, heads, n, m, d = 1, 2, 256, 512, 128
    g = torch.Generator(device="cpu").manual_seed(9)
    q = 0.02 * torch.randn(B, n, heads, d, generator=g)
    k = 0.02 * torch.randn(B, m, heads, d, generator=g)
    q[..., 0] = 4.0
    k[..., 0] = growth * torch.arange(m).float()[None, :, None] / 4.0
    v = torch.randn(B, m, heads, d, generator=g)
    q, k, v = (t.reshape(B, -1, heads * d).to("cuda", torch.bfloat16) for t in (q, k, v))
    scale = 1 / math.sqrt(d)
    out = ops.attention(q, k, v, heads, d, scale)
    qf, kf, vf = (t.float().view(B, -1, heads, d).transpose(1, 2) for t in (q, k, v))
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * scale, -1) @ vf).transpose(1, 2).reshape(B, n, heads * d)
    assert torch.isfinite(out.float()).all()
    assert rel_l2(out.float(), ref) < 1e-2


# ------------------------------------------------------------------ sampler kernels (bit-exact in fp32)
def test_timestep_embed():
    half = 80
    freq = torch.exp(math.log(10000) / half * torch.arange(0, end=half, dtype=torch.float32))
    t = torch.tensor([0, 1, 7, 49, 500, 999], dtype=torch.long)
    ref = torch.cat([torch.sin(t[:, None].float() * freq[None]), torch.cos(t[:, None].float() * freq[None])], -1)
    out = ops.timestep_embed(t.cuda(), freq.cuda())
    assert (out.cpu() - ref).abs().max() < 2e-6  # CPU (Sleef) vs CUDA sinf differ by <= 2 ulp of |result| <= 1
