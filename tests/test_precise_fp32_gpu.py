"""fp32 parity mode (north_star: "a single UNet forward within 1e-3 relative L2 in fp32"): the same drop-in
modules run with ``precision("fp32")`` -- fp32 activations, fp32 FMA GEMM / conv / attention kernels
(csrc/precise_f32.cu) -- against the committed outputs of the unmodified reference and the CPU oracle."""
import numpy as np
import pytest
import torch

from oracle import sd_oracle as O
from tests.golden import cases
from tests.test_parity_gpu import StubText, build_unet, rel_l2

pytestmark = pytest.mark.gpu
DEV = "cuda"
UNET_FP32_TOL = 1e-3      # north_star: fp32 forward vs the fp32 reference
SAMPLE_FP32_TOL = 5e-3    # final latents of a CFG sampling run in fp32 (SURVEY 8c proposal)


@pytest.fixture(scope="module", autouse=True)
def _gpu():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    yield
    import stable_diffusion_pytorch_b200 as sdb
    assert sdb.compute_dtype() == torch.bfloat16      # every test restores the product precision


def test_gemm_f32_epilogues_vs_torch():
    from stable_diffusion_pytorch_b200 import ops
    from stable_diffusion_pytorch_b200._lib import ACT_GEGLU, ACT_NONE, ACT_SILU
    from stable_diffusion_pytorch_b200._pack import geglu_w, precision
    g = torch.Generator().manual_seed(3)
    M, K1, K2, N = 203, 37, 22, 76          # nothing is a multiple of the 64x64x16 tile
    a = torch.randn(M, K1 + 5, generator=g)[:, :K1]      # row-strided view
    a2 = torch.randn(M, K2, generator=g)
    w = torch.randn(N, K1 + K2, generator=g) * 0.2
    b = torch.randn(N, generator=g)
    r = torch.randn(M, N, generator=g)
    ref = torch.cat([a, a2], 1).double() @ w.double().T + b.double()
    out = ops.gemm(a.to(DEV), w.to(DEV), b.to(DEV), residual=r.to(DEV), act=ACT_NONE, a2=a2.to(DEV))
    assert out.dtype == torch.float32 and rel_l2(out, ref + r.double()) < 2e-6
    out = ops.gemm(a.to(DEV), w.to(DEV), b.to(DEV), act=ACT_SILU, a2=a2.to(DEV))
    assert rel_l2(out, torch.nn.functional.silu(ref)) < 2e-6
    # GEGLU on the pair-interleaved rows the packer produces
    lin = torch.nn.Linear(K1 + K2, 2 * 40)
    with precision("fp32"):
        wi, bi = geglu_w(lin.weight, lin.bias)
        assert wi.dtype == torch.float32
        y = ops.gemm(torch.cat([a, a2], 1).contiguous().to(DEV), wi.to(DEV), bi.to(DEV), act=ACT_GEGLU)
    h = torch.cat([a, a2], 1).double() @ lin.weight.detach().double().T + lin.bias.detach().double()
    want = h[:, :40] * torch.nn.functional.gelu(h[:, 40:])
    assert tuple(y.shape) == (M, 40) and rel_l2(y, want) < 2e-6


@pytest.mark.parametrize("down2", [False, True])
def test_conv3x3_f32_vs_torch(down2):
    from stable_diffusion_pytorch_b200 import ops
    g = torch.Generator().manual_seed(4)
    B, H, W, Cin, Cout = 2, 10, 14, 13, 21
    x = torch.randn(B, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * 0.2
    b = torch.randn(Cout, generator=g)
    temb = torch.randn(B, Cout, generator=g)
    ref = torch.nn.functional.conv2d(x.double(), w.double(), b.double(), padding=1) + temb.double()[:, :, None, None]
    if down2:
        ref = torch.nn.functional.interpolate(ref, scale_factor=0.5, mode="nearest")      # resnet2d.py:122-123
    res = torch.randn(ref.shape, generator=g)
    out = ops.conv3x3(x.permute(0, 2, 3, 1).contiguous().to(DEV), w.permute(0, 2, 3, 1).contiguous().to(DEV),
                      b.to(DEV), temb=temb.to(DEV), residual=res.permute(0, 2, 3, 1).contiguous().to(DEV), down2=down2)
    assert out.dtype == torch.float32
    assert rel_l2(out.permute(0, 3, 1, 2), ref + res.double()) < 2e-6


@pytest.mark.parametrize("d,n,m", [(20, 300, 77), (20, 130, 130), (64, 70, 45), (128, 40, 33)])
def test_attention_f32_vs_torch(d, n, m):
    from stable_diffusion_pytorch_b200 import ops
    g = torch.Generator().manual_seed(5)
    B, heads = 2, 3
    q = torch.randn(B, n, heads * d, generator=g)
    kv = torch.randn(B, m, 2 * heads * d, generator=g)          # K | V as strided views of one GEMM output
    k, v = kv[:, :, :heads * d], kv[:, :, heads * d:]
    scale = d ** -0.5
    qd, kd, vd = (t.double().reshape(B, -1, heads, d).transpose(1, 2) for t in (q, k, v))
    ref = (torch.softmax(qd @ kd.transpose(-1, -2) * scale, -1) @ vd).transpose(1, 2).reshape(B, n, heads * d)
    kvd = kv.to(DEV)
    out = ops.attention(q.to(DEV), kvd[:, :, :heads * d], kvd[:, :, heads * d:], heads, d, scale)
    assert out.dtype == torch.float32 and rel_l2(out, ref) < 2e-6


@pytest.mark.parametrize("name", list(cases.UNET_CASES))
def test_unet_forward_fp32_vs_reference_golden(golden, name):
    from stable_diffusion_pytorch_b200 import precision
    cfg, seed, B, H, W = cases.UNET_CASES[name]
    m, _ = build_unet(cfg, seed)
    x, t, ctx = cases.unet_inputs(cfg, B, H, W, seed)
    with torch.no_grad(), precision("fp32"):
        y = m(x.to(DEV), t.to(DEV), ctx.to(DEV))
    assert y.dtype == torch.float32 and tuple(y.shape) == (B, cfg.latent_channels, H, W)
    err = rel_l2(y, golden[f"unet_{name}"])
    print(f"unet fp32 {name}: rel-L2 vs reference = {err:.3e}")
    assert err < UNET_FP32_TOL
    with torch.no_grad():                                  # and the same module goes back to the bf16 product path
        yb = m(x.to(DEV), t.to(DEV), ctx.to(DEV))
    eb = rel_l2(yb, golden[f"unet_{name}"])
    assert err < eb < 2e-2


def test_unet_forward_fp32_config2_size_vs_oracle():
    """One 64x64 latent (config-2 geometry) in fp32 against the oracle run here."""
    from stable_diffusion_pytorch_b200 import precision
    cfg = cases.DEFAULT
    m, sd = build_unet(cfg, 13)
    x, t, ctx = cases.unet_inputs(cfg, 1, 64, 64, 99)
    with torch.no_grad():
        ref = O.unet_forward(sd, cfg, x, t, ctx)
        with precision("fp32"):
            y = m(x.to(DEV), t.to(DEV), ctx.to(DEV))
    e = rel_l2(y, ref)
    print(f"unet fp32 default 64x64: {e:.3e}")
    assert e < UNET_FP32_TOL


@pytest.mark.parametrize("name", ["tiny_16", "default_16", "config1_32_50"])
def test_sample_fp32_vs_reference_golden(golden, name):
    """CFG sampling entirely in fp32 (BASELINE config 1 is the reference's own fp32 run): pred_noise and the final
    latents of the multi-step loop, identical host noise."""
    from stable_diffusion_pytorch_b200 import precision
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler, LatentDiffusion
    cfg, seed, B, H, W, steps, g = cases.SAMPLE_CASES[name]
    m, _ = build_unet(cfg, seed)
    x, ctx, uncond = cases.sample_inputs(cfg, B, H, W, seed)
    ld = LatentDiffusion(m, None, StubText(uncond), DDPMScheduler(DDPMConfig()))
    t = torch.full((B,), 7, dtype=torch.long)
    with torch.no_grad(), precision("fp32"):
        pn = ld.pred_noise(x.to(DEV), t.to(DEV), ctx.to(DEV), guidance_scale=g)
        e = rel_l2(pn, golden[f"prednoise_{name}"])
        print(f"pred_noise fp32 {name}: {e:.3e}")
        assert e < UNET_FP32_TOL
        torch.manual_seed(cases.SAMPLE_NOISE_SEED)
        xs = ld.sample(x.to(DEV), ctx.to(DEV), guidance_scale=g, time_steps=steps)
    e = rel_l2(xs, golden[f"sample_{name}"])
    print(f"sample fp32 {name}: {e:.3e}")
    assert e < SAMPLE_FP32_TOL


@pytest.mark.parametrize("name", list(cases.VAE_CASES))
def test_vae_decode_fp32_vs_reference_golden(golden, name):
    from stable_diffusion_pytorch_b200 import precision
    from stable_diffusion_pytorch_b200.models import AutoEncoderKL, AutoencoderConfig
    cfg, seed, B, H, W = cases.VAE_CASES[name]
    m = AutoEncoderKL(AutoencoderConfig(autoencoder_channels_list=list(cfg.autoencoder_channels_list),
                                        groups=cfg.groups)).eval()
    m.load_state_dict(O.seeded_state_dict(O.vae_decoder_param_shapes(cfg), seed), strict=False)
    m = m.to(DEV)
    z = cases.vae_inputs(cfg, B, H, W, seed)
    with torch.no_grad(), precision("fp32"):
        img = m.decode(z.to(DEV))
    e = rel_l2(img, golden[f"vae_{name}"])
    print(f"vae fp32 {name}: {e:.3e}")
    assert e < UNET_FP32_TOL


def test_fp32_mode_is_inference_only():
    from stable_diffusion_pytorch_b200 import ops, precision
    a = torch.randn(8, 16, device=DEV, requires_grad=True)
    w = torch.randn(8, 16, device=DEV)
    with precision("fp32"), pytest.raises(NotImplementedError):
        ops.gemm(a, w)
