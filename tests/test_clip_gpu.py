"""CLIP text encoder on the CUDA path against the committed outputs of the real
``transformers.CLIPTextModel`` (tests/golden/reference_clip.npz) -- bf16 product path and fp32 parity mode."""
import os

import numpy as np
import pytest
import torch

from oracle import sd_oracle as O
from tests.golden import cases

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def gold():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_clip.npz")))


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return float((a - b).norm() / b.norm())


def build(cfg, seed):
    from stable_diffusion_pytorch_b200.models import CLIPTextEncoder, ClipTextConfig
    m = CLIPTextEncoder(ClipTextConfig(**{k: getattr(cfg, k) for k in ClipTextConfig.__dataclass_fields__})).eval()
    m.load_state_dict(O.seeded_state_dict(O.clip_text_param_shapes(cfg), seed), strict=True)   # HF key names
    return m.to(DEV)


@pytest.mark.parametrize("name", list(cases.CLIP_CASES))
def test_clip_text_encoder_vs_transformers_golden(gold, name):
    from stable_diffusion_pytorch_b200 import precision
    cfg, seed, B = cases.CLIP_CASES[name]
    m = build(cfg, seed)
    ids = cases.clip_inputs(cfg, B, seed)
    with torch.no_grad():
        y = m(ids.to(DEV))
        with precision("fp32"):
            y32 = m(ids.to(DEV))
    assert y[0].dtype == torch.float32 and tuple(y.last_hidden_state.shape) == (B, 77, cfg.hidden_size)
    e16, e32 = rel(y[0], gold[f"clip_{name}"]), rel(y32[0], gold[f"clip_{name}"])
    print(f"clip {name}: bf16 {e16:.3e}, fp32 {e32:.3e}")
    assert e16 < 2e-2 and e32 < 1e-3


def test_causal_attention_small_vs_torch():
    from stable_diffusion_pytorch_b200 import _lib, ops
    g = torch.Generator().manual_seed(2)
    B, H, n, d = 2, 3, 77, 64
    qkv = torch.randn(B, n, 3 * H * d, generator=g).to(DEV)
    out = torch.empty(B, n, H * d, device=DEV)
    C = H * d
    _lib.check(_lib.lib().sdb_attention_causal_small(
        ops._ptr(qkv[:, :, :C]), 3 * C, ops._ptr(qkv[:, :, C:2 * C]), 3 * C, ops._ptr(qkv[:, :, 2 * C:]), 3 * C,
        ops._ptr(out), C, B, H, n, d, d ** -0.5, 0, ops._stream()), "causal")
    q, k, v = (t.double().cpu().reshape(B, n, H, d).transpose(1, 2) for t in qkv.split(C, dim=-1))
    mask = torch.full((n, n), float("-inf"), dtype=torch.float64).triu(1)
    ref = (torch.softmax(q @ k.transpose(-1, -2) * d ** -0.5 + mask, -1) @ v).transpose(1, 2).reshape(B, n, C)
    assert rel(out, ref) < 2e-6
    with pytest.raises(_lib.SdbError):      # outside the kernel: reported, never a silent fallback
        _lib.check(_lib.lib().sdb_attention_causal_small(ops._ptr(qkv), 3 * C, ops._ptr(qkv), 3 * C, ops._ptr(qkv), 3 * C,
                                                         ops._ptr(out), C, B, H, 200, d, 1.0, 0, ops._stream()), "causal")


def test_latent_diffusion_with_real_text_encoder():
    """pred_noise with guidance re-encodes the empty prompt through the text encoder (latent_diffusion.py:55-60)."""
    from stable_diffusion_pytorch_b200.models import (CLIPModel, ClipTextConfig, DDPMConfig, DDPMScheduler,
                                                      LatentDiffusion, UNetModel, UnetConfig)
    ccfg = cases.CLIP_TINY
    clip = CLIPModel(ClipTextConfig(**{k: getattr(ccfg, k) for k in ClipTextConfig.__dataclass_fields__}),
                     tokenizer=lambda prompts, L: torch.full((len(prompts), L), ccfg.vocab_size - 1)).to(DEV)
    clip.text_encoder.load_state_dict(O.seeded_state_dict(O.clip_text_param_shapes(ccfg), 5))
    ucfg = O.UnetCfg(groups=32, n_heads=4, channels_list=(64, 128), time_emb_dim=128, n_layers=1, context_dim=128)
    unet = UNetModel(4, 32, UnetConfig(n_heads=4, channels_list=[64, 128], time_emb_dim=128, n_layers=1,
                                       context_dim=128, dropout=0.0)).eval()
    sd = O.seeded_state_dict(O.unet_param_shapes(ucfg), 6)
    unet.load_state_dict(sd)
    ld = LatentDiffusion(unet.to(DEV), None, clip, DDPMScheduler(DDPMConfig()))
    g = torch.Generator().manual_seed(8)
    x = torch.randn(2, 4, 16, 16, generator=g)
    t = torch.tensor([3, 500])
    ids = cases.clip_inputs(ccfg, 2, 9)
    with torch.no_grad():
        ctx = clip.encode_text(ids.to(DEV))[0]
        pn = ld.pred_noise(x.to(DEV), t.to(DEV), ctx, guidance_scale=7.5)
        # oracle: the same chain on the CPU
        csd = O.seeded_state_dict(O.clip_text_param_shapes(ccfg), 5)
        ctx_ref = O.clip_text_forward(csd, ccfg, ids)
        unc_ref = O.clip_text_forward(csd, ccfg, torch.full((2, 77), ccfg.vocab_size - 1))
        ref = O.pred_noise(lambda a, b, c: O.unet_forward(sd, ucfg, a, b, c), x, t, ctx_ref, unc_ref, 7.5)
    assert rel(ctx, ctx_ref) < 2e-2
    assert rel(pn, ref) < 5e-2
