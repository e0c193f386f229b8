"""CPU: host-side logic of the drop-in modules -- state_dict compatibility, scheduler tables,
timestep order, error behaviour.  No kernel is launched."""
import os
import sys

import pytest
import torch

from oracle import sd_oracle as O
from stable_diffusion_pytorch_b200.models import (AutoEncoderKL, AutoencoderConfig, DDPMConfig, DDPMScheduler,
                                                  UNetModel, UnetConfig)
from tests.golden import cases


def _ucfg(c: O.UnetCfg) -> UnetConfig:
    return UnetConfig(num_res_blocks=c.num_res_blocks, n_heads=c.n_heads,
                      attention_resolutions=list(c.attention_resolutions), channels_list=list(c.channels_list),
                      time_emb_dim=c.time_emb_dim, dropout=0.0, n_layers=c.n_layers, context_dim=c.context_dim)


@pytest.mark.parametrize("cfg", [cases.DEFAULT, cases.TINY, cases.TINY3])
def test_unet_state_dict_keys_match_reference_layout(cfg):
    m = UNetModel(cfg.latent_channels, cfg.groups, _ucfg(cfg))
    shapes = O.unet_param_shapes(cfg)   # pinned to the reference by make_golden.py (strict load)
    sd = m.state_dict()
    assert set(sd) == set(shapes)
    assert all(tuple(sd[k].shape) == shapes[k] for k in sd)
    m.load_state_dict(O.seeded_state_dict(shapes, 3), strict=True)


def test_reference_zero_init_is_kept():
    m = UNetModel(4, 32, UnetConfig())
    assert float(m.input_blocks[0][0].out_layers[3].weight.abs().sum()) == 0.0   # resnet2d.py:194-198
    assert float(m.input_blocks[0][1].proj_out.weight.abs().sum()) == 0.0        # transformer.py:346-348
    assert m.middle_block[0].groups == 2 and m.middle_block[2].groups == 32      # models/utils.py:112
    assert m.middle_block[1].transformer_blocks[0].self_attn.d_head == 20        # d_head leak, unet.py:198


@pytest.mark.parametrize("cfg", [cases.VAE_DEFAULT, cases.VAE_KL_TINY])
def test_vae_state_dict(cfg):
    m = AutoEncoderKL(AutoencoderConfig(autoencoder_channels_list=list(cfg.autoencoder_channels_list),
                                        groups=cfg.groups))
    shapes = O.vae_decoder_param_shapes(cfg)
    sd = m.state_dict()
    assert all(k in sd and tuple(sd[k].shape) == shapes[k] for k in shapes)
    assert all(k.startswith(("encoder.", "decoder.", "quant_conv.", "post_quant_conv.")) for k in sd)


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="reference tree only exists in the build container")
def test_state_dict_keys_equal_the_live_reference():
    sys.path.insert(0, "/root/reference")
    from stable_diffusion.models.unet import UNetModel as RefUNet, UnetConfig as RefCfg
    from stable_diffusion.models.autoencoder import AutoEncoderKL as RefVae, AutoencoderConfig as RefVaeCfg
    ref, mine = RefUNet(4, 32, RefCfg()).state_dict(), UNetModel(4, 32, UnetConfig()).state_dict()
    assert list(ref) == list(mine)          # same keys in the same order
    assert all(ref[k].shape == mine[k].shape for k in ref)
    rv, mv = RefVae(RefVaeCfg()).state_dict(), AutoEncoderKL(AutoencoderConfig()).state_dict()
    assert list(rv) == list(mv) and all(rv[k].shape == mv[k].shape for k in rv)


def test_scheduler_tables_bit_exact(golden):
    s = DDPMScheduler(DDPMConfig())
    for name in O.TABLE_NAMES:
        assert (getattr(s, name).numpy() == golden[f"table_{name}"]).all(), name
    assert (s.noise_time_steps.numpy() == golden["table_noise_time_steps"]).all()
    rows = s.coef_rows([0, 49])
    assert rows[0, 4] == 0.0 and rows[1, 4] == (0.5 * s.log_var[49]).exp()
    assert rows[0, 5] == 1.0 and rows[0, 6] == 0.0


def test_errors_mirror_reference_asserts():
    m = UNetModel(4, 32, UnetConfig())
    with pytest.raises(AssertionError, match="context dim"):
        m(torch.zeros(1, 4, 8, 8), torch.zeros(1, dtype=torch.long), torch.zeros(1, 77, 64))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.zeros(1, 4, 8, 8), torch.zeros(1, dtype=torch.long), torch.zeros(1, 77, 768))
    s = DDPMScheduler(DDPMConfig())
    with pytest.raises(AssertionError, match="should be less than"):
        s.step(torch.zeros(1, 4, 8, 8), torch.zeros(1, 4, 8, 8), 1000)
    v = AutoEncoderKL(AutoencoderConfig())
    with pytest.raises(AssertionError, match="Expected latent"):
        v.decode(torch.zeros(1, 3, 8, 8))


def test_precision_context_and_pack_cache_follow_compute_dtype():
    """precision("fp32") (the parity mode) is scoped, validated, and re-packs cached operands."""
    import torch
    import stable_diffusion_pytorch_b200 as sdb
    from stable_diffusion_pytorch_b200 import _pack
    assert sdb.compute_dtype() == torch.bfloat16
    w = torch.nn.Parameter(torch.randn(8, 4))
    cache = _pack.PackCache()
    a = cache.get("w", [w], lambda: _pack.linear_w(w))
    assert a.dtype == torch.bfloat16 and cache.get("w", [w], lambda: _pack.linear_w(w)) is a
    with sdb.precision("fp32"):
        assert sdb.compute_dtype() == torch.float32
        b = cache.get("w", [w], lambda: _pack.linear_w(w))
        assert b.dtype == torch.float32 and torch.equal(b, w.detach())
        with sdb.precision(torch.bfloat16):
            assert sdb.compute_dtype() == torch.bfloat16
        assert sdb.compute_dtype() == torch.float32
    assert sdb.compute_dtype() == torch.bfloat16
    assert cache.get("w", [w], lambda: _pack.linear_w(w)).dtype == torch.bfloat16
    import pytest
    with pytest.raises(ValueError):
        sdb.set_precision("fp16")


def test_clip_text_encoder_keys_match_transformers_names():
    """CLIPTextEncoder's parameter tree carries the transformers.CLIPTextModel key names (oracle enumeration)."""
    from oracle import sd_oracle as O
    from stable_diffusion_pytorch_b200.models import CLIPModel, CLIPTextEncoder, ClipTextConfig
    cfg = O.ClipTextCfg(vocab_size=100, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=2)
    m = CLIPTextEncoder(ClipTextConfig(vocab_size=100, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                                       num_attention_heads=2))
    want = O.clip_text_param_shapes(cfg)
    got = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert got == want
    sd = O.seeded_state_dict(want, 1)
    sd["text_model.embeddings.position_ids"] = torch.arange(77)[None]     # buffer of older transformers releases
    m.load_state_dict(sd, strict=True)
    import pytest
    with pytest.raises(RuntimeError):
        CLIPModel(ClipTextConfig(vocab_size=100, hidden_size=64, intermediate_size=128, num_hidden_layers=1,
                                 num_attention_heads=2)).tokenize("a prompt")        # no tokenizer given
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 77, dtype=torch.long))                                        # CPU module: no fallback


def test_cosine_and_cubic_noise_schedules_match_the_oracle():
    """the schedules DDPMConfig advertises (scheduler.py:11-17; the reference itself only builds linear)."""
    from oracle import sd_oracle as O
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler
    for kind in ("linear", "cosine", "cubic"):
        s = DDPMScheduler(DDPMConfig(noise_schedule=kind))
        tab = O.ddpm_tables(O.DdpmCfg(noise_schedule=kind))
        assert torch.allclose(s.betas, tab["betas"], rtol=1e-5, atol=1e-9), kind
        assert torch.allclose(s.alphas_cumprod, tab["alphas_cumprod"], rtol=1e-4, atol=1e-12), kind
        assert float(s.betas.min()) > 0 and float(s.betas.max()) <= 0.9990001 and torch.all(s.alphas_cumprod[1:] < s.alphas_cumprod[:-1])
        rows = s.coef_rows(range(s.noise_steps))
        assert torch.isfinite(rows).all() and float(rows[0, 4]) == 0.0      # no noise at step 0
    lin = DDPMScheduler(DDPMConfig())
    assert torch.equal(lin.betas, torch.linspace(1e-4, 0.02, 1000))         # the default stays the reference's table
    import pytest
    with pytest.raises(ValueError):
        DDPMScheduler(DDPMConfig(noise_schedule="sigmoid"))
