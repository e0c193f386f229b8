"""Seeded input recipes shared by ``make_golden.py`` (which runs the reference) and the
parity tests (which run the oracle and the CUDA path).  No reference import here."""
import torch

from oracle import sd_oracle as O

TIMESTEP_CASES = [0, 1, 7, 49, 500, 999]

STEP_CASES = [0, 1, 49, 500, 999]
STEP_NOISE_SEED = 1234
STEP_SCALE = 0.9


def _g(seed):
    return torch.Generator().manual_seed(seed)


def step_inputs(step, shape=(2, 4, 16, 16)):
    g = _g(100 + step)
    return torch.randn(shape, generator=g), torch.randn(shape, generator=g)


def add_noise_inputs(shape=(5, 4, 8, 8)):
    g = _g(77)
    x0 = torch.randn(shape, generator=g)
    nz = torch.randn(shape, generator=g)
    ts = torch.tensor([0, 1, 49, 500, 999], dtype=torch.long)
    return x0, nz, ts


# name -> (cfg, seed, B, H, W)
TINY = O.UnetCfg(groups=32, n_heads=4, channels_list=(64, 128), time_emb_dim=128,
                 n_layers=1, context_dim=64)
TINY3 = O.UnetCfg(groups=16, n_heads=2, channels_list=(32, 64, 96), time_emb_dim=None,
                  n_layers=1, context_dim=96, attention_resolutions=(0, 1, 2))
DEFAULT = O.UnetCfg()

UNET_CASES = {
    "tiny_8": (TINY, 11, 2, 8, 8),
    "tiny_16": (TINY, 11, 3, 16, 16),
    "tiny3_16": (TINY3, 12, 2, 16, 16),
    "default_8": (DEFAULT, 13, 2, 8, 8),
    "default_32": (DEFAULT, 13, 2, 32, 32),
}


def unet_inputs(cfg, B, H, W, seed, ctx_len=77):
    g = _g(1000 + seed)
    x = torch.randn(B, cfg.latent_channels, H, W, generator=g)
    t = torch.randint(0, 1000, (B,), generator=g, dtype=torch.long)
    ctx = torch.randn(B, ctx_len, cfg.context_dim, generator=g)
    return x, t, ctx


# name -> (cfg, seed, B, H, W, steps, guidance)
SAMPLE_CASES = {
    "tiny_16": (TINY, 11, 2, 16, 16, 6, 7.5),
    "default_16": (DEFAULT, 13, 1, 16, 16, 4, 7.5),
    # BASELINE.json configs[0]: repo-default UNet, 1 prompt, 32x32x4 latent, 50 DDPM steps with CFG, fp32 on CPU
    "config1_32_50": (DEFAULT, 14, 1, 32, 32, 50, 7.5),
}
SAMPLE_NOISE_SEED = 4321


def sample_inputs(cfg, B, H, W, seed, ctx_len=77):
    g = _g(2000 + seed)
    x = torch.rand(B, cfg.latent_channels, H, W, generator=g)  # uniform start, scripts/txt2img.py:33
    ctx = torch.randn(B, ctx_len, cfg.context_dim, generator=g)
    uncond = torch.randn(1, ctx_len, cfg.context_dim, generator=g)
    return x, ctx, uncond


VAE_DEFAULT = O.VaeCfg()
VAE_KL_TINY = O.VaeCfg(autoencoder_channels_list=(32, 64, 64, 64), groups=16)
VAE_CASES = {
    "default_16": (VAE_DEFAULT, 21, 2, 16, 16),
    "kltiny_8": (VAE_KL_TINY, 22, 1, 8, 8),
}


def vae_inputs(cfg, B, H, W, seed):
    return torch.randn(B, cfg.latent_channels, H, W, generator=_g(3000 + seed))


# VAE encode -- name -> (cfg, seed, B, H, W) with H x W the IMAGE size
VAE_WIDE_MID = O.VaeCfg(autoencoder_channels_list=(32, 64, 256), groups=16)   # bottleneck head dim 256 (> 128)
VAE_ENC_CASES = {
    "default_32": (VAE_DEFAULT, 24, 2, 32, 32),      # [64, 128]: 16 x 16 = 256 bottleneck tokens, d = 128
    "kltiny_32": (VAE_KL_TINY, 25, 1, 32, 32),       # three DownSamples: 32 -> 4, d = 64
    "widemid_24": (VAE_WIDE_MID, 26, 2, 24, 24),     # 24 -> 6: 36 tokens (not a multiple of 8), d = 256
}
VAE_ENC_EPS_SEED = 777


def vae_encode_inputs(cfg, B, H, W, seed):
    return torch.randn(B, cfg.in_channels, H, W, generator=_g(4000 + seed))


# ---- train step (row T1) -------------------------------------------------------------------------
# name -> (cfg, seed, B, H, W, n_steps, guidance); lr raised so two steps move the weights measurably
TRAIN_CASES = {
    "tiny_8": (TINY, 11, 4, 8, 8, 2, 7.5),
    "tiny3_16_nocfg": (TINY3, 12, 2, 16, 16, 2, 1.0),   # heads 16 and 32 wide (dense-head attention backward)
}
TRAIN_LR = 1e-3
TRAIN_WD = 0.1
TRAIN_CLIP = 0.1
# parameters whose full gradient / updated value is stored (the rest is covered by per-tensor norms)
TRAIN_PROBE = ["out.2.bias", "time_embedding.0.bias", "conv_in.weight", "out.0.weight"]


def train_batches(cfg, B, H, W, seed, n_steps, ctx_len=77):
    """[(x_t, t, ctx, noise)] * n_steps and the empty-prompt embedding [1, L, D]."""
    g = _g(5000 + seed)
    uncond = torch.randn(1, ctx_len, cfg.context_dim, generator=g)
    out = []
    for _ in range(n_steps):
        x_t = torch.randn(B, cfg.latent_channels, H, W, generator=g)
        t = torch.randint(0, 1000, (B,), generator=g, dtype=torch.long)
        ctx = torch.randn(B, ctx_len, cfg.context_dim, generator=g)
        noise = torch.randn(B, cfg.latent_channels, H, W, generator=g)
        out.append((x_t, t, ctx, noise))
    return out, uncond


# ---- CLIP text encoder -- name -> (cfg, seed, B)
CLIP_TINY = O.ClipTextCfg(vocab_size=1000, hidden_size=128, intermediate_size=256, num_hidden_layers=2,
                          num_attention_heads=2)
CLIP_SD15 = O.ClipTextCfg()     # the text tower of runwayml/stable-diffusion-v1-5 (clip_model.py:11-18)
CLIP_CASES = {
    "tiny": (CLIP_TINY, 21, 3),
    "sd15": (CLIP_SD15, 22, 1),
}


def clip_inputs(cfg, B, seed):
    """Token ids shaped like the tokenizer's output (clip_model.py:72-85): BOS, random tokens, EOS padding."""
    g = _g(3000 + seed)
    L = cfg.max_position_embeddings
    ids = torch.randint(0, cfg.vocab_size - 2, (B, L), generator=g, dtype=torch.long)
    ids[:, 0] = cfg.vocab_size - 2
    for b in range(B):
        ids[b, 5 + 11 * b:] = cfg.vocab_size - 1
    return ids
