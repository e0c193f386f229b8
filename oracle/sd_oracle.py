"""CPU restatement (plain torch fp32, functional, state_dict-driven) of the
reference's latent-diffusion hot path.  TEST INFRASTRUCTURE -- see
``oracle/__init__.py``; never imported by the product package.

Every function cites the reference file:line (relative to the reference repo
root) whose arithmetic it follows.  Quirks are reproduced on purpose:
ascending timesteps, mislabelled CFG halves, positive-exponent sinusoid with a
CPU-built ``freq`` table, post-norm transformer, ``d_head`` leaking into the
bottleneck, ``groups=2`` in the first bottleneck ResBlock.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
StateDict = Dict[str, Tensor]


# ----------------------------------------------------------------------------
# configs (field names follow stable_diffusion/models/unet.py:29-57,
# autoencoder.py:22-48, scheduler.py:9-22)
# ----------------------------------------------------------------------------
@dataclass
class UnetCfg:
    latent_channels: int = 4
    groups: int = 32
    num_res_blocks: int = 2
    n_heads: int = 8
    attention_resolutions: Sequence[int] = (0, 1)
    channels_list: Sequence[int] = (160, 320)
    time_emb_dim: Optional[int] = 512
    n_layers: int = 2
    context_dim: int = 768


@dataclass
class VaeCfg:
    in_channels: int = 3
    latent_channels: int = 4
    out_channels: int = 3
    autoencoder_channels_list: Sequence[int] = (64, 128)
    autoencoder_num_res_blocks: int = 2
    groups: int = 32


@dataclass
class DdpmCfg:
    noise_steps: int = 1000
    beta_start: float = 1e-4
    beta_end: float = 0.02
    noise_schedule: str = "linear"      # the reference builds "linear" whatever this says (scheduler.py:89-93)


# ----------------------------------------------------------------------------
# topology: restates stable_diffusion/models/utils.py:35-190
# A block is a list of layers; a layer is ("res", prefix, cin, cout, groups,
# has_temb) | ("st", prefix, ch, n_heads, d_head, n_layers, groups) |
# ("attn", prefix, ch) | ("down", prefix, ch) | ("up", prefix, ch)
# ----------------------------------------------------------------------------
def _input_blocks(prefix, in_ch, num_res_blocks, attn_res, n_heads, n_layers,
                  levels, channels_list, groups, has_temb):
    """models/utils.py:35-97."""
    blocks, chans = [], [in_ch]
    attn_mult, d_head, idx = 1, None, 0
    for level in range(levels):
        for _ in range(num_res_blocks):
            out_ch = channels_list[level]
            layers = [("res", f"{prefix}.{idx}.0", in_ch, out_ch, groups, has_temb)]
            in_ch = out_ch
            if attn_res is not None and attn_mult in attn_res:
                d_head = in_ch // n_heads
                layers.append(("st", f"{prefix}.{idx}.1", in_ch, n_heads, d_head, n_layers, groups))
            blocks.append(layers)
            chans.append(in_ch)
            idx += 1
        if level != levels - 1:
            blocks.append([("down", f"{prefix}.{idx}.0", in_ch)])
            chans.append(in_ch)
            attn_mult *= 2
            idx += 1
    return blocks, chans, in_ch, d_head, attn_mult


def _output_blocks(prefix, num_res_blocks, attn_res, n_heads, n_layers, skip_chans,
                   levels, channels_list, in_ch, attn_mult, groups, has_temb):
    """models/utils.py:137-190."""
    skip_chans = list(skip_chans) if skip_chans else []
    blocks, idx = [], 0
    for level in reversed(range(levels)):
        for rb in range(num_res_blocks + 1):
            out_ch = channels_list[level]
            cin = in_ch + skip_chans.pop() if skip_chans else in_ch
            layers = [("res", f"{prefix}.{idx}.0", cin, out_ch, groups, has_temb)]
            in_ch = out_ch
            if attn_res is not None and attn_mult in attn_res:
                layers.append(("st", f"{prefix}.{idx}.{len(layers)}", in_ch, n_heads,
                               in_ch // n_heads, n_layers, groups))
            if level != 0 and rb == num_res_blocks:
                layers.append(("up", f"{prefix}.{idx}.{len(layers)}.0", in_ch))
                if attn_mult:
                    attn_mult //= 2
            blocks.append(layers)
            idx += 1
    return blocks, in_ch


def unet_topology(cfg: UnetCfg):
    """models/unet.py:164-246."""
    ch0 = cfg.channels_list[0]
    levels = len(cfg.channels_list)
    inp, chans, mid_ch, d_head, attn_mult = _input_blocks(
        "input_blocks", ch0, cfg.num_res_blocks, cfg.attention_resolutions, cfg.n_heads,
        cfg.n_layers, levels, cfg.channels_list, cfg.groups, True)
    # models/utils.py:100-134: first ResBlock omits `groups` -> default 2 (resnet2d.py:159)
    mid = [("res", "middle_block.0", mid_ch, mid_ch, 2, True),
           ("st", "middle_block.1", mid_ch, cfg.n_heads, d_head, cfg.n_layers, cfg.groups),
           ("res", "middle_block.2", mid_ch, mid_ch, cfg.groups, True)]
    out, out_ch = _output_blocks(
        "output_blocks", cfg.num_res_blocks, cfg.attention_resolutions, cfg.n_heads,
        cfg.n_layers, chans, levels, cfg.channels_list, mid_ch, attn_mult, cfg.groups, True)
    return inp, mid, out, out_ch


def vae_decoder_topology(cfg: VaeCfg):
    """models/autoencoder.py:202-232."""
    cl = cfg.autoencoder_channels_list
    c0 = cl[0]
    mid = [("res", "decoder.bottleneck.0", c0, c0, 2, False),
           ("attn", "decoder.bottleneck.1", c0),
           ("res", "decoder.bottleneck.2", c0, c0, cfg.groups, False)]
    up, out_ch = _output_blocks("decoder.up", cfg.autoencoder_num_res_blocks, None, 1, 1,
                                None, len(cl), cl, c0, None, cfg.groups, False)
    return mid, up, out_ch


def vae_encoder_topology(cfg: VaeCfg):
    """models/autoencoder.py:164-199: conv_in -> build_input_blocks (ResBlocks without time embedding, a DownSample
    after every level but the last, no attention: attention_resolutions is None) -> bottleneck [ResBlock(groups
    default 2), bare 1-head CrossAttention with d_head = mid_ch, ResBlock(groups)] -> GN + SiLU + conv head."""
    cl = cfg.autoencoder_channels_list
    down, _, mid_ch, _, _ = _input_blocks("encoder.down", cl[0], cfg.autoencoder_num_res_blocks, None, 1, 1,
                                          len(cl), cl, cfg.groups, False)
    mid = [("res", "encoder.bottleneck.0", mid_ch, mid_ch, 2, False),
           ("attn", "encoder.bottleneck.1", mid_ch),
           ("res", "encoder.bottleneck.2", mid_ch, mid_ch, cfg.groups, False)]
    return down, mid, mid_ch


# ----------------------------------------------------------------------------
# parameter shapes + seeded init (our own recipe; every tensor non-zero so that
# zero_module'd convs -- models/utils.py:22-28 -- do not make parity vacuous)
# ----------------------------------------------------------------------------
def _res_shapes(p, cin, cout, temb_dim):
    s = {f"{p}.in_layers.0.weight": (cin,), f"{p}.in_layers.0.bias": (cin,),
         f"{p}.in_layers.2.weight": (cout, cin, 3, 3), f"{p}.in_layers.2.bias": (cout,)}
    if temb_dim:
        s[f"{p}.time_embedding.1.weight"] = (cout, temb_dim)
        s[f"{p}.time_embedding.1.bias"] = (cout,)
    s.update({f"{p}.out_layers.0.weight": (cout,), f"{p}.out_layers.0.bias": (cout,),
              f"{p}.out_layers.3.weight": (cout, cout, 3, 3), f"{p}.out_layers.3.bias": (cout,)})
    if cin != cout:
        s[f"{p}.skip_connection.weight"] = (cout, cin, 1, 1)
        s[f"{p}.skip_connection.bias"] = (cout,)
    return s


def _attn_shapes(p, qdim, cdim, dmodel):
    return {f"{p}.to_q.weight": (dmodel, qdim), f"{p}.to_k.weight": (dmodel, cdim),
            f"{p}.to_v.weight": (dmodel, cdim), f"{p}.out.0.weight": (qdim, dmodel),
            f"{p}.out.0.bias": (qdim,)}


def _st_shapes(p, ch, n_heads, d_head, n_layers, ctx_dim):
    s = {f"{p}.norm.weight": (ch,), f"{p}.norm.bias": (ch,),
         f"{p}.proj_in.weight": (ch, ch, 1, 1), f"{p}.proj_in.bias": (ch,)}
    dm = n_heads * d_head
    for i in range(n_layers):
        b = f"{p}.transformer_blocks.{i}"
        s.update(_attn_shapes(f"{b}.self_attn", ch, ch, dm))
        s.update({f"{b}.norm1.weight": (ch,), f"{b}.norm1.bias": (ch,)})
        s.update(_attn_shapes(f"{b}.cross_attn", ch, ctx_dim, dm))
        s.update({f"{b}.norm2.weight": (ch,), f"{b}.norm2.bias": (ch,)})
        s.update({f"{b}.ffn.net.0.proj.weight": (8 * ch, ch), f"{b}.ffn.net.0.proj.bias": (8 * ch,),
                  f"{b}.ffn.net.2.weight": (ch, 4 * ch), f"{b}.ffn.net.2.bias": (ch,)})
        s.update({f"{b}.norm3.weight": (ch,), f"{b}.norm3.bias": (ch,)})
    s.update({f"{p}.proj_out.weight": (ch, ch, 1, 1), f"{p}.proj_out.bias": (ch,)})
    return s


def _layers_shapes(layers, temb_dim, ctx_dim):
    s = {}
    for L in layers:
        if L[0] == "res":
            s.update(_res_shapes(L[1], L[2], L[3], temb_dim if L[5] else None))
        elif L[0] == "st":
            s.update(_st_shapes(L[1], L[2], L[3], L[4], L[5], ctx_dim))
        elif L[0] == "attn":
            s.update(_attn_shapes(L[1], L[2], L[2], L[2]))
        else:  # down / up
            s.update({f"{L[1]}.conv.weight": (L[2], L[2], 3, 3), f"{L[1]}.conv.bias": (L[2],)})
    return s


def unet_param_shapes(cfg: UnetCfg) -> Dict[str, Tuple[int, ...]]:
    ch0 = cfg.channels_list[0]
    td = cfg.time_emb_dim or 4 * ch0
    s = {"time_embedding.0.weight": (td, ch0), "time_embedding.0.bias": (td,),
         "time_embedding.2.weight": (td, td), "time_embedding.2.bias": (td,),
         "conv_in.weight": (ch0, cfg.latent_channels, 3, 3), "conv_in.bias": (ch0,)}
    inp, mid, out, out_ch = unet_topology(cfg)
    for blk in inp:
        s.update(_layers_shapes(blk, td, cfg.context_dim))
    s.update(_layers_shapes(mid, td, cfg.context_dim))
    for blk in out:
        s.update(_layers_shapes(blk, td, cfg.context_dim))
    s.update({"out.0.weight": (out_ch,), "out.0.bias": (out_ch,),
              "out.2.weight": (cfg.latent_channels, out_ch, 3, 3), "out.2.bias": (cfg.latent_channels,)})
    return s


def vae_decoder_param_shapes(cfg: VaeCfg) -> Dict[str, Tuple[int, ...]]:
    c0 = cfg.autoencoder_channels_list[0]
    lc = cfg.latent_channels
    s = {"post_quant_conv.weight": (lc, lc, 1, 1), "post_quant_conv.bias": (lc,),
         "decoder.conv_in.weight": (c0, lc, 3, 3), "decoder.conv_in.bias": (c0,)}
    mid, up, out_ch = vae_decoder_topology(cfg)
    s.update(_layers_shapes(mid, None, None))
    for blk in up:
        s.update(_layers_shapes(blk, None, None))
    s.update({"decoder.out.0.weight": (out_ch,), "decoder.out.0.bias": (out_ch,),
              "decoder.out.2.weight": (cfg.out_channels, out_ch, 3, 3),
              "decoder.out.2.bias": (cfg.out_channels,)})
    return s


def vae_encoder_param_shapes(cfg: VaeCfg) -> Dict[str, Tuple[int, ...]]:
    """Keys of ``AutoEncoderKL.encoder`` and ``quant_conv`` (models/autoencoder.py:95-121,164-199)."""
    c0 = cfg.autoencoder_channels_list[0]
    lc = cfg.latent_channels
    s = {"quant_conv.weight": (2 * lc, 2 * lc, 1, 1), "quant_conv.bias": (2 * lc,),
         "encoder.conv_in.weight": (c0, cfg.in_channels, 3, 3), "encoder.conv_in.bias": (c0,)}
    down, mid, mid_ch = vae_encoder_topology(cfg)
    for blk in down:
        s.update(_layers_shapes(blk, None, None))
    s.update(_layers_shapes(mid, None, None))
    s.update({"encoder.out.0.weight": (mid_ch,), "encoder.out.0.bias": (mid_ch,),
              "encoder.out.2.weight": (2 * lc, mid_ch, 3, 3), "encoder.out.2.bias": (2 * lc,)})
    return s


def seeded_state_dict(shapes: Dict[str, Tuple[int, ...]], seed: int) -> StateDict:
    """Deterministic non-degenerate init from a CPU ``torch.Generator``.

    1-D ``*.weight`` (GroupNorm/LayerNorm gains) ~ 1 + 0.1 N(0,1); biases ~ 0.05 N(0,1);
    matrices/filters ~ N(0,1)/sqrt(fan_in).  Keys are visited in sorted order so
    the recipe does not depend on dict construction order.
    """
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k in sorted(shapes):
        shp = shapes[k]
        r = torch.randn(shp, generator=g, dtype=torch.float32)
        if k.endswith(".bias"):
            sd[k] = 0.05 * r
        elif len(shp) == 1:
            sd[k] = 1.0 + 0.1 * r
        else:
            fan_in = 1
            for d in shp[1:]:
                fan_in *= d
            sd[k] = r / math.sqrt(fan_in)
    return sd


# ----------------------------------------------------------------------------
# modules (functional)
# ----------------------------------------------------------------------------
def sinusoidal_freq(emb_dim: int, max_len: int = 10000) -> Tensor:
    """modules/timestep_embedding.py:93-99: POSITIVE exponent, built on the CPU in fp32."""
    half = emb_dim // 2
    return torch.exp(math.log(max_len) / half * torch.arange(0, end=half, dtype=torch.float32))


def sinusoidal_time_proj(time_steps: Tensor, emb_dim: int, max_len: int = 10000) -> Tensor:
    """modules/timestep_embedding.py:73-105: cat[sin, cos] of t.float() * freq."""
    freq = sinusoidal_freq(emb_dim, max_len).to(time_steps.device)      # :99 (CPU-built table, then moved)
    args = time_steps[:, None].float() * freq[None]
    return torch.cat([torch.sin(args), torch.cos(args)], dim=-1)


def res_block(sd: StateDict, p: str, x: Tensor, temb: Optional[Tensor], groups: int) -> Tensor:
    """modules/resnet2d.py:208-238 (eval mode: Dropout is the identity)."""
    h = F.group_norm(x, groups, sd[f"{p}.in_layers.0.weight"], sd[f"{p}.in_layers.0.bias"], eps=1e-5)
    h = F.conv2d(F.silu(h), sd[f"{p}.in_layers.2.weight"], sd[f"{p}.in_layers.2.bias"], padding=1)
    if temb is not None and f"{p}.time_embedding.1.weight" in sd:
        t = F.linear(F.silu(temb), sd[f"{p}.time_embedding.1.weight"], sd[f"{p}.time_embedding.1.bias"])
        h = h + t[:, :, None, None]
    h = F.group_norm(h, groups, sd[f"{p}.out_layers.0.weight"], sd[f"{p}.out_layers.0.bias"], eps=1e-5)
    h = F.conv2d(F.silu(h), sd[f"{p}.out_layers.3.weight"], sd[f"{p}.out_layers.3.bias"], padding=1)
    if f"{p}.skip_connection.weight" in sd:
        x = F.conv2d(x, sd[f"{p}.skip_connection.weight"], sd[f"{p}.skip_connection.bias"])
    return h + x


def attention(sd: StateDict, p: str, q_in: Tensor, ctx: Optional[Tensor], n_heads: int) -> Tensor:
    """modules/transformer.py:60-119 on [B, n, C] tokens (mask path is dead code there)."""
    if ctx is None:
        ctx = q_in
    Q = F.linear(q_in, sd[f"{p}.to_q.weight"])
    K = F.linear(ctx, sd[f"{p}.to_k.weight"])
    V = F.linear(ctx, sd[f"{p}.to_v.weight"])
    B, n, dm = Q.shape
    m = K.shape[1]
    d = dm // n_heads
    scale = 1 / (d ** 0.5)
    Q = Q.view(B, n, n_heads, d).permute(0, 2, 1, 3).reshape(B * n_heads, n, d)
    K = K.view(B, m, n_heads, d).permute(0, 2, 1, 3).reshape(B * n_heads, m, d)
    V = V.view(B, m, n_heads, d).permute(0, 2, 1, 3).reshape(B * n_heads, m, d)
    sim = torch.einsum("bnd,bmd->bnm", Q, K) * scale
    attn = sim.softmax(dim=-1)
    o = torch.einsum("bnm,bmd->bnd", attn, V)
    o = o.view(B, n_heads, n, d).permute(0, 2, 1, 3).reshape(B, n, dm)
    return F.linear(o, sd[f"{p}.out.0.weight"], sd[f"{p}.out.0.bias"])


def transformer_block(sd: StateDict, p: str, x: Tensor, ctx: Optional[Tensor], n_heads: int) -> Tensor:
    """modules/transformer.py:250-279: POST-norm residual blocks; GEGLU FFN :179-192."""
    C = x.shape[-1]
    x = F.layer_norm(x + attention(sd, f"{p}.self_attn", x, None, n_heads), (C,),
                     sd[f"{p}.norm1.weight"], sd[f"{p}.norm1.bias"], 1e-5)
    x = F.layer_norm(x + attention(sd, f"{p}.cross_attn", x, ctx, n_heads), (C,),
                     sd[f"{p}.norm2.weight"], sd[f"{p}.norm2.bias"], 1e-5)
    a, gate = F.linear(x, sd[f"{p}.ffn.net.0.proj.weight"], sd[f"{p}.ffn.net.0.proj.bias"]).chunk(2, dim=-1)
    ff = F.linear(a * F.gelu(gate), sd[f"{p}.ffn.net.2.weight"], sd[f"{p}.ffn.net.2.bias"])
    return F.layer_norm(x + ff, (C,), sd[f"{p}.norm3.weight"], sd[f"{p}.norm3.bias"], 1e-5)


def spatial_transformer(sd: StateDict, p: str, x: Tensor, ctx: Optional[Tensor], n_heads: int,
                        n_layers: int, groups: int) -> Tensor:
    """modules/transformer.py:350-383: GN (no SiLU) -> 1x1 -> tokens -> blocks -> 1x1 -> +x_in."""
    B, C, H, W = x.shape
    h = F.group_norm(x, groups, sd[f"{p}.norm.weight"], sd[f"{p}.norm.bias"], eps=1e-5)
    h = F.conv2d(h, sd[f"{p}.proj_in.weight"], sd[f"{p}.proj_in.bias"])
    h = h.permute(0, 2, 3, 1).reshape(B, H * W, C)
    for i in range(n_layers):
        h = transformer_block(sd, f"{p}.transformer_blocks.{i}", h, ctx, n_heads)
    h = h.reshape(B, H, W, C).permute(0, 3, 1, 2)
    h = F.conv2d(h, sd[f"{p}.proj_out.weight"], sd[f"{p}.proj_out.bias"])
    return h + x


def _run_layers(sd, layers, x, temb, ctx):
    """modules/timestep_embedding.py:60-70 routing."""
    for L in layers:
        kind, p = L[0], L[1]
        if kind == "res":
            x = res_block(sd, p, x, temb, L[4])
        elif kind == "st":
            x = spatial_transformer(sd, p, x, ctx, L[3], L[5], L[6])
        elif kind == "attn":
            # VAE bottleneck: bare CrossAttention on a 4-D map, 1 head, NO residual
            # (modules/transformer.py:85-88,117-118; models/utils.py:113-117)
            B, C, H, W = x.shape
            t = x.permute(0, 2, 3, 1).reshape(B, H * W, C)
            t = attention(sd, p, t, None, 1)
            x = t.reshape(B, H, W, C).permute(0, 3, 1, 2)
        elif kind == "down":
            # modules/resnet2d.py:108-124: conv THEN nearest x0.5
            x = F.conv2d(x, sd[f"{p}.conv.weight"], sd[f"{p}.conv.bias"], padding=1)
            x = F.interpolate(x, scale_factor=0.5, mode="nearest")
        elif kind == "up":
            # modules/resnet2d.py:46-63: nearest x2 THEN conv
            x = F.interpolate(x, scale_factor=2, mode="nearest")
            x = F.conv2d(x, sd[f"{p}.conv.weight"], sd[f"{p}.conv.bias"], padding=1)
    return x


def unet_forward(sd: StateDict, cfg: UnetCfg, x: Tensor, timesteps: Tensor,
                 ctx: Optional[Tensor]) -> Tensor:
    """models/unet.py:264-310."""
    if ctx is not None:
        assert ctx.shape[-1] == cfg.context_dim
    inp, mid, out, _ = unet_topology(cfg)
    temb = sinusoidal_time_proj(timesteps, cfg.channels_list[0]).to(x.dtype)
    temb = F.linear(temb, sd["time_embedding.0.weight"], sd["time_embedding.0.bias"])
    temb = F.linear(F.silu(temb), sd["time_embedding.2.weight"], sd["time_embedding.2.bias"])
    x = F.conv2d(x, sd["conv_in.weight"], sd["conv_in.bias"], padding=1)
    skips = [x]
    for blk in inp:
        x = _run_layers(sd, blk, x, temb, ctx)
        skips.append(x)
    x = _run_layers(sd, mid, x, temb, ctx)
    for blk in out:
        x = torch.cat([x, skips.pop()], dim=1)
        x = _run_layers(sd, blk, x, temb, ctx)
    x = F.group_norm(x, cfg.groups, sd["out.0.weight"], sd["out.0.bias"], eps=1e-5)
    return F.conv2d(F.silu(x), sd["out.2.weight"], sd["out.2.bias"], padding=1)


def vae_decode(sd: StateDict, cfg: VaeCfg, latent: Tensor) -> Tensor:
    """models/autoencoder.py:148-161 + Decoder.forward :234-241."""
    assert latent.shape[1] == cfg.latent_channels
    mid, up, _ = vae_decoder_topology(cfg)
    z = F.conv2d(latent, sd["post_quant_conv.weight"], sd["post_quant_conv.bias"])
    x = F.conv2d(z, sd["decoder.conv_in.weight"], sd["decoder.conv_in.bias"], padding=1)
    x = _run_layers(sd, mid, x, None, None)
    for blk in up:
        x = _run_layers(sd, blk, x, None, None)
    x = F.group_norm(x, cfg.groups, sd["decoder.out.0.weight"], sd["decoder.out.0.bias"], eps=1e-5)
    return F.conv2d(F.silu(x), sd["decoder.out.2.weight"], sd["decoder.out.2.bias"], padding=1)


def vae_encode(sd: StateDict, cfg: VaeCfg, img: Tensor) -> Tensor:
    """models/autoencoder.py:132-146 + Encoder.forward :191-199 -> the MOMENTS [B, 2 * latent_channels, h, w] that
    ``GaussianDistribution`` (modules/distributions.py:5-19) splits into mean | log_var."""
    down, mid, _ = vae_encoder_topology(cfg)
    x = F.conv2d(img, sd["encoder.conv_in.weight"], sd["encoder.conv_in.bias"], padding=1)
    for blk in down:
        x = _run_layers(sd, blk, x, None, None)
    x = _run_layers(sd, mid, x, None, None)
    x = F.group_norm(x, cfg.groups, sd["encoder.out.0.weight"], sd["encoder.out.0.bias"], eps=1e-5)
    z = F.conv2d(F.silu(x), sd["encoder.out.2.weight"], sd["encoder.out.2.bias"], padding=1)
    return F.conv2d(z, sd["quant_conv.weight"], sd["quant_conv.bias"])


def gaussian_moments(moments: Tensor) -> Tuple[Tensor, Tensor]:
    """modules/distributions.py:8: mean, log_var = chunk(moments, 2, dim=1)."""
    mean, log_var = torch.chunk(moments, 2, dim=1)
    return mean, log_var


def gaussian_sample(moments: Tensor, eps: Tensor) -> Tensor:
    """modules/distributions.py:10-13 with the ``randn_like`` draw passed in: mean + eps * exp(0.5 * log_var)."""
    mean, log_var = gaussian_moments(moments)
    return mean + eps * torch.exp(0.5 * log_var)


def gaussian_kl(moments: Tensor) -> Tensor:
    """modules/distributions.py:15-19: 0.5 * sum(mean^2 + var - 1 - log_var) over dims [1, 2, 3]."""
    mean, log_var = gaussian_moments(moments)
    return 0.5 * torch.sum(torch.pow(mean, 2) + torch.exp(log_var) - 1.0 - log_var, dim=[1, 2, 3])


# ----------------------------------------------------------------------------
# scheduler / pipeline
# ----------------------------------------------------------------------------
TABLE_NAMES = ("betas", "alpha", "alphas_cumprod", "sqrt_alpha_bar", "sqrt_1m_alpha_bar",
               "sqrt_recip_alpha_bar", "sqrt_recip_m1_alpha_bar", "log_var", "mean_x0_coef",
               "mean_xt_coef")


def noise_schedule_betas(cfg: DdpmCfg) -> Tensor:
    """"linear": models/scheduler.py:89-93.  "cosine" / "cubic" are advertised by DDPMConfig (:11-17) but never built
    by the reference -- restated here from their published definitions (Nichol & Dhariwal 2021, eq. 17; the cube-root
    analogue of the scaled-linear schedule), "parity unpinned"."""
    T = cfg.noise_steps
    if cfg.noise_schedule == "linear":
        return torch.linspace(cfg.beta_start, cfg.beta_end, T)
    if cfg.noise_schedule == "cosine":
        f = [math.cos(((i / T) + 0.008) / 1.008 * math.pi / 2) ** 2 for i in range(T + 1)]
        return torch.tensor([min(1 - f[i + 1] / f[i], 0.999) for i in range(T)], dtype=torch.float32)
    if cfg.noise_schedule == "cubic":
        return torch.linspace(cfg.beta_start ** (1 / 3), cfg.beta_end ** (1 / 3), T) ** 3
    raise ValueError(cfg.noise_schedule)


def ddpm_tables(cfg: DdpmCfg) -> Dict[str, Tensor]:
    """models/scheduler.py:52-93: fp32 CPU tables built with the same torch ops, same order."""
    T = cfg.noise_steps
    betas = noise_schedule_betas(cfg)
    alpha = 1.0 - betas
    ac = torch.cumprod(alpha, dim=0)
    ac_prev = torch.cat([ac.new_tensor([1.0]), ac[:-1]])
    variance = betas * (1.0 - ac_prev) / (1.0 - ac)
    return {
        "noise_time_steps": torch.arange(T - 1, -1, -1),
        "betas": betas, "alpha": alpha, "alphas_cumprod": ac,
        "sqrt_alpha_bar": ac ** 0.5,
        "sqrt_1m_alpha_bar": (1.0 - ac) ** 0.5,
        "sqrt_recip_alpha_bar": ac ** -0.5,
        "sqrt_recip_m1_alpha_bar": (1 / ac - 1) ** 0.5,
        "log_var": torch.log(torch.clamp(variance, min=1e-20)),
        "mean_x0_coef": betas * (ac_prev ** 0.5) / (1.0 - ac),
        "mean_xt_coef": (1.0 - ac_prev) * ((1 - betas) ** 0.5) / (1.0 - ac),
    }


def add_noise(tab, original: Tensor, noise: Tensor, timesteps: Tensor) -> Tensor:
    """models/scheduler.py:95-139 (tables cast to the sample dtype first, then **0.5)."""
    assert timesteps.max().item() < tab["betas"].numel()
    ac = tab["alphas_cumprod"].to(dtype=original.dtype)
    a = (ac[timesteps] ** 0.5).flatten()
    b = ((1 - ac[timesteps]) ** 0.5).flatten()
    while a.dim() < original.dim():
        a = a.unsqueeze(-1)
        b = b.unsqueeze(-1)
    return a * original + b * noise


def ddpm_step(tab, pred_noise: Tensor, x_t: Tensor, time_step: int, noise: Optional[Tensor],
              scale_factor: float = 1.0) -> Tuple[Tensor, Tensor]:
    """models/scheduler.py:141-219.  ``noise`` is the tensor the reference would have drawn
    with ``torch.randn(x_t.shape)`` on the CPU (ignored when time_step == 0: zeros, :200-201).
    Coefficients are cast to ``x_t.dtype`` exactly like ``x_t.new_full`` does."""
    assert time_step < tab["betas"].numel()
    bsz = x_t.shape[0]
    full = lambda name: x_t.new_full((bsz, 1, 1, 1), tab[name][time_step])
    x0 = full("sqrt_recip_alpha_bar") * x_t - full("sqrt_recip_m1_alpha_bar") * pred_noise
    mean = full("mean_x0_coef") * x0 + full("mean_xt_coef") * x_t
    log_var = full("log_var")
    if time_step == 0 or noise is None:
        noise = torch.zeros(x_t.shape)
    noise = noise.to(mean.device, dtype=mean.dtype) * scale_factor
    x_prev = mean + (0.5 * log_var).exp() * noise
    return x_prev, x0


def ddim_step(tab, pred_noise: Tensor, x_t: Tensor, time_step: int) -> Tuple[Tensor, Tensor]:
    """DDIM eta=0 -- NOT in the reference ("parity unpinned").  x0 follows scheduler.py:184;
    x_prev = sqrt(abar_prev) x0 + sqrt(1 - abar_prev) eps with abar_prev = alphas_cumprod[t-1]
    (1.0 at t == 0), using the same dense index convention as ``sample`` (t = loop index)."""
    bsz = x_t.shape[0]
    full = lambda v: x_t.new_full((bsz, 1, 1, 1), v)
    x0 = full(tab["sqrt_recip_alpha_bar"][time_step]) * x_t - full(tab["sqrt_recip_m1_alpha_bar"][time_step]) * pred_noise
    ac_prev = tab["alphas_cumprod"][time_step - 1] if time_step > 0 else torch.tensor(1.0)
    x_prev = full(ac_prev ** 0.5) * x0 + full((1.0 - ac_prev) ** 0.5) * pred_noise
    return x_prev, x0


def cfg_combine(eps2: Tensor, guidance_scale: float) -> Tensor:
    """models/latent_diffusion.py:62-65: first half is the UNCOND batch (c_in = [uncond, ctx],
    :61) but is named 'cond'; result = first + g * (first - second)."""
    first, second = torch.chunk(eps2, 2, dim=0)
    return first + guidance_scale * (first - second)


def pred_noise(unet_fn, x: Tensor, t: Tensor, ctx: Tensor, uncond_ctx: Tensor,
               guidance_scale: float) -> Tensor:
    """models/latent_diffusion.py:27-65.  ``uncond_ctx`` stands for the text encoder's
    embedding of "" (CLIP is out of scope): shape [B, L, D]."""
    if not guidance_scale > 1:
        return unet_fn(x, t, ctx)
    t_in = torch.cat([t] * 2)
    x_in = torch.cat([x] * 2)
    c_in = torch.cat([uncond_ctx, ctx])
    return cfg_combine(unet_fn(x_in, t_in, c_in), guidance_scale)


def sample_timesteps(tab, time_steps: Optional[int]) -> List[int]:
    """models/latent_diffusion.py:103-108: reversed() of a descending range => ASCENDING."""
    if time_steps is not None:
        seq = range(time_steps - 1, -1, -1)
        return [int(s) for s in reversed(seq)]
    return [int(s) for s in reversed(tab["noise_time_steps"])]


def sample(unet_fn, tab, x: Tensor, ctx: Tensor, uncond_ctx: Tensor, guidance_scale: float = 7.5,
           scale_factor: float = 1.0, time_steps: Optional[int] = None,
           noises: Optional[Sequence[Optional[Tensor]]] = None, sampler: str = "ddpm") -> Tensor:
    """models/latent_diffusion.py:67-126.  ``noises[i]`` is the CPU draw of iteration i
    (reference draws ``torch.randn(x.shape)`` from the global CPU generator unless step == 0);
    pass None to draw here in the same order."""
    bsz = x.shape[0]
    for i, step in enumerate(sample_timesteps(tab, time_steps)):
        t = x.new_full((bsz,), step, dtype=torch.long)
        eps = pred_noise(unet_fn, x, t, ctx, uncond_ctx, guidance_scale)
        if sampler == "ddim":
            x, _ = ddim_step(tab, eps, x, step)
            continue
        if noises is not None:
            nz = noises[i]
        else:
            nz = None if step == 0 else torch.randn(x.shape)
        x, _ = ddpm_step(tab, eps, x, step, nz, scale_factor)
    return x


# ----------------------------------------------------------------------------
# train step (row T1): train_unet.py:489-529 (__one_step) and :359-375
# ----------------------------------------------------------------------------
def train_loss(sd: StateDict, cfg: UnetCfg, x_t: Tensor, timesteps: Tensor, ctx: Tensor, uncond_ctx: Tensor,
               noise: Tensor, guidance_scale: float = 7.5) -> Tensor:
    """train_unet.py:524-529: F.mse_loss(pred_noise(x_t, t, ctx, guidance).float(), noise.float(), 'mean');
    pred_noise doubles the batch for guidance > 1 exactly as in sampling (latent_diffusion.py:27-65)."""
    fn = lambda x, t, c: unet_forward(sd, cfg, x, t, c)  # noqa: E731
    pred = pred_noise(fn, x_t, timesteps, ctx, uncond_ctx, guidance_scale)
    return ((pred.float() - noise.float()) ** 2).mean()


def clip_grad_norm(grads: Sequence[Tensor], max_norm: float) -> Tensor:
    """torch.nn.utils.clip_grad_norm_ (train_unet.py:369-371): global L2 norm over all tensors,
    every gradient scaled in place by min(1, max_norm / (norm + 1e-6)).  Returns the pre-clip norm."""
    total = torch.sqrt(sum((g.double() ** 2).sum() for g in grads)).float()
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    for g in grads:
        g.mul_(coef)
    return total


def adamw_update(p: Tensor, g: Tensor, m: Tensor, v: Tensor, step: int, lr: float = 4e-5, beta1: float = 0.9,
                 beta2: float = 0.999, eps: float = 1e-8, weight_decay: float = 0.1) -> None:
    """torch.optim.AdamW single-tensor update (train_unet.py:231-236 builds it with lr and weight_decay only,
    trainer_args.py:71-76 defaults), in place on p, m, v; ``step`` counts from 1."""
    p.mul_(1.0 - lr * weight_decay)
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


def train_steps(sd: StateDict, cfg: UnetCfg, batches, uncond_ctx: Tensor, guidance_scale: float = 7.5,
                max_grad_norm: float = 0.1, lr: float = 4e-5, weight_decay: float = 0.1):
    """Runs len(batches) optimiser steps on a copy of ``sd``; batches = [(x_t, t, ctx, noise), ...].
    Returns (losses, pre-clip grad norms, per-step list of {name: grad}, final params)."""
    params = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    ms = {k: torch.zeros_like(v) for k, v in params.items()}
    vs = {k: torch.zeros_like(v) for k, v in params.items()}
    losses, norms, all_grads = [], [], []
    for i, (x_t, t, ctx, noise) in enumerate(batches, start=1):
        loss = train_loss(params, cfg, x_t, t, ctx, uncond_ctx.expand(x_t.shape[0], -1, -1), noise, guidance_scale)
        names = list(params)
        grads = torch.autograd.grad(loss, [params[k] for k in names])
        grads = [g.clone() for g in grads]
        all_grads.append({k: g.clone() for k, g in zip(names, grads)})
        norms.append(float(clip_grad_norm(grads, max_grad_norm)))
        with torch.no_grad():
            for k, g in zip(names, grads):
                adamw_update(params[k], g, ms[k], vs[k], i, lr=lr, weight_decay=weight_decay)
        losses.append(float(loss.detach()))
    return losses, norms, all_grads, {k: v.detach() for k, v in params.items()}


# ----------------------------------------------------------------------------
# CLIP text encoder (the context producer in front of the path)
# ----------------------------------------------------------------------------
# The reference does not implement it: models/clip_model.py:56-96 wraps the third-party
# ``transformers.CLIPTextModel`` (requirements.txt:8, unpinned; the build container has transformers 5.5.0) and
# returns its output object, of which the hot path uses ``[0]`` = last_hidden_state (latent_diffusion.py:55-60).
# This is a restatement of that published architecture (Radford et al. 2021 text tower as implemented in
# transformers/models/clip/modeling_clip.py: CLIPTextEmbeddings, CLIPEncoderLayer, CLIPTextTransformer.forward):
# token + learned position embeddings; pre-LayerNorm blocks with CAUSAL multi-head self-attention (q, k, v, out
# projections with bias, scale d^-1/2) and a quick-GELU MLP x * sigmoid(1.702 x); final LayerNorm.  Pinned by
# tests/golden/reference_clip.npz, produced by the real CLIPTextModel on seeded weights (make_golden_clip.py).
@dataclass(frozen=True)
class ClipTextCfg:
    vocab_size: int = 49408
    hidden_size: int = 768
    intermediate_size: int = 3072
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    max_position_embeddings: int = 77
    layer_norm_eps: float = 1e-5


def clip_text_param_shapes(cfg: ClipTextCfg) -> Dict[str, Tuple[int, ...]]:
    C, F = cfg.hidden_size, cfg.intermediate_size
    s = {"text_model.embeddings.token_embedding.weight": (cfg.vocab_size, C),
         "text_model.embeddings.position_embedding.weight": (cfg.max_position_embeddings, C),
         "text_model.final_layer_norm.weight": (C,), "text_model.final_layer_norm.bias": (C,)}
    for i in range(cfg.num_hidden_layers):
        p = f"text_model.encoder.layers.{i}."
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            s[p + f"self_attn.{n}.weight"] = (C, C)
            s[p + f"self_attn.{n}.bias"] = (C,)
        for n in ("layer_norm1", "layer_norm2"):
            s[p + n + ".weight"] = (C,)
            s[p + n + ".bias"] = (C,)
        s[p + "mlp.fc1.weight"], s[p + "mlp.fc1.bias"] = (F, C), (F,)
        s[p + "mlp.fc2.weight"], s[p + "mlp.fc2.bias"] = (C, F), (C,)
    return s


def clip_text_forward(sd: StateDict, cfg: ClipTextCfg, input_ids: Tensor) -> Tensor:
    """last_hidden_state [B, L, C] of CLIPTextModel(input_ids) (no padding mask: clip_model.py:87-96 passes ids only)."""
    B, L = input_ids.shape
    C, H = cfg.hidden_size, cfg.num_attention_heads
    d = C // H
    x = sd["text_model.embeddings.token_embedding.weight"][input_ids] \
        + sd["text_model.embeddings.position_embedding.weight"][:L][None]
    causal = torch.full((L, L), float("-inf")).triu(1)
    for i in range(cfg.num_hidden_layers):
        p = f"text_model.encoder.layers.{i}."
        h = F.layer_norm(x, (C,), sd[p + "layer_norm1.weight"], sd[p + "layer_norm1.bias"], cfg.layer_norm_eps)
        q, k, v = (F.linear(h, sd[p + f"self_attn.{n}.weight"], sd[p + f"self_attn.{n}.bias"])
                   .view(B, L, H, d).transpose(1, 2) for n in ("q_proj", "k_proj", "v_proj"))
        a = torch.softmax(q @ k.transpose(-1, -2) * d ** -0.5 + causal, dim=-1) @ v
        a = a.transpose(1, 2).reshape(B, L, C)
        x = x + F.linear(a, sd[p + "self_attn.out_proj.weight"], sd[p + "self_attn.out_proj.bias"])
        h = F.layer_norm(x, (C,), sd[p + "layer_norm2.weight"], sd[p + "layer_norm2.bias"], cfg.layer_norm_eps)
        h = F.linear(h, sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"])
        h = h * torch.sigmoid(1.702 * h)
        x = x + F.linear(h, sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"])
    return F.layer_norm(x, (C,), sd["text_model.final_layer_norm.weight"], sd["text_model.final_layer_norm.bias"],
                        cfg.layer_norm_eps)
