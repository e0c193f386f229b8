"""TimestepEmbedSequential and the sinusoidal projection
(stable_diffusion/modules/timestep_embedding.py:19-105)."""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch
from torch import nn

from .. import _pack, ops
from .resnet2d import ResBlock
from .transformer import SpatialTransformer

Tensor = torch.Tensor

_freq_cache: Dict[Tuple[int, int, str], Tensor] = {}


def sinusoidal_freq(emb_dim: int, max_len: int, device) -> Tensor:
    """The reference's table (timestep_embedding.py:93-99): POSITIVE exponent, evaluated on the
    CPU with the same torch expression so it is bit-identical, uploaded once per device."""
    key = (emb_dim, max_len, str(device))
    f = _freq_cache.get(key)
    if f is None:
        half = emb_dim // 2
        f = torch.exp(math.log(max_len) / half * torch.arange(0, end=half, dtype=torch.float32)).to(device)
        _freq_cache[key] = f
    return f


def sinusoidal_time_proj(time_steps: Tensor, emb_dim: int, max_len: int = 10000) -> Tensor:
    """cat[sin(t * freq), cos(t * freq)] -> fp32 [B, emb_dim] (timestep_embedding.py:73-105)."""
    _pack.require_cuda(time_steps, "sinusoidal_time_proj")
    t = time_steps if time_steps.dtype == torch.int64 else time_steps.to(torch.int64)
    return ops.timestep_embed(t.contiguous(), sinusoidal_freq(emb_dim, max_len, t.device), torch.float32)


class TimestepBlock(nn.Module):
    """Marker base class (timestep_embedding.py:19-26)."""

    def forward(self, x, time_emb):  # pragma: no cover - interface only
        raise NotImplementedError


class TimestepEmbedSequential(nn.Sequential, TimestepBlock):
    """Routes (x, time_emb, context) by child type exactly like timestep_embedding.py:60-70."""

    def forward(self, x: Tensor, time_emb: Optional[Tensor] = None, context: Optional[Tensor] = None):
        for module in self:
            if isinstance(module, ResBlock):
                x = module(x, time_emb)
            elif isinstance(module, SpatialTransformer):
                x = module(x, context)
            else:
                x = module(x)
        return x

    def forward_nhwc(self, x: Tensor, temb_of=None, ctx: Optional[Tensor] = None, kv_of=None,
                     x2: Optional[Tensor] = None, expand: int = 1) -> Tensor:
        """Channels-last chain.  ``temb_of(resblock)`` -> projected time embedding [B, Cout] or None;
        ``kv_of(spatial_transformer)`` -> per-layer precomputed context (K, V) or None; ``x2`` is the
        UNet skip tensor virtually concatenated in front of the first ResBlock.
        ``expand`` > 1 (classifier-free guidance, UNetModel.forward): ``x`` holds every distinct latent ONCE although
        the batch is that latent repeated ``expand`` times; the layers run on the distinct latents until the first
        cross-attention needs the per-copy context (``SpatialTransformer.forward_nhwc``); the result has the full batch."""
        for module in self:
            if expand > 1 and not isinstance(module, (ResBlock, SpatialTransformer)):
                x = x.repeat(expand, 1, 1, 1)                # nothing left to share: materialise the copies
                expand = 1
            if isinstance(module, ResBlock):
                temb = None if temb_of is None else temb_of(module)
                if expand > 1 and temb is not None:
                    temb = temb[:x.shape[0]]                 # (the time steps are repeated like the latents)
                x = module.forward_nhwc(x, temb, x2)
                x2 = None
            elif isinstance(module, SpatialTransformer):
                x = module.forward_nhwc(x, ctx, None if kv_of is None else kv_of(module), expand=expand)
                expand = 1
            elif isinstance(module, TimestepEmbedSequential):
                x = module.forward_nhwc(x, temb_of, ctx, kv_of)
            elif hasattr(module, "forward_nhwc"):
                x = module.forward_nhwc(x)
            elif hasattr(module, "forward_tokens"):  # bare CrossAttention on a feature map (VAE bottleneck)
                B, H, W, C = x.shape
                x = module.forward_tokens(x.view(B, H * W, C)).view(B, H, W, C)
            else:
                raise TypeError(f"no channels-last path for {type(module).__name__}")
        if expand > 1:
            x = x.repeat(expand, 1, 1, 1)
        assert x2 is None, "skip tensor was not consumed by a ResBlock"
        return x
