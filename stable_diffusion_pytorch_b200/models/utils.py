"""Block builders shared by the UNet and the VAE (topology + state_dict key layout of
stable_diffusion/models/utils.py:31-200, including its quirks: attention only where the running
``attn_mult`` is listed in ``attention_resolutions``; ``d_head`` of the last attention level leaking
into the bottleneck; the first bottleneck ResBlock built with the ResBlock default ``groups=2``)."""
from __future__ import annotations

from typing import List, Optional, Sequence

from torch import nn

from .._pack import zero_module  # noqa: F401  (re-exported: reference exposes it here)
from ..modules.resnet2d import DownSample, ResBlock, UpSample
from ..modules.timestep_embedding import TimestepEmbedSequential
from ..modules.transformer import CrossAttention, SpatialTransformer


def build_conv_in(in_channels: int, out_channels: int):
    return nn.Conv2d(in_channels, out_channels, 3, padding=1)


def _wants_attention(attention_resolutions, attn_mult) -> bool:
    return attention_resolutions is not None and attn_mult in attention_resolutions


def build_input_blocks(in_channels: int, num_res_blocks: int, attention_resolutions: Optional[Sequence[int]] = None,
                       n_heads: int = 1, n_layers: int = 1, dropout: float = 0.0,
                       context_dim: Optional[int] = None, levels: int = 1, time_emb_dim: Optional[int] = None,
                       channels_list: Sequence[int] = None, groups: int = 4):
    blocks = nn.ModuleList()
    skip_channels: List[int] = [in_channels]
    ch, attn_mult, d_head = in_channels, 1, None
    for level in range(levels):
        for _ in range(num_res_blocks):
            # NB: like the reference, input-level ResBlocks are built WITHOUT dropout
            stage = [ResBlock(in_channels=ch, out_channels=channels_list[level], time_emb_dim=time_emb_dim,
                              groups=groups)]
            ch = channels_list[level]
            if _wants_attention(attention_resolutions, attn_mult):
                d_head = ch // n_heads
                stage.append(SpatialTransformer(ch, n_heads=n_heads, d_head=d_head, n_layers=n_layers,
                                                dropout=dropout, context_dim=context_dim, groups=groups))
            blocks.append(TimestepEmbedSequential(*stage))
            skip_channels.append(ch)
        if level != levels - 1:
            blocks.append(TimestepEmbedSequential(DownSample(ch)))
            skip_channels.append(ch)
            attn_mult *= 2
    return blocks, skip_channels, ch, d_head, attn_mult


def build_bottleneck(in_ch: int, time_emb_dim: Optional[int] = None, n_heads: int = 1, d_head: int = 1,
                     n_layers: int = 1, dropout: float = 0.0, context_dim: Optional[int] = None,
                     use_attn_only: bool = False, groups: int = 4):
    if use_attn_only:
        mid = CrossAttention(query_dim=in_ch, n_heads=n_heads, d_head=d_head, dropout=dropout)
    else:
        mid = SpatialTransformer(in_ch, n_heads=n_heads, d_head=d_head, n_layers=n_layers, dropout=dropout,
                                 context_dim=context_dim, groups=groups)
    return TimestepEmbedSequential(
        ResBlock(in_channels=in_ch, time_emb_dim=time_emb_dim, dropout=dropout),  # groups: ResBlock default (2)
        mid,
        ResBlock(in_channels=in_ch, time_emb_dim=time_emb_dim, dropout=dropout, groups=groups))


def build_output_blocks(num_res_blocks: int, attention_resolutions: Optional[Sequence[int]] = None,
                        n_heads: int = 1, n_layers: int = 1, dropout: float = 0.0,
                        context_dim: Optional[int] = None, input_block_channels: Optional[List[int]] = None,
                        levels: int = 1, time_emb_dim: Optional[int] = None, channels_list: Sequence[int] = None,
                        in_ch: int = 1, attn_mult: Optional[int] = None, groups: int = 4):
    blocks = nn.ModuleList()
    ch = in_ch
    for level in reversed(range(levels)):
        for i in range(num_res_blocks + 1):
            skip = input_block_channels.pop() if input_block_channels else 0
            stage = [ResBlock(in_channels=ch + skip, out_channels=channels_list[level], time_emb_dim=time_emb_dim,
                              dropout=dropout, groups=groups)]
            ch = channels_list[level]
            if _wants_attention(attention_resolutions, attn_mult):
                stage.append(SpatialTransformer(ch, n_heads=n_heads, d_head=ch // n_heads, n_layers=n_layers,
                                                dropout=dropout, context_dim=context_dim, groups=groups))
            if level != 0 and i == num_res_blocks:
                stage.append(TimestepEmbedSequential(UpSample(ch)))
                if attn_mult:
                    attn_mult //= 2
            blocks.append(TimestepEmbedSequential(*stage))
    return blocks, ch


def build_final_output(out_ch: int, out_channels: int, groups: int = 4):
    return nn.Sequential(nn.GroupNorm(groups, out_ch), nn.SiLU(),
                         nn.Conv2d(in_channels=out_ch, out_channels=out_channels, kernel_size=3, padding=1))
