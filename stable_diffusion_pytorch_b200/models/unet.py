"""UNetModel -- eps_theta(x_t, t, context) -- with the reference's constructor, ``forward(x,
timesteps, context_emb)`` signature and state_dict keys (stable_diffusion/models/unet.py:29-310).

forward() plan (all kernels in libsdb200, activations NHWC bf16):
  1. K6 sinusoid -> time MLP (two tcgen05 GEMMs with SiLU epilogues) -> ONE batched GEMM that yields
     every ResBlock's ``Linear(SiLU(temb))`` projection (resnet2d.py:179-183, 12 of them).
  2. ONE batched GEMM computes K and V of the text context for all cross-attention layers.
  3. conv_in (CUDA-core, reads NCHW), then the block chain on channels-last tensors; the skip
     ``torch.cat`` (unet.py:308) is virtual: GroupNorm and the 1x1 skip GEMM read both halves.
  4. GroupNorm+SiLU, head conv (CUDA-core, writes NCHW in the caller's dtype).
"""
from __future__ import annotations

from dataclasses import dataclass, field
import os
from typing import Dict, List, Optional

import torch
from torch import nn

from .. import _pack, ops
from .._lib import ACT_SILU
from ..modules.resnet2d import ResBlock
from ..modules.timestep_embedding import TimestepEmbedSequential, sinusoidal_freq, sinusoidal_time_proj
from ..modules.transformer import SpatialTransformer
from .utils import (build_bottleneck, build_conv_in, build_final_output, build_input_blocks,
                    build_output_blocks)

_CFG_SHARE = os.environ.get("SDB_CFG_SHARE", "1") != "0"   # A/B switch of the CFG prefix sharing in UNetModel.forward

Tensor = torch.Tensor


@dataclass
class UnetConfig:
    """Field names and defaults of the reference's UnetConfig (unet.py:29-57)."""
    num_res_blocks: int = 2
    n_heads: int = 8
    attention_resolutions: List[int] = field(default_factory=lambda: [0, 1])
    channels_list: List[int] = field(default_factory=lambda: [160, 320])
    time_emb_dim: Optional[int] = 512
    dropout: float = 0.1
    n_layers: int = 2
    context_dim: int = 768


class UNetModel(nn.Module):
    def __init__(self, latent_channels, groups, cfg):
        super().__init__()
        channels_list = cfg.channels_list
        self.context_dim = cfg.context_dim
        self.channels = channels_list[0]
        self.latent_channels = latent_channels
        self.groups = groups
        time_emb_dim = cfg.time_emb_dim or channels_list[0] * 4
        self.time_emb_dim = time_emb_dim
        self.time_embedding = nn.Sequential(nn.Linear(channels_list[0], time_emb_dim), nn.SiLU(),
                                            nn.Linear(time_emb_dim, time_emb_dim))
        self.conv_in = build_conv_in(latent_channels, self.channels)
        levels = len(channels_list)
        self.input_blocks, skip_channels, mid_ch, d_head, attn_mult = build_input_blocks(
            in_channels=self.channels, num_res_blocks=cfg.num_res_blocks,
            attention_resolutions=cfg.attention_resolutions, n_heads=cfg.n_heads, n_layers=cfg.n_layers,
            dropout=cfg.dropout, context_dim=cfg.context_dim, levels=levels, time_emb_dim=time_emb_dim,
            channels_list=channels_list, groups=groups)
        self.middle_block = build_bottleneck(
            in_ch=mid_ch, time_emb_dim=time_emb_dim, n_heads=cfg.n_heads, d_head=d_head, n_layers=cfg.n_layers,
            dropout=cfg.dropout, context_dim=cfg.context_dim, groups=groups)
        self.output_blocks, out_ch = build_output_blocks(
            num_res_blocks=cfg.num_res_blocks, attention_resolutions=cfg.attention_resolutions,
            n_heads=cfg.n_heads, n_layers=cfg.n_layers, dropout=cfg.dropout, context_dim=cfg.context_dim,
            input_block_channels=skip_channels, levels=levels, time_emb_dim=time_emb_dim,
            channels_list=channels_list, in_ch=mid_ch, attn_mult=attn_mult, groups=groups)
        self.out = build_final_output(out_ch=out_ch, out_channels=latent_channels, groups=groups)
        self._cache = _pack.PackCache()
        self._ctx_cache = None  # (key, per-transformer K/V views) of the last context tensor
        self._rbs = tuple(m for m in self.modules() if isinstance(m, ResBlock) and m.time_emb_dim)
        self._sts = tuple(m for m in self.modules() if isinstance(m, SpatialTransformer))

    # ------------------------------------------------------------------------------------------
    def time_proj(self, time_steps: Tensor, max_len: int = 10000) -> Tensor:
        return sinusoidal_time_proj(time_steps, self.channels, max_len)

    def _resblocks(self):
        return self._rbs

    def _transformers(self):
        return self._sts

    def _packed(self):
        c = self._cache
        l0, l2 = self.time_embedding[0], self.time_embedding[2]
        gn, head = self.out[0], self.out[2]
        rbs = self._resblocks()
        tw = [rb.time_proj_params() for rb in rbs]
        d = {
            "w0": c.get("w0", [l0.weight], lambda: _pack.linear_w(l0.weight)),
            "b0": c.get("b0", [l0.bias], lambda: _pack.f32(l0.bias)),
            "w2": c.get("w2", [l2.weight], lambda: _pack.linear_w(l2.weight)),
            "b2": c.get("b2", [l2.bias], lambda: _pack.f32(l2.bias)),
            "wt": c.get("wt", [w for w, _ in tw], lambda: _pack.linear_w(torch.cat([_pack.d(w) for w, _ in tw], 0))),
            "bt": c.get("bt", [b for _, b in tw], lambda: _pack.f32(torch.cat([_pack.d(b) for _, b in tw], 0))),
            "wci": c.get("wci", [self.conv_in.weight], lambda: _pack.f32(self.conv_in.weight)),
            "bci": c.get("bci", [self.conv_in.bias], lambda: _pack.f32(self.conv_in.bias)),
            "go": c.get("go", [gn.weight], lambda: _pack.f32(gn.weight)),
            "bo": c.get("bo", [gn.bias], lambda: _pack.f32(gn.bias)),
            "wh": c.get("wh", [head.weight, head.bias], lambda: _pack.head_w16(head.weight, head.bias)),
        }
        offs, o = {}, 0
        for rb in rbs:
            offs[id(rb)] = (o, rb.out_channels)
            o += rb.out_channels
        d["temb_offsets"] = offs
        # context K/V weights of every cross-attention layer, concatenated along N
        kvw, packed, kv_offs, o = [], [], {}, 0
        for st in self._transformers():
            lay = []
            for blk in st.transformer_blocks:
                ca = blk.cross_attn
                kvw += list(ca.kv_weight())
                packed.append(ca.kv_packed)
                lay.append((o, ca.kv_width))
                o += 2 * ca.kv_width
            kv_offs[id(st)] = lay
        d["kv_offsets"] = kv_offs
        d["wkv"] = c.get("wkv", kvw, lambda: torch.cat([f()[0] for f in packed], 0)) if kvw else None
        d["bkv"] = c.get("bkv", kvw, lambda: torch.cat([f()[1] for f in packed], 0)) if kvw else None
        return d

    def _context_kv(self, p, ctx: Tensor) -> Dict[int, list]:
        """All cross-attention K/V of ``ctx`` in one GEMM; reused while the same context tensor
        (same storage and version) comes back, i.e. across the steps of a sampling loop."""
        key = (ctx.data_ptr(), ctx._version, tuple(ctx.shape), ctx.dtype, id(p["wkv"]), _pack.epoch())
        cacheable = not (torch.cuda.is_current_stream_capturing() or _pack.in_graph_mode())
        if cacheable and self._ctx_cache is not None and self._ctx_cache[0] == key:
            return self._ctx_cache[1]
        B, m, cd = ctx.shape
        act = _pack.compute_dtype()
        c16 = ctx.contiguous() if ctx.dtype == act else ops.cast(ctx.contiguous(), act)
        kv = ops.gemm(c16.view(B * m, cd), p["wkv"], p["bkv"]).view(B, m, -1)
        views = {sid: [(kv[:, :, o:o + dm], kv[:, :, o + dm:o + 2 * dm]) for (o, dm) in lay]
                 for sid, lay in p["kv_offsets"].items()}
        if cacheable:
            self._ctx_cache = (key, views, ctx)
        return views

    def forward(self, x: Tensor, timesteps: Tensor, context_emb: Optional[Tensor] = None, x_repeat: int = 1):
        """``x_repeat`` (extension, default 1): the effective batch is ``x`` repeated that many times
        along dim 0 -- the ``torch.cat([x] * 2)`` of classifier-free guidance without the copy;
        ``timesteps`` / ``context_emb`` carry the full repeated batch, and ``timesteps`` repeats like ``x``
        (row i + k * B equals row i, as ``torch.cat([time_step] * 2)`` of latent_diffusion.py:54 does): the layers ahead
        of the first cross-attention are evaluated once per distinct latent."""
        if context_emb is not None:
            assert context_emb.shape[-1] == self.context_dim, (
                f"context dim from passed in context({context_emb.shape}) should be equal to "
                f"self.context_dim({self.context_dim})")
        _pack.require_cuda(x, "UNetModel.forward")
        p = self._packed()
        assert timesteps.shape[0] == x.shape[0] * x_repeat, "timesteps must cover the (repeated) batch"
        # 1. time embedding chain
        t = timesteps if timesteps.dtype == torch.int64 else timesteps.to(torch.int64)
        emb = ops.timestep_embed(t.contiguous(), sinusoidal_freq(self.channels, 10000, x.device), _pack.compute_dtype())
        h1 = ops.gemm(emb, p["w0"], p["b0"], act=ACT_SILU)
        st = ops.gemm(h1, p["w2"], p["b2"], act=ACT_SILU)          # SiLU(time_emb), shared by all ResBlocks
        temb_all = ops.gemm(st, p["wt"], p["bt"], out_dtype=torch.float32)
        offs = p["temb_offsets"]

        def temb_of(rb):
            o, n = offs[id(rb)]
            return temb_all[:, o:o + n]

        # 2. context K/V for all cross-attention layers
        kv_of = None
        if context_emb is not None and p["wkv"] is not None:
            views = self._context_kv(p, context_emb)
            kv_of = lambda st_mod: views[id(st_mod)]  # noqa: E731
        # 3. block chain
        # (inside the scope the conv / GEMM epilogues leave per-channel sums for the GroupNorm that reads them next)
        # classifier-free guidance (x_repeat > 1): the copies of a latent differ only in their context, so everything
        # ahead of the first cross-attention -- conv_in, the first ResBlock, the first transformer's GroupNorm, proj_in
        # and self-attention (one of ten 4096 x 4096 attentions at config 2) -- is evaluated once per distinct latent.
        # Same arithmetic per sample; SDB_CFG_SHARE=0 evaluates every copy (A/B).  Inference only.
        share = (x_repeat > 1 and _CFG_SHARE and not torch.is_grad_enabled() and context_emb is not None
                 and len(self.input_blocks) > 0 and isinstance(self.input_blocks[0], TimestepEmbedSequential))
        with ops.gn_sums_scope(x.device):
            h = ops.conv3x3_in(x, p["wci"], p["bci"], _pack.compute_dtype(), repeat=1 if share else x_repeat)
            skips = [h.repeat(x_repeat, 1, 1, 1) if share else h]
            for i, module in enumerate(self.input_blocks):
                if share and i == 0:
                    h = module.forward_nhwc(h, temb_of, None, kv_of, expand=x_repeat)
                else:
                    h = module.forward_nhwc(h, temb_of, None, kv_of)
                skips.append(h)
            h = self.middle_block.forward_nhwc(h, temb_of, None, kv_of)
            for module in self.output_blocks:
                h = module.forward_nhwc(h, temb_of, None, kv_of, x2=skips.pop())
            # 4. head
            h = ops.group_norm(h, p["go"], p["bo"], self.groups, ACT_SILU)
        return ops.conv3x3_head(h, p["wh"][0], p["wh"][1], self.latent_channels, x.dtype)
