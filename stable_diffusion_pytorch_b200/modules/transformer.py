"""CrossAttention / GEGLU / FeedForward / BasicTransformerBlock / SpatialTransformer with the
reference's signatures and state_dict keys (stable_diffusion/modules/transformer.py:18-383).

Tokens are [B, n, C] bf16 -- the same memory as the NHWC feature map, so the reference's two
``rearrange`` calls (:378,:381) vanish.  Kernel plan per BasicTransformerBlock (post-norm, :274-278):
fused-QKV tcgen05 GEMM -> fused attention -> out-proj GEMM with bias+residual epilogue -> LayerNorm;
Q GEMM + (precomputed) context K/V -> fused attention -> out-proj(+res) -> LayerNorm;
GEGLU GEMM (gate applied in the epilogue) -> FFN-out GEMM(+res) -> LayerNorm.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import nn

from .. import _pack, ops
from .._lib import ACT_GEGLU, ACT_NONE

Tensor = torch.Tensor


def _tokens_bf16(x: Tensor) -> Tensor:
    _pack.require_cuda(x, "transformer")
    x = x.contiguous()
    cd = _pack.compute_dtype()
    return x if x.dtype == cd else ops.cast(x, cd)


class CrossAttention(nn.Module):
    """softmax(Q K^T / sqrt(d)) V with bias-free q/k/v projections and a biased output projection
    (transformer.py:18-119).  ``context_emb=None`` => self-attention.  4-D inputs [B, C, H, W] are
    accepted like the reference's ``convert`` path (used by the VAE bottleneck)."""

    def __init__(self, query_dim: int, context_dim: Optional[int] = None, n_heads: int = 1, d_head: int = 1,
                 dropout: float = 0.0):
        super().__init__()
        if not context_dim:
            context_dim = query_dim
        d_model = n_heads * d_head
        self.n_heads = n_heads
        self.d_head = d_head
        self.scale = 1 / (d_head ** 0.5)
        self.to_q = nn.Linear(query_dim, d_model, bias=False)
        self.to_k = nn.Linear(context_dim, d_model, bias=False)
        self.to_v = nn.Linear(context_dim, d_model, bias=False)
        self.out = nn.Sequential(nn.Linear(d_model, query_dim), nn.Dropout(dropout))
        self._cache = _pack.PackCache()
        self._self_qkv_ok = context_dim == query_dim

    @property
    def use_p32(self) -> bool:
        """Small head dims run on the tcgen05 attention kernel over head-padded operands."""
        return self.d_head <= 20 and self.d_head % 4 == 0 and _pack.compute_dtype() == torch.bfloat16

    @property
    def kv_width(self) -> int:
        """Columns of one of K / V as produced by the packed context projection."""
        return self.n_heads * (_pack.HEAD_PAD if self.use_p32 else self.d_head)

    def kv_weight(self) -> Tuple[Tensor, Tensor]:
        return self.to_k.weight, self.to_v.weight

    def kv_packed(self) -> Tuple[Tensor, Tensor]:
        """(weight bf16 [2*kv_width, Cc], bias fp32 [2*kv_width]) of the fused K|V projection."""
        c = self._cache
        k, v = self.to_k.weight, self.to_v.weight
        H, d = self.n_heads, self.d_head

        def build_w():
            if self.use_p32:
                return _pack.linear_w(torch.cat([_pack.head_pad_w(k, H, d), _pack.head_pad_w(v, H, d)], 0))
            return _pack.linear_w(torch.cat([_pack.d(k), _pack.d(v)], 0))

        def build_b():
            kw = self.kv_width
            b = torch.zeros(2 * kw, dtype=torch.float32, device=k.device)
            if self.use_p32:
                b[kw:] = _pack.head_pad_ones_bias(H, d, k.device)
            return b

        return c.get("wkv", [k, v], build_w), c.get("bkv", [k, v], build_b)

    def _w(self):
        c = self._cache
        q, k, v, o = self.to_q.weight, self.to_k.weight, self.to_v.weight, self.out[0]
        H, dh = self.n_heads, self.d_head
        pad = (lambda w: _pack.head_pad_w(w, H, dh)) if self.use_p32 else (lambda w: _pack.d(w))
        d = {
            "wq": c.get("wq", [q], lambda: _pack.linear_w(pad(q))),
            "wo": c.get("wo", [o.weight], lambda: _pack.linear_w(o.weight)),
            "bo": c.get("bo", [o.bias], lambda: _pack.f32(o.bias)),
        }
        d["wkv"], d["bkv"] = self.kv_packed()
        if self._self_qkv_ok:
            d["wqkv"] = c.get("wqkv", [q, k, v], lambda: _pack.linear_w(torch.cat([pad(q), pad(k), pad(v)], 0)))
            kw = self.kv_width
            d["bqkv"] = c.get("bqkv", [q, k, v],
                              lambda: torch.cat([torch.zeros(kw, dtype=torch.float32, device=q.device), d["bkv"]]))
        return d

    def forward_tokens(self, x: Tensor, ctx: Optional[Tensor] = None, kv: Optional[Tuple[Tensor, Tensor]] = None,
                       residual: Optional[Tensor] = None, ln=None) -> Tensor:
        """x [B, n, C] bf16.  ctx [B, m, Cc] bf16 or ``kv`` = precomputed (K, V) [B, m, kv_width] views (the
        output layout of ``kv_packed``).  Returns out-proj(attn) (+ residual, fused in the GEMM epilogue); with
        ``ln`` = (gamma, beta) the post-norm LayerNorm of the block (transformer.py:274-278) is applied to that sum,
        inside the same epilogue when the row fits one tile."""
        B, n, C = x.shape
        dm = self.n_heads * self.d_head
        kw = self.kv_width
        w = self._w()
        x2d = x.view(B * n, C)
        if ctx is None and kv is None:
            if "wqkv" not in w:
                raise ValueError("CrossAttention: context_emb is None but context_dim != query_dim "
                                 "(the reference fails with a shape mismatch here too)")
            qkv = ops.gemm(x2d, w["wqkv"], w["bqkv"] if self.use_p32 else None).view(B, n, 3 * kw)
            q, k, v = qkv[:, :, :kw], qkv[:, :, kw:2 * kw], qkv[:, :, 2 * kw:]
            packed = qkv if self.use_p32 else None
        else:
            packed = None
            q = ops.gemm(x2d, w["wq"]).view(B, n, kw)
            if kv is None:
                m = ctx.shape[1]
                kvt = ops.gemm(ctx.view(B * m, ctx.shape[2]), w["wkv"],
                               w["bkv"] if self.use_p32 else None).view(B, m, 2 * kw)
                k, v = kvt[:, :, :kw], kvt[:, :, kw:]
            else:
                k, v = kv
        if packed is not None:
            a = ops.attention_p32_packed(packed, self.n_heads, self.d_head, self.scale)
        elif self.use_p32:
            a = ops.attention_p32(q, k, v, self.n_heads, self.d_head, self.scale)
        else:
            a = ops.attention(q, k, v, self.n_heads, self.d_head, self.scale)
        res2d = None if residual is None else residual.view(B * n, C)
        if self.training and self.out[1].p:
            # out = Sequential(Linear, Dropout) (transformer.py:58): the mask sits between the projection and the
            # residual add, so the add leaves the GEMM epilogue
            o = ops.dropout(ops.gemm(a.view(B * n, dm), w["wo"], w["bo"]), self.out[1].p, True)
            o = (o if res2d is None else o + res2d).view(B, n, C)
            return o if ln is None else ops.add_layernorm(o, None, *ln)
        return ops.gemm(a.view(B * n, dm), w["wo"], w["bo"], residual=res2d, ln=ln).view(B, n, C)

    def forward(self, query: Tensor, context_emb: Optional[Tensor] = None, mask: Optional[Tensor] = None):
        if mask is not None:
            raise NotImplementedError("CrossAttention: the reference's mask path is dead code (transformer.py:102-107)")
        in_dtype = query.dtype
        if query.dim() == 4:
            B, C, H, W = query.shape
            t = ops.nchw_to_nhwc(query, _pack.compute_dtype()).view(B, H * W, C)
            ctx = None if context_emb is None else _tokens_bf16(context_emb)
            o = self.forward_tokens(t, ctx)
            return ops.nhwc_to_nchw(o.view(B, H, W, C), in_dtype)
        t = _tokens_bf16(query)
        ctx = None if context_emb is None else _tokens_bf16(context_emb)
        o = self.forward_tokens(t, ctx)
        return o if in_dtype == o.dtype else ops.cast(o, in_dtype)


class GEGLU(nn.Module):
    """x, gate = Linear(C -> 2F).chunk(2); x * gelu_erf(gate) (transformer.py:158-192), one GEMM:
    the weight rows are interleaved at pack time so the gate is applied in the epilogue."""

    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        self.proj = nn.Linear(in_features, out_features * 2)
        self.gelu = nn.GELU()
        self._cache = _pack.PackCache()

    def forward_tokens(self, x2d: Tensor) -> Tensor:
        w, b = self._cache.get("wb", [self.proj.weight, self.proj.bias],
                               lambda: _pack.geglu_w(self.proj.weight, self.proj.bias))
        return ops.gemm(x2d, w, b, act=ACT_GEGLU)

    def forward(self, x: Tensor) -> Tensor:
        t = _tokens_bf16(x)
        o = self.forward_tokens(t.view(-1, t.shape[-1])).view(*x.shape[:-1], -1)
        return o if x.dtype == o.dtype else ops.cast(o, x.dtype)


class FeedForward(nn.Module):
    """GEGLU -> Dropout -> Linear(4C -> C) (transformer.py:122-155)."""

    def __init__(self, d_model: int, dim_mult: int = 4, dropout: float = 0.0):
        super().__init__()
        self.net = nn.Sequential(GEGLU(d_model, dim_mult * d_model), nn.Dropout(p=dropout),
                                 nn.Linear(d_model * dim_mult, d_model))
        self._cache = _pack.PackCache()

    def forward_tokens(self, x2d: Tensor, residual: Optional[Tensor] = None, ln=None) -> Tensor:
        lin = self.net[2]
        w = self._cache.get("w", [lin.weight], lambda: _pack.linear_w(lin.weight))
        b = self._cache.get("b", [lin.bias], lambda: _pack.f32(lin.bias))
        g = ops.dropout(self.net[0].forward_tokens(x2d), self.net[1].p, self.training)      # net[1] (transformer.py:147)
        return ops.gemm(g, w, b, residual=residual, ln=ln)

    def forward(self, x: Tensor) -> Tensor:
        t = _tokens_bf16(x)
        o = self.forward_tokens(t.view(-1, t.shape[-1])).view(x.shape)
        return o if x.dtype == o.dtype else ops.cast(o, x.dtype)


class BasicTransformerBlock(nn.Module):
    """Post-norm block: x = LN1(x + SA(x)); x = LN2(x + CA(x, ctx)); x = LN3(x + FFN(x))
    (transformer.py:195-279)."""

    def __init__(self, d_model: int, n_heads: int, d_head: int, dropout: float = 0.0, context_dim: int = 768):
        super().__init__()
        self.d_model = d_model
        self.context_dim = context_dim
        self.self_attn = CrossAttention(d_model, context_dim=d_model, n_heads=n_heads, d_head=d_head, dropout=dropout)
        self.norm1 = nn.LayerNorm(d_model)
        self.cross_attn = CrossAttention(d_model, context_dim=context_dim, n_heads=n_heads, d_head=d_head,
                                         dropout=dropout)
        self.norm2 = nn.LayerNorm(d_model)
        self.ffn = FeedForward(d_model, dropout=dropout)
        self.norm3 = nn.LayerNorm(d_model)
        self._cache = _pack.PackCache()

    def _ln(self, i: int):
        ln = getattr(self, f"norm{i}")
        return (self._cache.get(f"g{i}", [ln.weight], lambda: _pack.f32(ln.weight)),
                self._cache.get(f"b{i}", [ln.bias], lambda: _pack.f32(ln.bias)))

    def forward_tokens(self, x: Tensor, ctx: Optional[Tensor] = None,
                       kv: Optional[Tuple[Tensor, Tensor]] = None, expand: int = 1) -> Tensor:
        B, n, C = x.shape
        # every residual add AND the LayerNorm after it ride in the epilogue of the GEMM that ends the sub-layer
        x = self.self_attn.forward_tokens(x, None, None, residual=x, ln=self._ln(1))      # LN1(x + SA(x))
        if expand > 1:                                       # the cross-attention is the first per-copy layer
            x = x.repeat(expand, 1, 1)
            B = B * expand
        if ctx is None and kv is None:
            x = self.cross_attn.forward_tokens(x, None, None, residual=x, ln=self._ln(2))  # reference: ctx None => self-attn
        else:
            x = self.cross_attn.forward_tokens(x, ctx, kv, residual=x, ln=self._ln(2))    # LN2(x + CA(x, ctx))
        return self.ffn.forward_tokens(x.view(B * n, C), residual=x.view(B * n, C),
                                       ln=self._ln(3)).view(B, n, C)                        # LN3(x + FFN(x))

    def forward(self, x: Tensor, context_emb: Optional[Tensor] = None) -> Tensor:
        assert x.shape[-1] == self.d_model, f"input dim {x.shape[-1]} should be equal to d_model {self.d_model}"
        if context_emb is not None:
            assert context_emb.shape[-1] == self.context_dim, (
                f"context dim {context_emb.shape[-1]} should be equal to context_dim {self.context_dim}")
        t = _tokens_bf16(x)
        ctx = None if context_emb is None else _tokens_bf16(context_emb)
        o = self.forward_tokens(t, ctx)
        return o if x.dtype == o.dtype else ops.cast(o, x.dtype)


class SpatialTransformer(nn.Module):
    """GN (no SiLU) -> 1x1 conv -> n_layers blocks -> 1x1 conv (zero-init) -> + x_in
    (transformer.py:282-383)."""

    @property
    def act_dtype(self):
        return _pack.compute_dtype()

    def __init__(self, in_channels: int, n_heads: int, d_head: int, n_layers: int = 1, dropout: float = 0.0,
                 context_dim: int = None, groups: int = 2):
        super().__init__()
        assert n_heads > 0, f"n_heads({n_heads}) should be greater than 0 for SpatialTransformer"
        assert in_channels % groups == 0, (
            f"in_channels({in_channels}) should be divisible by num_groups({groups}) for GroupNorm")
        self.in_channels = in_channels
        self.context_dim = context_dim
        self.groups = groups
        self.norm = nn.GroupNorm(groups, in_channels)
        self.proj_in = nn.Conv2d(in_channels, in_channels, kernel_size=1, padding=0)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(in_channels, n_heads, d_head, dropout=dropout, context_dim=context_dim)
             for _ in range(n_layers)])
        self.proj_out = _pack.zero_module(nn.Conv2d(in_channels, in_channels, kernel_size=1, padding=0))
        self._cache = _pack.PackCache()

    def forward_nhwc(self, x: Tensor, ctx: Optional[Tensor] = None, kvs=None, expand: int = 1) -> Tensor:
        """x NHWC bf16; ctx [B, m, Cc] bf16; kvs = optional per-layer precomputed (K, V).
        ``expand`` > 1: ``x`` holds B / expand distinct latents of a batch that repeats them ``expand`` times (classifier-free
        guidance: ``torch.cat([x] * 2)``, latent_diffusion.py:53); GroupNorm, proj_in and the first block's self-attention
        -- everything ahead of the first cross-attention -- are evaluated once per distinct latent and copied."""
        B, H, W, C = x.shape
        assert C == self.in_channels, f"input channels {C} should be equal to in_channels {self.in_channels}"
        c = self._cache
        g = c.get("g", [self.norm.weight], lambda: _pack.f32(self.norm.weight))
        b = c.get("b", [self.norm.bias], lambda: _pack.f32(self.norm.bias))
        wi = c.get("wi", [self.proj_in.weight], lambda: _pack.conv1x1_w(self.proj_in.weight))
        bi = c.get("bi", [self.proj_in.bias], lambda: _pack.f32(self.proj_in.bias))
        wo = c.get("wo", [self.proj_out.weight], lambda: _pack.conv1x1_w(self.proj_out.weight))
        bo = c.get("bo", [self.proj_out.bias], lambda: _pack.f32(self.proj_out.bias))
        h = ops.group_norm(x, g, b, self.groups, ACT_NONE)
        t = ops.gemm(h.view(B * H * W, C), wi, bi).view(B, H * W, C)
        for i, blk in enumerate(self.transformer_blocks):
            t = blk.forward_tokens(t, ctx, None if kvs is None else kvs[i], expand=expand if i == 0 else 1)
        if expand > 1:
            if len(self.transformer_blocks) == 0:
                t = t.repeat(expand, 1, 1)
            x = x.repeat(expand, 1, 1, 1)                    # the residual of proj_out, per copy
            B = B * expand
        o = ops.gemm(t.view(B * H * W, C), wo, bo, residual=x.view(B * H * W, C), sums_rows=H * W)
        return ops.carry_sums(o, o.view(B, H, W, C))      # the next ResBlock's GroupNorm reads it

    def forward(self, x: Tensor, context_emb: Tensor = None) -> Tensor:
        assert x.shape[1] == self.in_channels, (
            f"input channels {x.shape[1]} should be equal to in_channels {self.in_channels}")
        if context_emb is not None:
            assert context_emb.shape[-1] == self.context_dim, (
                f"context dim {context_emb.shape[-1]} should be equal to context_dim {self.context_dim}")
        _pack.require_cuda(x, "SpatialTransformer")
        ctx = None if context_emb is None else _tokens_bf16(context_emb)
        o = self.forward_nhwc(ops.nchw_to_nhwc(x, self.act_dtype), ctx)
        return ops.nhwc_to_nchw(o, x.dtype)
