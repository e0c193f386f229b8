"""Autograd wrappers of the fused ops (row T1: train_unet.py:489-529).

``ops.gemm`` / ``conv3x3`` / ``group_norm`` / ``add_layernorm`` / ``attention_p32`` / ... dispatch here when
grad mode is on and an input requires grad.  Every forward is the inference kernel; every backward runs this
library's kernels as well:

* dgrad of a GEMM / 3x3 conv = the same tcgen05 kernel with transposed (and flipped) weights,
* wgrad = ``csrc/wgrad_umma.cu``: dY^T . X on the tensor cores with both operands read as MN-major tiles straight
  from their channels-last rows (conv taps through shifted TMA boxes), split over the pixel dimension with fp32
  atomics; shapes outside it (thin conv_in / head) go through transposed copies / im2col and the forward GEMM,
* attention: flash-style recomputation (``csrc/attention_bwd.cu``); norms, GEGLU, SiLU, resampling:
  ``csrc/backward.cu``.

torch is used for tensor views, allocation and the (tiny) weight-layout permutes only.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
from torch.autograd import Function

from . import _lib
from ._lib import ACT_GEGLU, ACT_NONE, ACT_SILU, check

Tensor = torch.Tensor
_CTX_BWD = os.environ.get("SDB_BWD_CTX", "1") != "0"          # 0 = cross-attention backward on the generic flash kernels
_DKV_UMMA = os.environ.get("SDB_BWD_DKV_UMMA", "1") != "0"
_DQ_UMMA = os.environ.get("SDB_BWD_DQ_UMMA", "1") != "0"      # tuning / A-B aid: 0 = dQ on the mma.sync kernel


def wants(*tensors) -> bool:
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


def refuse(op: str, *tensors) -> None:
    """Ops without a backward must not silently cut the graph."""
    if wants(*tensors):
        raise NotImplementedError(f"{op}: no backward is built for this op (it is not on the UNet train path); "
                                  f"call it under torch.no_grad() or detach its inputs")


def _ops():
    from . import ops
    return ops


def _bf16(t: Tensor) -> Tensor:
    t = t.contiguous()
    return t if t.dtype == torch.bfloat16 else _ops().cast(t, torch.bfloat16)


def _like(g: Tensor, ref: Tensor) -> Tensor:
    return g if g.dtype == ref.dtype else _ops().cast(g.contiguous(), ref.dtype)


def _t2d(x: Tensor) -> Tensor:
    """[R, C] -> contiguous [C, R] with the library's transpose kernel; R padded to a multiple of 8 with zeros
    so the result is a legal K-major GEMM operand."""
    ops = _ops()
    R, Cc = x.shape
    xt = ops.transpose2d(x.contiguous())
    if R % 8:
        xt = torch.cat([xt, xt.new_zeros(Cc, 8 - R % 8)], dim=1).contiguous()
    return xt


def colsum(x2d: Tensor, groups: int = 1) -> Tensor:
    ops = _ops()
    rows, cols = x2d.shape
    out = torch.zeros(groups, cols, device=x2d.device, dtype=torch.float32)
    check(_lib.lib().sdb_colsum(ops._ptr(x2d), ops._dt(x2d), x2d.stride(0), groups, rows // groups, cols,
                                ops._ptr(out), ops._stream()), "colsum")
    return out


# ------------------------------------------------------------------------------------------ GEMM
class _Gemm(Function):
    @staticmethod
    def forward(ctx, a, w, bias, residual, a2, out_dtype):
        ops = _ops()
        out = ops.gemm(a, w, bias, residual, ACT_NONE, a2, out_dtype)
        ctx.save_for_backward(a, w, a2)
        ctx.has_bias, ctx.res_dtype = bias is not None, (residual.dtype if residual is not None else None)
        return out

    @staticmethod
    def backward(ctx, dy):
        ops = _ops()
        a, w, a2 = ctx.saved_tensors
        dyb = _bf16(dy)
        K1 = a.shape[1]
        need = ctx.needs_input_grad
        da = dw = db = dres = da2 = None
        if need[0]:
            da = ops.gemm(dyb, w[:, :K1].t().contiguous())
        if a2 is not None and need[4]:
            da2 = ops.gemm(dyb, w[:, K1:].t().contiguous())
        if need[1]:
            # tcgen05 weight-gradient kernel: dy^T . a straight from the row-major operands (split over the rows)
            dw32 = torch.zeros(w.shape[0], w.shape[1], device=w.device, dtype=torch.float32)
            done = ops.wgrad_linear(dyb, a, out=dw32[:, :K1]) is not None
            if done and a2 is not None:
                done = ops.wgrad_linear(dyb, a2, out=dw32[:, K1:]) is not None
            if done:
                dw = dw32.to(w.dtype)
            else:   # shapes outside that kernel: K-major copies + the forward GEMM kernel
                dyt = _t2d(dyb)
                parts = [ops.gemm(dyt, _t2d(a), out_dtype=torch.float32)]
                if a2 is not None:
                    parts.append(ops.gemm(dyt, _t2d(a2), out_dtype=torch.float32))
                dw = (parts[0] if len(parts) == 1 else torch.cat(parts, dim=1)).to(w.dtype)
        if ctx.has_bias and need[2]:
            db = colsum(dyb)[0]
        if ctx.res_dtype is not None and need[3]:
            dres = dy if dy.dtype == ctx.res_dtype else ops.cast(dy.contiguous(), ctx.res_dtype)
        return da, dw, db, dres, da2, None


class _Geglu(Function):
    @staticmethod
    def forward(ctx, h):
        ops = _ops()
        M, F2 = h.shape
        out = torch.empty(M, F2 // 2, device=h.device, dtype=h.dtype)
        check(_lib.lib().sdb_geglu_fwd(ops._ptr(h), ops._ptr(out), M, F2 // 2, ops._stream()), "geglu_fwd")
        ctx.save_for_backward(h)
        return out

    @staticmethod
    def backward(ctx, dout):
        ops = _ops()
        (h,) = ctx.saved_tensors
        dh = torch.empty_like(h)
        check(_lib.lib().sdb_geglu_bwd(ops._ptr(h), ops._ptr(_bf16(dout)), ops._ptr(dh), h.shape[0], h.shape[1] // 2,
                                       ops._stream()), "geglu_bwd")
        return dh


class _Silu(Function):
    @staticmethod
    def forward(ctx, x, out_dtype):
        ctx.save_for_backward(x)
        return _ops().silu(x, out_dtype)

    @staticmethod
    def backward(ctx, dy):
        ops = _ops()
        (x,) = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        check(_lib.lib().sdb_silu_bwd(ops._ptr(x), ops._dt(x), ops._ptr(dy), ops._ptr(dx), ops._dt(dy), x.numel(),
                                      ops._stream()), "silu_bwd")
        return _like(dx, x), None


class _Dropout(Function):
    @staticmethod
    def forward(ctx, x, p, site):
        ctx.p, ctx.site = p, site
        return _ops().dropout_raw(x, p, site)

    @staticmethod
    def backward(ctx, dy):
        return _ops().dropout_raw(dy, ctx.p, ctx.site), None, None      # same (state, site) -> same mask


def dropout(x, p, site):
    return _Dropout.apply(x, p, site)


class _Cast(Function):
    @staticmethod
    def forward(ctx, x, dtype):
        ctx.src = x.dtype
        return _ops().cast(x, dtype)

    @staticmethod
    def backward(ctx, dy):
        return _ops().cast(dy.contiguous(), ctx.src), None


def gemm(a, w, bias=None, residual=None, act=ACT_NONE, a2=None, out_dtype=None):
    if act == ACT_NONE:
        return _Gemm.apply(a, w, bias, residual, a2, out_dtype)
    assert residual is None
    h = _Gemm.apply(a, w, bias, None, a2, None)          # pre-activation kept for the backward
    if act == ACT_GEGLU:
        o = _Geglu.apply(h)
        return o if out_dtype in (None, o.dtype) else _Cast.apply(o, out_dtype)
    if act == ACT_SILU:
        return _Silu.apply(h, out_dtype)
    raise ValueError(f"unknown activation {act}")


# ------------------------------------------------------------------------------------------ conv 3x3
class _Conv3x3(Function):
    @staticmethod
    def forward(ctx, x, w, bias, temb, residual, down2, out_dtype):
        out = _ops().conv3x3(x, w, bias, temb, residual, down2=down2, out_dtype=out_dtype)
        ctx.save_for_backward(x, w)
        ctx.down2 = down2
        ctx.has_bias, ctx.has_temb = bias is not None, temb is not None
        ctx.res_dtype = residual.dtype if residual is not None else None
        return out

    @staticmethod
    def backward(ctx, dy):
        ops = _ops()
        x, w = ctx.saved_tensors
        B, H, W, Cin = x.shape
        Cout = w.shape[0]
        need = ctx.needs_input_grad
        dyb = _bf16(dy)                                   # [B, Ho, Wo, Cout]
        Ho, Wo = dyb.shape[1], dyb.shape[2]
        dyf = dyb
        if ctx.down2:                                     # gradient of conv-then-nearest-x0.5 sits on the even pixels
            dyf = torch.empty(B, H, W, Cout, device=x.device, dtype=torch.bfloat16)
            check(_lib.lib().sdb_scatter_even(ops._ptr(dyb), ops._ptr(dyf), ops._dt(dyb), B, H, W, Cout, ops._stream()),
                  "scatter_even")
        dx = dw = db = dtemb = dres = None
        if need[0]:
            w_flip = w.flip(1, 2).permute(3, 1, 2, 0).contiguous()          # [Cin, 3, 3, Cout]
            dx = ops.conv3x3(dyf, w_flip, None)
        if need[1]:
            dw32 = ops.wgrad_conv3x3(dyf, x)              # tcgen05 implicit weight-gradient GEMM (no im2col)
            if dw32 is None:                              # geometry outside that kernel: explicit im2col + GEMM
                M = B * H * W
                col = torch.empty(M, 9 * Cin, device=x.device, dtype=torch.bfloat16)
                check(_lib.lib().sdb_im2col3x3(ops._ptr(x), ops._ptr(col), B, H, W, Cin, 9 * Cin, ops._stream()), "im2col3x3")
                dw32 = ops.gemm(_t2d(dyf.view(M, Cout)), _t2d(col), out_dtype=torch.float32).view(Cout, 3, 3, Cin)
            dw = dw32.to(w.dtype)
        if ctx.has_bias and need[2]:
            db = colsum(dyb.view(B * Ho * Wo, Cout))[0]
        if ctx.has_temb and need[3]:
            dtemb = colsum(dyb.view(B * Ho * Wo, Cout), groups=B)
        if ctx.res_dtype is not None and need[4]:
            dres = dy if dy.dtype == ctx.res_dtype else ops.cast(dy.contiguous(), ctx.res_dtype)
        return dx, dw, db, dtemb, dres, None, None


def conv3x3(x, w, bias, temb=None, residual=None, down2=False, out_dtype=None):
    return _Conv3x3.apply(x, w, bias, temb, residual, down2, out_dtype)


class _ConvIn(Function):
    """conv_in (models/utils.py:31-32): data gradient is never needed (x is the noised latent)."""

    @staticmethod
    def forward(ctx, x_nchw, w, bias, out_dtype, repeat):
        ctx.save_for_backward(x_nchw, w)
        ctx.repeat = repeat
        return _ops().conv3x3_in(x_nchw, w, bias, out_dtype, repeat)

    @staticmethod
    def backward(ctx, dy):
        ops = _ops()
        x, w = ctx.saved_tensors
        Bs, Cin, H, W = x.shape
        Cout = w.shape[0]
        dyb = _bf16(dy)                                   # [Bs*repeat, H, W, Cout]
        xp = torch.empty(Bs, H, W, 16, device=x.device, dtype=torch.bfloat16)
        check(_lib.lib().sdb_nchw_to_nhwc_pad(ops._ptr(x.contiguous()), ops._dt(x), ops._ptr(xp), ops._dt(xp), Bs, Cin, 16,
                                              H, W, ops._stream()), "nchw_to_nhwc_pad")
        M = Bs * H * W
        col = torch.empty(M, 144, device=x.device, dtype=torch.bfloat16)
        check(_lib.lib().sdb_im2col3x3(ops._ptr(xp), ops._ptr(col), Bs, H, W, 16, 144, ops._stream()), "im2col3x3")
        colt = _t2d(col)
        dw = None
        for r in range(ctx.repeat):                      # the virtual batch repeat reads every latent `repeat` times
            part = ops.gemm(_t2d(dyb[r * Bs:(r + 1) * Bs].reshape(M, Cout)), colt, out_dtype=torch.float32)
            dw = part if dw is None else dw + part
        dw = dw.view(Cout, 3, 3, 16)[..., :Cin].permute(0, 3, 1, 2).contiguous().to(w.dtype)   # -> OIHW
        db = colsum(dyb.view(-1, Cout))[0]
        return None, dw, db, None, None


def conv3x3_in(x_nchw, w, bias, out_dtype=torch.bfloat16, repeat=1):
    return _ConvIn.apply(x_nchw, w, bias, out_dtype, repeat)


class _SliceToNCHW(Function):
    @staticmethod
    def forward(ctx, y, cout, out_dtype):
        ops = _ops()
        B, H, W, Cp = y.shape
        ctx.cp = Cp
        out = torch.empty(B, cout, H, W, device=y.device, dtype=out_dtype)
        check(_lib.lib().sdb_nhwc_slice_to_nchw(ops._ptr(y), Cp, ops._ptr(out), ops._dt(out), B, cout, H, W, ops._stream()),
              "nhwc_slice_to_nchw")
        return out

    @staticmethod
    def backward(ctx, dout):
        ops = _ops()
        dout = dout.contiguous()
        B, C, H, W = dout.shape
        dy = torch.empty(B, H, W, ctx.cp, device=dout.device, dtype=torch.float32)
        check(_lib.lib().sdb_nchw_to_nhwc_pad(ops._ptr(dout), ops._dt(dout), ops._ptr(dy), ops._dt(dy), B, C, ctx.cp, H, W,
                                              ops._stream()), "nchw_to_nhwc_pad")
        return dy, None, None


def conv3x3_head(x_nhwc, w16, bias16, cout, out_dtype=torch.float32):
    y = _Conv3x3.apply(x_nhwc, w16, bias16, None, None, False, torch.float32)
    return _SliceToNCHW.apply(y, cout, out_dtype)


# ------------------------------------------------------------------------------------------ norms
class _GroupNorm(Function):
    @staticmethod
    def forward(ctx, x1, x2, gamma, beta, groups, act, out_dtype, eps):
        ops = _ops()
        B, H, W, C1 = x1.shape
        C2 = x2.shape[-1] if x2 is not None else 0
        hw = H * W
        L = _lib.lib()
        chunks = L.sdb_groupnorm_chunks(hw)
        partial = torch.empty(B * chunks * (C1 + C2) * 2, device=x1.device, dtype=torch.float32)
        stats = torch.empty(B * groups * 2, device=x1.device, dtype=torch.float32)
        out = torch.empty(B, H, W, C1 + C2, device=x1.device, dtype=out_dtype or x1.dtype)
        check(L.sdb_groupnorm_stats(ops._ptr(x1), C1, ops._ptr(x2), C2, ops._dt(x1), B, hw, groups, eps, ops._ptr(partial),
                                    ops._ptr(stats), ops._stream()), "groupnorm_stats")
        check(L.sdb_groupnorm_apply(ops._ptr(x1), C1, ops._ptr(x2), C2, ops._dt(x1), ops._ptr(stats), ops._ptr(gamma),
                                    ops._ptr(beta), act, ops._ptr(out), ops._dt(out), B, hw, groups, ops._stream()),
              "groupnorm_apply")
        ctx.save_for_backward(x1, x2, gamma, beta, stats)
        ctx.groups, ctx.act = groups, act
        return out

    @staticmethod
    def backward(ctx, dy):
        ops = _ops()
        x1, x2, gamma, beta, stats = ctx.saved_tensors
        B, H, W, C1 = x1.shape
        C2 = x2.shape[-1] if x2 is not None else 0
        dyc = _like(dy.contiguous(), x1)
        dx1 = torch.empty_like(x1)
        dx2 = torch.empty_like(x2) if x2 is not None else None
        dg = torch.zeros(C1 + C2, device=x1.device, dtype=torch.float32)
        db = torch.zeros(C1 + C2, device=x1.device, dtype=torch.float32)
        sums = torch.empty(B * ctx.groups * 2, device=x1.device, dtype=torch.float32)
        colbuf = torch.zeros(B * 2 * (C1 + C2), device=x1.device, dtype=torch.float32)   # per-(image, channel) sums
        check(_lib.lib().sdb_groupnorm_bwd_vec(ops._ptr(x1), C1, ops._ptr(x2), C2, ops._dt(x1), ops._ptr(stats),
                                               ops._ptr(gamma), ops._ptr(beta), ctx.act, ops._ptr(dyc), ops._ptr(dx1),
                                               ops._ptr(dx2), ops._ptr(dg), ops._ptr(db), ops._ptr(sums), ops._ptr(colbuf),
                                               B, H * W, ctx.groups, ops._stream()), "groupnorm_bwd")
        return dx1, dx2, dg.to(gamma.dtype), db.to(beta.dtype), None, None, None, None


def group_norm(x1, gamma, beta, groups, act=ACT_NONE, x2=None, out_dtype=None, eps=1e-5):
    return _GroupNorm.apply(x1, x2, gamma, beta, groups, act, out_dtype, eps)


class _LayerNorm(Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        ctx.save_for_backward(x, gamma)
        ctx.eps = eps
        return _ops().add_layernorm(x, None, gamma, beta, eps)

    @staticmethod
    def backward(ctx, dy):
        ops = _ops()
        x, gamma = ctx.saved_tensors
        Cc = x.shape[-1]
        rows = x.numel() // Cc
        dyc = _like(dy.contiguous(), x)
        dx = torch.empty_like(x)
        dg = torch.zeros(Cc, device=x.device, dtype=torch.float32)
        db = torch.zeros(Cc, device=x.device, dtype=torch.float32)
        check(_lib.lib().sdb_layernorm_bwd(ops._ptr(x), ops._ptr(gamma), ops._ptr(dyc), ops._ptr(dx), ops._ptr(dg), ops._ptr(db),
                                           rows, Cc, ctx.eps, ops._dt(x), ops._stream()), "layernorm_bwd")
        return dx, dg.to(gamma.dtype), db, None


def add_layernorm(x, res, gamma, beta, eps=1e-5):
    if res is not None:
        raise NotImplementedError("training path: the residual is fused into the preceding GEMM, not into LayerNorm")
    return _LayerNorm.apply(x.contiguous(), gamma, beta, eps)


# ------------------------------------------------------------------------------------------ attention / resampling
class _AttentionP32(Function):
    @staticmethod
    def forward(ctx, q, k, v, heads, d, scale):
        # the tcgen05 forward keeps the row log-sum-exp, so the backward does not rebuild it
        o, lse = _ops().attention_p32_lse(q, k, v, heads, d, scale)
        ctx.has_lse = lse is not None
        ctx.save_for_backward(q, k, v, o, *([lse] if lse is not None else []))
        ctx.heads, ctx.d, ctx.scale = heads, d, scale
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o = ctx.saved_tensors[:4]
        B, n, m = q.shape[0], q.shape[1], k.shape[1]
        dq = torch.empty(B, n, ctx.heads * 32, device=q.device, dtype=torch.bfloat16)
        dk = torch.empty(B, m, ctx.heads * 32, device=q.device, dtype=torch.bfloat16)
        dv = torch.empty_like(dk)
        _attention_p32_backward(ctx, q, k, v, o, ctx.saved_tensors[4] if ctx.has_lse else None, do, dq, dk, dv)
        return dq, dk, dv, None, None, None


class _SelfAttentionP32(Function):
    """``_AttentionP32`` over the packed projection output qkv [B, n, 3*heads*32]: Q / K / V are column slices of one
    tensor, and the three backward kernels write their slices of ONE packed gradient (row stride 3*heads*32) -- as
    three autograd inputs every slice cost a full-size zero fill, a strided copy and an add per layer."""

    @staticmethod
    def forward(ctx, qkv, heads, d, scale):
        kw = heads * 32
        q, k, v = qkv[:, :, :kw], qkv[:, :, kw:2 * kw], qkv[:, :, 2 * kw:]
        o, lse = _ops().attention_p32_lse(q, k, v, heads, d, scale)
        ctx.has_lse = lse is not None
        ctx.save_for_backward(qkv, o, *([lse] if lse is not None else []))
        ctx.heads, ctx.d, ctx.scale = heads, d, scale
        return o

    @staticmethod
    def backward(ctx, do):
        qkv, o = ctx.saved_tensors[:2]
        kw = ctx.heads * 32
        # the tcgen05 kernels write every padded column of their slice; the mma.sync fall-backs leave the pad alone
        full = ctx.has_lse and ctx.d <= 20 and _DQ_UMMA and _DKV_UMMA and qkv.shape[1] % 4 == 0
        dqkv = torch.empty_like(qkv) if full else torch.zeros_like(qkv)
        _attention_p32_backward(ctx, qkv[:, :, :kw], qkv[:, :, kw:2 * kw], qkv[:, :, 2 * kw:], o,
                                ctx.saved_tensors[2] if ctx.has_lse else None, do, dqkv[:, :, :kw],
                                dqkv[:, :, kw:2 * kw], dqkv[:, :, 2 * kw:])
        return dqkv, None, None, None


def _attention_p32_backward(ctx, q, k, v, o, lse, do, dq, dk, dv):
    """Fills dq / dk / dv (head-padded, row-strided views allowed)."""
    ops = _ops()
    B, n, m = q.shape[0], q.shape[1], k.shape[1]
    hp = 32
    do = _bf16(do)
    if ctx.has_lse and ctx.d <= 20 and _DQ_UMMA:
        # long sequences (the forward ran on tcgen05 and kept lse): dQ and dK / dV on the tensor cores
        H, d = ctx.heads, ctx.d
        L = _lib.lib()
        do = do if do.stride(2) == 1 and do.stride(0) == n * do.stride(1) else do.contiguous()
        do_pad = torch.empty(B, n, H, hp, device=q.device, dtype=torch.bfloat16)
        stats = torch.empty(3, B, H, n, device=q.device, dtype=torch.float32)
        delta, ndelta, nlse = stats[0], stats[1], stats[2]
        check(L.sdb_attention_bwd_prep(ops._ptr(do), do.stride(1), ops._ptr(o), o.stride(1), ops._ptr(lse),
                                       ops._ptr(do_pad), ops._ptr(delta), ops._ptr(ndelta), ops._ptr(nlse), B, H, n, d,
                                       ops._stream()), "attention_bwd_prep")
        check(L.sdb_attention_bwd_dq_p32(ops._ptr(q), q.stride(1), ops._ptr(k), k.stride(1), ops._ptr(v), v.stride(1),
                                         ops._ptr(do_pad), H * hp, ops._ptr(lse), ops._ptr(delta), ops._ptr(dq),
                                         dq.stride(1), B, H, n, m, d, ctx.scale, ops._stream()), "attention_bwd_dq_p32")
        if n % 4 == 0 and _DKV_UMMA:
            check(L.sdb_attention_bwd_dkv_p32(ops._ptr(q), q.stride(1), ops._ptr(k), k.stride(1), ops._ptr(v),
                                              v.stride(1), ops._ptr(do_pad), H * hp, ops._ptr(nlse), ops._ptr(ndelta),
                                              ops._ptr(dk), dk.stride(1), ops._ptr(dv), dv.stride(1), B, H, n, m, d,
                                              ctx.scale, ops._stream()), "attention_bwd_dkv_p32")
        else:
            check(L.sdb_attention_bwd_dkv(ops._ptr(q), q.stride(1), ops._ptr(k), k.stride(1), ops._ptr(v), v.stride(1),
                                          ops._ptr(do), do.stride(1), ops._ptr(dk), dk.stride(1), ops._ptr(dv),
                                          dv.stride(1), ops._ptr(lse), ops._ptr(delta), B, H, n, m, d, hp, ctx.scale,
                                          ops._stream()), "attention_bwd_dkv")
        return
    delta = torch.empty(B * ctx.heads * n, device=q.device, dtype=torch.float32)
    lse = lse if ctx.has_lse else torch.empty_like(delta)
    if m <= 128 and ctx.d <= 20 and _CTX_BWD:
        # short context (cross-attention over CLIP's 77 tokens): one-pass dQ kernel that also leaves lse / delta for dK / dV
        L = _lib.lib()
        do = do if do.stride(2) == 1 and do.stride(0) == n * do.stride(1) else do.contiguous()
        check(L.sdb_attention_ctx_bwd_dq(ops._ptr(q), q.stride(1), ops._ptr(k), k.stride(1), ops._ptr(v), v.stride(1),
                                         ops._ptr(do), do.stride(1), ops._ptr(dq), dq.stride(1), ops._ptr(lse), ops._ptr(delta),
                                         B, ctx.heads, n, m, ctx.d, ctx.scale, ops._stream()), "attention_ctx_bwd_dq")
        check(L.sdb_attention_bwd_dkv(ops._ptr(q), q.stride(1), ops._ptr(k), k.stride(1), ops._ptr(v), v.stride(1),
                                      ops._ptr(do), do.stride(1), ops._ptr(dk), dk.stride(1), ops._ptr(dv), dv.stride(1),
                                      ops._ptr(lse), ops._ptr(delta), B, ctx.heads, n, m, ctx.d, hp, ctx.scale,
                                      ops._stream()), "attention_bwd_dkv")
        return
    check(_lib.lib().sdb_attention_bwd(
        ops._ptr(q), q.stride(1), ops._ptr(k), k.stride(1), ops._ptr(v), v.stride(1), ops._ptr(o), o.stride(1),
        ops._ptr(do), do.stride(1), ops._ptr(dq), dq.stride(1), ops._ptr(dk), dk.stride(1), ops._ptr(dv), dv.stride(1),
        ops._ptr(lse), ops._ptr(delta), int(ctx.has_lse), B, ctx.heads, n, m, ctx.d, hp, ctx.scale, ops._stream()),
        "attention_bwd")


class _Attention(Function):
    """Dense-head layout (``ops.attention``): head h at columns [h*d, (h+1)*d) of q / k / v."""

    @staticmethod
    def forward(ctx, q, k, v, heads, d, scale):
        o = _ops().attention(q, k, v, heads, d, scale)
        ctx.save_for_backward(q, k, v, o)
        ctx.heads, ctx.d, ctx.scale = heads, d, scale
        return o

    @staticmethod
    def backward(ctx, do):
        ops = _ops()
        q, k, v, o = ctx.saved_tensors
        B, n, m = q.shape[0], q.shape[1], k.shape[1]
        do = _bf16(do)
        dq = torch.empty(B, n, ctx.heads * ctx.d, device=q.device, dtype=torch.bfloat16)
        dk = torch.empty(B, m, ctx.heads * ctx.d, device=q.device, dtype=torch.bfloat16)
        dv = torch.empty_like(dk)
        scratch = torch.empty(2, B * ctx.heads * n, device=q.device, dtype=torch.float32)
        check(_lib.lib().sdb_attention_bwd(
            ops._ptr(q), q.stride(1), ops._ptr(k), k.stride(1), ops._ptr(v), v.stride(1), ops._ptr(o), o.stride(1),
            ops._ptr(do), do.stride(1), ops._ptr(dq), dq.stride(1), ops._ptr(dk), dk.stride(1), ops._ptr(dv), dv.stride(1),
            ops._ptr(scratch[0]), ops._ptr(scratch[1]), 0, B, ctx.heads, n, m, ctx.d, ctx.d, ctx.scale, ops._stream()),
            "attention_bwd")
        return dq, dk, dv, None, None, None


def attention(q, k, v, heads, d, scale):
    if d > 32:
        raise NotImplementedError(f"attention backward: head dim {d} > 32 is not built (the repo-default UNet trains "
                                  f"with d_head 20)")
    return _Attention.apply(q, k, v, heads, d, scale)


def attention_p32(q, k, v, heads, d, scale):
    return _AttentionP32.apply(q, k, v, heads, d, scale)


def attention_p32_packed(qkv, heads, d, scale):
    return _SelfAttentionP32.apply(qkv, heads, d, scale)


class _Upsample2x(Function):
    @staticmethod
    def forward(ctx, x):
        return _ops().upsample2x(x)

    @staticmethod
    def backward(ctx, dy):
        ops = _ops()
        dy = dy.contiguous()
        B, H2, W2, Cc = dy.shape
        dx = torch.empty(B, H2 // 2, W2 // 2, Cc, device=dy.device, dtype=dy.dtype)
        check(_lib.lib().sdb_upsample2x_bwd(ops._ptr(dy), ops._ptr(dx), ops._dt(dy), B, H2 // 2, W2 // 2, Cc, ops._stream()),
              "upsample2x_bwd")
        return dx


def upsample2x(x):
    return _Upsample2x.apply(x)


def silu(x, out_dtype=None):
    return _Silu.apply(x, out_dtype)


def cast(x, dtype):
    return _Cast.apply(x, dtype)


# ------------------------------------------------------------------------------------------ loss / guidance
class _CfgCombine(Function):
    @staticmethod
    def forward(ctx, eps2, g):
        ctx.g = g
        return _ops().cfg_combine(eps2, g)

    @staticmethod
    def backward(ctx, d):
        ops = _ops()
        d = d.contiguous()
        de2 = torch.empty((2 * d.shape[0], *d.shape[1:]), device=d.device, dtype=d.dtype)
        check(_lib.lib().sdb_cfg_combine_bwd(ops._ptr(d), ops._ptr(de2), ops._dt(d), d.numel(), ctx.g, ops._stream()),
              "cfg_combine_bwd")
        return de2, None


def cfg_combine(eps2, guidance):
    return _CfgCombine.apply(eps2.contiguous(), guidance)


class _Mse(Function):
    """F.mse_loss(pred.float(), target.float(), reduction='mean') (train_unet.py:517), gradient in the same pass."""

    @staticmethod
    def forward(ctx, pred, target):
        ops = _ops()
        pred, target = pred.contiguous(), target.contiguous()
        loss = torch.zeros(1, device=pred.device, dtype=torch.float32)
        dpred = torch.empty_like(pred)
        check(_lib.lib().sdb_mse(ops._ptr(pred), ops._ptr(target), ops._dt(pred), pred.numel(), 1.0, ops._ptr(loss),
                                 ops._ptr(dpred), ops._stream()), "mse")
        ctx.save_for_backward(dpred)
        return loss[0]

    @staticmethod
    def backward(ctx, dloss):
        (dpred,) = ctx.saved_tensors
        return dpred * dloss.to(dpred.dtype), None


def mse_loss(pred: Tensor, target: Tensor) -> Tensor:
    assert pred.shape == target.shape
    ops = _ops()
    if pred.dtype != torch.float32:                       # .float() of train_unet.py:527-528
        pred = ops.cast(pred, torch.float32)
    if target.dtype != torch.float32:
        target = ops.cast(target, torch.float32)
    return _Mse.apply(pred, target)
