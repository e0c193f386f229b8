"""Thin torch-tensor front end of the C ABI: torch supplies device memory and the current stream,
every computation happens in ``libsdb200.so``.  All tensors must be CUDA tensors (no CPU path)."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from . import _lib
from . import autograd_ops as _ag
from ._lib import ACT_GEGLU, ACT_NONE, ACT_SILU, SDB_BF16, SDB_F32, SDB_U8, SdbConvArgs, SdbGemmArgs, check

Tensor = torch.Tensor


def _dt(t_or_dtype) -> int:
    d = t_or_dtype.dtype if isinstance(t_or_dtype, torch.Tensor) else t_or_dtype
    if d == torch.float32:
        return SDB_F32
    if d == torch.bfloat16:
        return SDB_BF16
    if d == torch.uint8:
        return SDB_U8
    raise TypeError(f"unsupported dtype {d}: the B200 path computes in float32 or bfloat16")


def _ptr(t: Optional[Tensor]) -> Optional[int]:
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("stable_diffusion_pytorch_b200 has no CPU fallback: tensor is not on a CUDA device")
    return t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _no_f32_grad(x: Tensor) -> None:
    if x.dtype == torch.float32:
        raise NotImplementedError("the fp32 parity mode is inference-only: backward kernels exist for the bf16 path")


def timestep_embed(t: Tensor, freq: Tensor, out_dtype=torch.float32) -> Tensor:
    assert t.dtype == torch.int64 and freq.dtype == torch.float32
    B, half = t.shape[0], freq.shape[0]
    out = torch.empty(B, 2 * half, device=t.device, dtype=out_dtype)
    check(_lib.lib().sdb_timestep_embed(_ptr(t), _ptr(freq), _ptr(out), B, half, _dt(out_dtype), _stream()),
          "timestep_embed")
    return out


def cfg_sampler_step(eps: Tensor, x_t: Tensor, noise: Optional[Tensor], coef: Tensor, slot: int = 0,
                     slot_dev: Optional[Tensor] = None, kind: int = 0, cfg_mode: int = 1,
                     guidance: float = 7.5, scale_factor: float = 1.0, seed: int = 0,
                     x_prev: Optional[Tensor] = None, x0: Optional[Tensor] = None, want_x0: bool = True):
    B = x_t.shape[0]
    n_per = x_t.numel() // B
    assert eps.is_contiguous() and x_t.is_contiguous() and eps.dtype == x_t.dtype
    assert eps.shape[0] == (2 * B if cfg_mode else B)
    if noise is not None:
        assert noise.is_contiguous() and noise.dtype == x_t.dtype
        # device-slot mode: noise is the per-slot table [n_slots, B, ...]
        assert noise.numel() == x_t.numel() or (slot_dev is not None and noise.numel() % x_t.numel() == 0)
    if x_prev is None:
        x_prev = torch.empty_like(x_t)
    if x0 is None and want_x0:
        x0 = torch.empty_like(x_t)
    check(_lib.lib().sdb_cfg_sampler_step(_ptr(eps), _ptr(x_t), _ptr(noise), _ptr(coef), _ptr(slot_dev), slot,
                                          kind, cfg_mode, guidance, scale_factor, seed, _ptr(x_prev), _ptr(x0),
                                          B, n_per, _dt(x_t), _stream()), "cfg_sampler_step")
    return x_prev, x0


def sampler_advance(slot_dev: Tensor, t_table: Tensor, t_out: Tensor) -> None:
    check(_lib.lib().sdb_sampler_advance(_ptr(slot_dev), _ptr(t_table), t_table.numel(), _ptr(t_out),
                                         t_out.numel(), _stream()), "sampler_advance")


def add_noise(x0: Tensor, noise: Tensor, t: Tensor, alphas_cumprod: Tensor) -> Tensor:
    _ag.refuse("add_noise", x0, noise)
    B = x0.shape[0]
    out = torch.empty_like(x0)
    check(_lib.lib().sdb_add_noise(_ptr(x0.contiguous()), _ptr(noise.contiguous()), _ptr(t), _ptr(alphas_cumprod),
                                   alphas_cumprod.numel(), _ptr(out), B, x0.numel() // B, _dt(x0), _stream()),
          "add_noise")
    return out


def nchw_to_nhwc(x: Tensor, out_dtype=None) -> Tensor:
    _ag.refuse("nchw_to_nhwc", x)
    B, Cc, H, W = x.shape
    out = torch.empty(B, H, W, Cc, device=x.device, dtype=out_dtype or x.dtype)
    check(_lib.lib().sdb_nchw_to_nhwc(_ptr(x.contiguous()), _dt(x), _ptr(out), _dt(out), B, Cc, H, W, _stream()),
          "nchw_to_nhwc")
    return out


def nhwc_to_nchw(x: Tensor, out_dtype=None) -> Tensor:
    _ag.refuse("nhwc_to_nchw", x)
    B, H, W, Cc = x.shape
    out = torch.empty(B, Cc, H, W, device=x.device, dtype=out_dtype or x.dtype)
    check(_lib.lib().sdb_nhwc_to_nchw(_ptr(x.contiguous()), _dt(x), _ptr(out), _dt(out), B, Cc, H, W, _stream()),
          "nhwc_to_nchw")
    return out


def transpose2d(x: Tensor, out_dtype=None) -> Tensor:
    """[R, C] -> [C, R] (the batch-1 case of the layout kernel); feeds the weight-gradient GEMMs."""
    R, Cc = x.shape
    out = torch.empty(Cc, R, device=x.device, dtype=out_dtype or x.dtype)
    check(_lib.lib().sdb_nchw_to_nhwc(_ptr(x.contiguous()), _dt(x), _ptr(out), _dt(out), 1, R, Cc, 1, _stream()),
          "transpose2d")
    return out


class GroupNormScratch:
    """Caller-owned scratch of the two-pass GroupNorm (grown on demand, reused across calls)."""

    def __init__(self):
        self.partial = None
        self.stats = None

    def get(self, device, B, hw, Cc, groups):
        chunks = _lib.lib().sdb_groupnorm_chunks(hw)
        need_p, need_s = B * chunks * Cc * 2, B * groups * 2
        if self.partial is None or self.partial.numel() < need_p or self.partial.device != device:
            self.partial = torch.empty(need_p, device=device, dtype=torch.float32)
        if self.stats is None or self.stats.numel() < need_s or self.stats.device != device:
            self.stats = torch.empty(need_s, device=device, dtype=torch.float32)
        return self.partial, self.stats


_default_gn_scratch = GroupNormScratch()
# A/B switch: "0" = always the three-kernel route, "cluster" = the cluster-resident single-pass kernel where it fits,
# anything else = the one-launch cooperative kernel
_GN_FUSED = {"0": False, "cluster": "cluster"}.get(os.environ.get("SDB_GN_FUSED", "1"), "coop")
_LN_FUSED = os.environ.get("SDB_LN_FUSED", "1") != "0"    # A/B switch: 0 = LayerNorm always as its own launch
_gn_coop_state = {}


def _gn_coop_scratch(device, n_floats: int):
    """Per-device scratch of the cooperative GroupNorm: partial sums (grown on demand) and the two barrier words
    (zeroed once; the kernel leaves the counter at zero).  Calls on one stream run in order, so they can share it."""
    key = str(device)
    st = _gn_coop_state.get(key)
    if st is None or st[0].numel() < n_floats:
        bar = st[1] if st is not None else torch.zeros(2, device=device, dtype=torch.int32)
        st = (torch.empty(max(n_floats, 1 << 16), device=device, dtype=torch.float32), bar)
        _gn_coop_state[key] = st
    return st


# ---- GroupNorm statistics from the producer's epilogue ---------------------------------------------------------
# Inside a ``gn_sums_scope`` (one model forward), conv3x3 / gemm calls marked ``sums=`` also accumulate per-channel
# (sum, sum of squares) of their output -- 64-bit fixed point, so exact and order-independent -- into a slice of a
# zeroed per-device arena and tag the output tensor with it; ``group_norm`` then skips its statistics pass
# (sdb_groupnorm_apply_sums).  Outside a scope, for shapes the epilogue cannot cover, under autograd or in fp32 mode
# nothing is tagged and group_norm takes its usual route.  Only tensors that do not stay in L2 between the two passes
# of the cooperative kernel are worth it (measured: config 3 VAE decode 39.7 -> 37.1 ms; the UNet's 42 MB level-0
# tensors re-read from L2 for free, and the heavier epilogue costs what the saved pass gives).
_GN_SUMS = os.environ.get("SDB_GN_SUMS", "1") != "0"      # A/B switch
_GN_SUMS_MIN_BYTES = int(os.environ.get("SDB_GN_SUMS_MIN_MB", "64")) << 20   # output tensors smaller than this: no sums
_SUMS_CAPACITY = 1 << 21                                  # int64 slots (16 MB)
SUM_SCALE, SQ_SCALE = 2.0 ** 30, 2.0 ** 24                # csrc/common.cuh kChanSumScale / kChanSqScale


class _SumsArena:
    def __init__(self, device):
        self.buf = torch.zeros(_SUMS_CAPACITY, device=device, dtype=torch.int64)
        self.off = 0          # bump pointer (slots) of the current scope
        self.dirty = 0        # slots handed out since the last zeroing
        self.depth = 0
        self.epoch = 0

    def alloc(self, n: int) -> Optional[Tensor]:
        n = (n + 1) // 2 * 2                               # 16-byte aligned slices
        if self.depth == 0 or self.off + n > self.buf.numel():
            return None
        v = self.buf[self.off:self.off + n]
        self.off += n
        self.dirty = max(self.dirty, self.off)
        return v


_sums_arenas = {}
_sums_unsupported = set()
sums_counters = {"tagged": 0, "consumed": 0}             # producers that left sums / GroupNorms that used them


class gn_sums_scope:
    """``with ops.gn_sums_scope(device):`` around one forward pass (re-entrant: an inner scope joins the outer one)."""

    def __init__(self, device):
        self.arena = None
        if _GN_SUMS and torch.device(device).type == "cuda":
            key = str(device)
            if key not in _sums_arenas and not torch.cuda.is_current_stream_capturing():
                _sums_arenas[key] = _SumsArena(device)
            self.arena = _sums_arenas.get(key)

    def __enter__(self):
        a = self.arena
        if a is not None:
            if a.depth == 0:
                if a.dirty:
                    a.buf[:a.dirty].zero_()
                a.off = a.dirty = 0
                a.epoch += 1
            a.depth += 1
        return self

    def __exit__(self, *exc):
        if self.arena is not None:
            self.arena.depth -= 1
        return False


def _sums_new(device, n: int, out_bytes: int):
    a = _sums_arenas.get(str(device))
    if a is None or out_bytes < _GN_SUMS_MIN_BYTES:
        return None, None
    return a.alloc(n), a


def _sums_of(x: Optional[Tensor]) -> Optional[Tensor]:
    """The channel sums a producer attached to ``x`` during the current scope, or None."""
    tag = getattr(x, "_gn_sums", None) if x is not None else None
    if tag is None:
        return None
    a = _sums_arenas.get(str(x.device))
    return tag[1] if a is not None and a.depth > 0 and tag[0] == a.epoch else None


def carry_sums(src: Tensor, dst: Tensor) -> Tensor:
    """``dst`` is a view / reshape of ``src`` with the same (image, channel) layout: keep the tag."""
    tag = getattr(src, "_gn_sums", None)
    if tag is not None:
        dst._gn_sums = tag
    return dst


def group_norm(x1: Tensor, gamma: Tensor, beta: Tensor, groups: int, act: int = ACT_NONE,
               x2: Optional[Tensor] = None, out_dtype=None, eps: float = 1e-5,
               scratch: Optional[GroupNormScratch] = None, out: Optional[Tensor] = None) -> Tensor:
    """GroupNorm(+SiLU) of the virtual channel concat [x1, x2]; NHWC in, dense NHWC out."""
    if _ag.wants(x1, x2, gamma, beta):
        return _ag.group_norm(x1, gamma, beta, groups, act, x2, out_dtype, eps)
    B, H, W, C1 = x1.shape
    C2 = x2.shape[-1] if x2 is not None else 0
    assert x1.is_contiguous() and (x2 is None or (x2.is_contiguous() and x2.dtype == x1.dtype))
    hw = H * W
    if out is None:
        out = torch.empty(B, H, W, C1 + C2, device=x1.device, dtype=out_dtype or x1.dtype)
    L = _lib.lib()
    s1, s2 = _sums_of(x1), _sums_of(x2)
    if s1 is not None and (x2 is None or s2 is not None) and groups <= 128:
        # the producers' epilogues already summed every channel: normalise only (one read + one write)
        check(L.sdb_groupnorm_apply_sums(_ptr(x1), C1, _ptr(s1), _ptr(x2), C2, _ptr(s2), _dt(x1), _ptr(gamma), _ptr(beta),
                                         eps, act, _ptr(out), _dt(out), B, hw, groups, _stream()), "groupnorm_apply_sums")
        sums_counters["consumed"] += 1
        return out
    if _GN_FUSED and x1.dtype == torch.bfloat16 and out.dtype == torch.bfloat16:
        if _GN_FUSED == "cluster":
            # single pass (cluster per image, image resident in shared memory); -3 = the image does not fit a cluster
            rc = L.sdb_groupnorm_fused(_ptr(x1), C1, _ptr(x2), C2, _ptr(gamma), _ptr(beta), act, _ptr(out), B, hw,
                                       groups, eps, _stream())
            if rc == 0:
                return out
            if rc != -3:
                check(rc, "groupnorm_fused")
        else:
            # one cooperative launch: statistics pass, grid barrier, normalise pass (second read out of L2)
            part, bar = _gn_coop_scratch(x1.device, L.sdb_groupnorm_coop_scratch(B, hw, groups))
            check(L.sdb_groupnorm_coop(_ptr(x1), C1, _ptr(x2), C2, _ptr(gamma), _ptr(beta), act, _ptr(out), B, hw,
                                       groups, eps, _ptr(part), _ptr(bar), _stream()), "groupnorm_coop")
            return out
    if scratch is not None:
        partial, stats = scratch.get(x1.device, B, hw, C1 + C2, groups)
    else:  # per-call scratch from torch's caching allocator: safe under CUDA-graph capture
        chunks = L.sdb_groupnorm_chunks(hw)
        partial = torch.empty(B * chunks * (C1 + C2) * 2, device=x1.device, dtype=torch.float32)
        stats = torch.empty(B * groups * 2, device=x1.device, dtype=torch.float32)
    check(L.sdb_groupnorm_stats(_ptr(x1), C1, _ptr(x2), C2, _dt(x1), B, hw, groups, eps, _ptr(partial),
                                _ptr(stats), _stream()), "groupnorm_stats")
    check(L.sdb_groupnorm_apply(_ptr(x1), C1, _ptr(x2), C2, _dt(x1), _ptr(stats), _ptr(gamma), _ptr(beta), act,
                                _ptr(out), _dt(out), B, hw, groups, _stream()), "groupnorm_apply")
    return out


def add_layernorm(x: Tensor, res: Optional[Tensor], gamma: Tensor, beta: Tensor, eps: float = 1e-5,
                  out: Optional[Tensor] = None) -> Tensor:
    Cc = x.shape[-1]
    if _ag.wants(x, res, gamma, beta):
        return _ag.add_layernorm(x, res, gamma, beta, eps)
    rows = x.numel() // Cc
    assert x.is_contiguous() and (res is None or (res.is_contiguous() and res.dtype == x.dtype))
    if out is None:
        out = torch.empty_like(x)
    check(_lib.lib().sdb_add_layernorm(_ptr(x), _ptr(res), _ptr(gamma), _ptr(beta), _ptr(out), rows, Cc, eps,
                                       _dt(x), _stream()), "add_layernorm")
    return out


def gemm(a: Tensor, w: Tensor, bias: Optional[Tensor] = None, residual: Optional[Tensor] = None,
         act: int = ACT_NONE, a2: Optional[Tensor] = None, out_dtype=None, out: Optional[Tensor] = None,
         ln=None, sums_rows: int = 0) -> Tensor:
    """out[M, N] = act(a[M, K] @ w[N, K]^T + bias) + residual; ``a``/``a2`` may be row-strided 2-D views.
    ``ln`` = (gamma, beta) fp32 [N]: LayerNorm (eps 1e-5) of every output row, in the GEMM epilogue when the row fits
    one tile (bf16, N <= 256), as a separate ``add_layernorm`` launch otherwise.
    ``sums_rows`` = rows per image: the output feeds a GroupNorm -- inside a ``gn_sums_scope`` the epilogue also
    accumulates its channel sums (see ``group_norm``)."""
    if _ag.wants(a, w, bias, residual, a2) or (ln is not None and _ag.wants(*ln)):
        _no_f32_grad(a)
        o = _ag.gemm(a, w, bias, residual, act, a2, out_dtype)
        return o if ln is None else add_layernorm(o, None, ln[0], ln[1])
    assert a.dim() == 2 and w.dim() == 2 and a.dtype in (torch.bfloat16, torch.float32) and w.dtype == a.dtype
    assert a.stride(1) == 1 and w.is_contiguous()
    M, K1 = a.shape
    N = w.shape[0]
    K2 = 0
    if a2 is not None:
        assert a2.dim() == 2 and a2.shape[0] == M and a2.stride(1) == 1 and a2.dtype == a.dtype
        K2 = a2.shape[1]
    assert w.shape[1] == K1 + K2, f"weight K {w.shape[1]} != {K1}+{K2}"
    n_out = N // 2 if act == ACT_GEGLU else N
    if out is None:
        out = torch.empty(M, n_out, device=a.device, dtype=out_dtype or a.dtype)
    assert out.stride(1) == 1 and out.shape == (M, n_out)
    args = SdbGemmArgs()
    args.a, args.lda, args.K1 = _ptr(a), a.stride(0), K1
    args.a2, args.lda2, args.K2 = (_ptr(a2), a2.stride(0), K2) if a2 is not None else (None, 0, 0)
    args.w, args.bias = _ptr(w), _ptr(bias)
    if residual is not None:
        assert residual.shape == (M, n_out) and residual.stride(1) == 1
        args.residual, args.ldr, args.res_dtype = _ptr(residual), residual.stride(0), _dt(residual)
    else:
        args.residual, args.ldr, args.res_dtype = None, 0, 0
    args.out, args.ldo, args.out_dtype = _ptr(out), out.stride(0), _dt(out)
    args.M, args.N, args.act = M, N, act
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == N
    fn = _lib.lib().sdb_gemm_f32 if a.dtype == torch.float32 else _lib.lib().sdb_gemm   # fp32 = parity mode
    if sums_rows and ln is None and a.dtype == torch.bfloat16 and out.dtype == torch.bfloat16 and act == ACT_NONE:
        key = ("gemm", M, N, K1 + K2, sums_rows, residual is not None)
        sums, arena = (None, None) if key in _sums_unsupported else _sums_new(a.device, (M // sums_rows) * N * 2, out.numel() * 2)
        if sums is not None:
            args.chan_sums, args.rows_per_image = _ptr(sums), sums_rows
            rc = fn(C.byref(args), _stream())
            if rc == 0:
                out._gn_sums = (arena.epoch, sums)
                sums_counters["tagged"] += 1
                return out
            if rc != -3:
                check(rc, "gemm")
            _sums_unsupported.add(key)
            args.chan_sums, args.rows_per_image = None, 0
    fuse_ln = (ln is not None and _LN_FUSED and a.dtype == torch.bfloat16 and out.dtype == torch.bfloat16
               and residual is not None and residual.dtype == torch.bfloat16 and act == ACT_NONE and N <= 256)
    if fuse_ln:
        args.ln_gamma, args.ln_beta, args.ln_eps = _ptr(ln[0]), _ptr(ln[1]), 1e-5
        rc = fn(C.byref(args), _stream())
        if rc == 0:
            return out
        if rc != -3:
            check(rc, "gemm")
        args.ln_gamma = args.ln_beta = None
    check(fn(C.byref(args), _stream()), "gemm")
    return out if ln is None else add_layernorm(out, None, ln[0], ln[1], out=out)


def conv3x3(x: Tensor, w: Tensor, bias: Optional[Tensor], temb: Optional[Tensor] = None,
            residual: Optional[Tensor] = None, down2: bool = False, out_dtype=None,
            out: Optional[Tensor] = None, up2: bool = False, sums: bool = False) -> Tensor:
    """3x3 / pad 1 convolution, NHWC bf16 ``x`` [B,H,W,Cin], ``w`` bf16 [Cout,3,3,Cin].
    ``up2``: nearest x2 upsampling folded in -- ``x`` is the low-resolution tensor, the result is [B,2H,2W,Cout] and
    ``w`` the phase-collapsed weights [4,Cout,2,2,Cin] of ``_pack.conv3x3_up_w`` (bf16 tensor-core path only).
    ``sums``: the output feeds a GroupNorm (see ``gemm(sums_rows=)``)."""
    if _ag.wants(x, w, bias, temb, residual):
        _no_f32_grad(x)
        assert not up2, "conv3x3(up2=True) is inference-only: under autograd use upsample2x + conv3x3"
        return _ag.conv3x3(x, w, bias, temb, residual, down2, out_dtype)
    B, H, W, Cin = x.shape
    Cout = w.shape[1] if up2 else w.shape[0]
    assert x.is_contiguous() and w.is_contiguous() and x.dtype in (torch.bfloat16, torch.float32) and w.dtype == x.dtype
    if up2:
        assert x.dtype == torch.bfloat16 and tuple(w.shape) == (4, Cout, 2, 2, Cin) and not down2 and temb is None
    else:
        assert tuple(w.shape) == (Cout, 3, 3, Cin)
    Ho, Wo = (H // 2, W // 2) if down2 else ((2 * H, 2 * W) if up2 else (H, W))
    if out is not None and out.dtype == torch.uint8:      # image head with the detransform in the epilogue
        assert x.dtype == torch.bfloat16 and Cout == 16 and tuple(out.shape[:3]) == (B, H, W) and out.is_contiguous()
        assert residual is None and temb is None and not down2 and not up2
    if out is None:
        out = torch.empty(B, Ho, Wo, Cout, device=x.device, dtype=out_dtype or x.dtype)
    a = SdbConvArgs()
    a.x, a.w, a.bias = _ptr(x), _ptr(w), _ptr(bias)
    if temb is not None:
        assert temb.dtype == torch.float32 and temb.stride(1) == 1 and temb.shape[0] == B
        a.temb, a.ld_temb = _ptr(temb), temb.stride(0)
    else:
        a.temb, a.ld_temb = None, 0
    if residual is not None:
        assert residual.is_contiguous() and tuple(residual.shape) == (B, Ho, Wo, Cout)
        a.residual, a.res_dtype = _ptr(residual), _dt(residual)
    else:
        a.residual, a.res_dtype = None, 0
    a.out, a.out_dtype = _ptr(out), _dt(out)
    a.B, a.H, a.W, a.Cin, a.Cout = B, H, W, Cin, Cout
    a.down2, a.up2 = int(down2), int(up2)
    a.out_channels = out.shape[-1] if out.dtype == torch.uint8 else 0
    fn = _lib.lib().sdb_conv3x3_f32 if x.dtype == torch.float32 else _lib.lib().sdb_conv3x3
    if sums and x.dtype == torch.bfloat16 and out.dtype == torch.bfloat16 and not down2:
        key = ("conv", H, W, Cin, Cout, up2)
        sm, arena = (None, None) if key in _sums_unsupported else _sums_new(x.device, B * Cout * 2, out.numel() * 2)
        if sm is not None:
            a.chan_sums = _ptr(sm)
            rc = fn(C.byref(a), _stream())
            if rc == 0:
                out._gn_sums = (arena.epoch, sm)
                sums_counters["tagged"] += 1
                return out
            if rc != -3:
                check(rc, "conv3x3")
            _sums_unsupported.add(key)
            a.chan_sums = None
    check(fn(C.byref(a), _stream()), "conv3x3")
    return out


def conv3x3_in(x_nchw: Tensor, w: Tensor, bias: Tensor, out_dtype=torch.bfloat16, repeat: int = 1) -> Tensor:
    """``repeat`` > 1: output batch = repeat * input batch, image b reads input b % B_in (virtual cat)."""
    _ag.refuse("conv3x3_in (gradient w.r.t. the latent)", x_nchw)
    if _ag.wants(w, bias):
        return _ag.conv3x3_in(x_nchw, w, bias, out_dtype, repeat)
    Bs, Cin, H, W = x_nchw.shape
    B = Bs * repeat
    Cout = w.shape[0]
    assert w.dtype == torch.float32 and w.is_contiguous() and bias.dtype == torch.float32
    out = torch.empty(B, H, W, Cout, device=x_nchw.device, dtype=out_dtype)
    check(_lib.lib().sdb_conv3x3_in(_ptr(x_nchw.contiguous()), _dt(x_nchw), _ptr(w), _ptr(bias), _ptr(out),
                                    _dt(out), B, Bs, H, W, Cin, Cout, _stream()), "conv3x3_in")
    return out


def cfg_combine(eps2: Tensor, guidance: float) -> Tensor:
    """first + g * (first - second) over the two batch halves (latent_diffusion.py:62-65)."""
    if _ag.wants(eps2):
        return _ag.cfg_combine(eps2, guidance)
    assert eps2.is_contiguous() and eps2.shape[0] % 2 == 0
    out = torch.empty((eps2.shape[0] // 2, *eps2.shape[1:]), device=eps2.device, dtype=eps2.dtype)
    check(_lib.lib().sdb_cfg_combine(_ptr(eps2), guidance, _ptr(out), out.numel(), _dt(eps2), _stream()),
          "cfg_combine")
    return out


def conv3x3_out(x_nhwc: Tensor, w: Tensor, bias: Tensor, out_dtype=torch.float32) -> Tensor:
    _ag.refuse("conv3x3_out", x_nhwc, w, bias)
    B, H, W, Cin = x_nhwc.shape
    Cout = w.shape[0]
    assert w.dtype == torch.float32 and w.is_contiguous() and x_nhwc.is_contiguous()
    out = torch.empty(B, Cout, H, W, device=x_nhwc.device, dtype=out_dtype)
    check(_lib.lib().sdb_conv3x3_out(_ptr(x_nhwc), _dt(x_nhwc), _ptr(w), _ptr(bias), _ptr(out), _dt(out), B, H,
                                     W, Cin, Cout, _stream()), "conv3x3_out")
    return out


def conv3x3_head(x_nhwc: Tensor, w16: Tensor, bias16: Tensor, cout: int, out_dtype=torch.float32) -> Tensor:
    """Output head on the tensor cores: 3x3 conv with Cout padded to 16 (``_pack.head_w16``), fp32 channels-last
    result, then the padding is dropped and the layout restored to NCHW [B, cout, H, W]."""
    if _ag.wants(x_nhwc, w16, bias16):
        return _ag.conv3x3_head(x_nhwc, w16, bias16, cout, out_dtype)
    B, H, W, _ = x_nhwc.shape
    y = conv3x3(x_nhwc, w16, bias16, out_dtype=torch.float32)
    out = torch.empty(B, cout, H, W, device=x_nhwc.device, dtype=out_dtype)
    check(_lib.lib().sdb_nhwc_slice_to_nchw(_ptr(y), y.shape[-1], _ptr(out), _dt(out), B, cout, H, W, _stream()),
          "nhwc_slice_to_nchw")
    return out


def conv3x3_head_uint8(x_nhwc: Tensor, w16: Tensor, bias16: Tensor, cout: int) -> Tensor:
    """The decoder's image head with the reference's ``detransform`` (utils/prepare_dataset.py:145-150) in the conv
    epilogue: uint8 NHWC [B, H, W, cout], PNG-ready -- no fp32 image is written."""
    _ag.refuse("conv3x3_head_uint8", x_nhwc, w16, bias16)
    B, H, W, _ = x_nhwc.shape
    out = torch.empty(B, H, W, cout, device=x_nhwc.device, dtype=torch.uint8)
    return conv3x3(x_nhwc, w16, bias16, out=out)


def conv1x1_nchw_small(x: Tensor, w: Tensor, bias: Tensor) -> Tensor:
    _ag.refuse("conv1x1_nchw_small", x, w, bias)
    B, Cin, H, W = x.shape
    Cout = w.shape[0]
    out = torch.empty(B, Cout, H, W, device=x.device, dtype=x.dtype)
    check(_lib.lib().sdb_conv1x1_nchw_small(_ptr(x.contiguous()), _ptr(w), _ptr(bias), _ptr(out), _dt(x), B, H, W,
                                            Cin, Cout, _stream()), "conv1x1_nchw_small")
    return out


def upsample2x(x: Tensor) -> Tensor:
    if _ag.wants(x):
        return _ag.upsample2x(x)
    B, H, W, Cc = x.shape
    assert x.is_contiguous()
    out = torch.empty(B, 2 * H, 2 * W, Cc, device=x.device, dtype=x.dtype)
    check(_lib.lib().sdb_upsample2x(_ptr(x), _ptr(out), _dt(x), B, H, W, Cc, _stream()), "upsample2x")
    return out


def attention(q: Tensor, k: Tensor, v: Tensor, heads: int, d: int, scale: float,
              out: Optional[Tensor] = None) -> Tensor:
    """q [B, n, >=heads*d] (last-dim stride 1, row-strided views allowed), k/v [B, m, ...]."""
    if _ag.wants(q, k, v):
        _no_f32_grad(q)
        return _ag.attention(q, k, v, heads, d, scale)
    B, n = q.shape[0], q.shape[1]
    m = k.shape[1]
    for t in (q, k, v):
        assert t.dtype == q.dtype and t.stride(2) == 1 and t.stride(0) == t.shape[1] * t.stride(1)
    assert q.dtype in (torch.bfloat16, torch.float32)
    if out is None:
        out = torch.empty(B, n, heads * d, device=q.device, dtype=q.dtype)
    if d > ATTENTION_MAX_FUSED_D:
        return _attention_wide(q, k, v, heads, d, scale, out)
    check(_lib.lib().sdb_attention(_ptr(q), q.stride(1), _ptr(k), k.stride(1), _ptr(v), v.stride(1), _ptr(out),
                                   out.stride(1), B, heads, n, m, d, scale, _dt(q), _stream()), "attention")
    return out


ATTENTION_MAX_FUSED_D = 128     # head dims the fused (flash-style) kernels hold in registers / shared memory


def _attention_wide(q: Tensor, k: Tensor, v: Tensor, heads: int, d: int, scale: float, out: Tensor) -> Tensor:
    """Head dims beyond the fused kernels (the VAE encoder bottleneck has ONE head of d = mid_ch, e.g. 512 for the
    AutoencoderKL shape, models/autoencoder.py:181-183): per (batch, head) S = Q K^T on the GEMM kernel (fp32 out),
    ``sdb_softmax_rows``, O = P V on the GEMM kernel.  `sim` is materialised -- the sequence there is (H/8)^2 tokens.
    Keys are zero-padded to a multiple of 16 (the GEMM's N / K granularity); the softmax zeroes the padding."""
    assert d % 8 == 0, f"attention: head dim {d} > {ATTENTION_MAX_FUSED_D} must be a multiple of 8"
    B, n, m = q.shape[0], q.shape[1], k.shape[1]
    m16 = (m + 15) // 16 * 16
    L = _lib.lib()
    kp = torch.zeros(m16, d, device=q.device, dtype=q.dtype)
    vt = torch.zeros(d, m16, device=q.device, dtype=q.dtype)
    s = torch.empty(n, m16, device=q.device, dtype=torch.float32)
    p = torch.empty(n, m16, device=q.device, dtype=q.dtype)
    for b in range(B):
        for h in range(heads):
            cols = slice(h * d, (h + 1) * d)
            kp[:m].copy_(k[b, :, cols])
            vt[:, :m].copy_(v[b, :, cols].t())
            gemm(q[b, :, cols], kp, out=s)
            check(L.sdb_softmax_rows(_ptr(s), m16, _ptr(p), m16, _dt(p), n, m16, m, scale, _stream()), "softmax_rows")
            gemm(p, vt, out=out[b, :, cols])
    return out


def attention_p32(q: Tensor, k: Tensor, v: Tensor, heads: int, d: int, scale: float,
                  out: Optional[Tensor] = None) -> Tensor:
    """tcgen05 attention over head-padded operands: q [B, n, >=heads*32], k/v [B, m, >=heads*32] (row-strided
    views allowed); head h at columns [32h, 32h+32), q/k zero beyond d, v[..., 32h+d] == 1."""
    if _ag.wants(q, k, v):
        return _ag.attention_p32(q, k, v, heads, d, scale)
    B, n = q.shape[0], q.shape[1]
    m = k.shape[1]
    for t in (q, k, v):
        assert t.dtype == torch.bfloat16 and t.stride(2) == 1 and t.stride(0) == t.shape[1] * t.stride(1)
    if out is None:
        out = torch.empty(B, n, heads * d, device=q.device, dtype=q.dtype)
    check(_lib.lib().sdb_attention_p32(_ptr(q), q.stride(1), _ptr(k), k.stride(1), _ptr(v), v.stride(1), _ptr(out),
                                       out.stride(1), B, heads, n, m, d, scale, _stream()), "attention_p32")
    return out


def attention_p32_packed(qkv: Tensor, heads: int, d: int, scale: float) -> Tensor:
    """Self-attention over the packed, head-padded projection output qkv [B, n, 3*heads*32] (Q | K | V column blocks).
    Same kernels as ``attention_p32``; under autograd the gradient comes back as one packed tensor."""
    if _ag.wants(qkv):
        return _ag.attention_p32_packed(qkv, heads, d, scale)
    kw = heads * 32
    return attention_p32(qkv[:, :, :kw], qkv[:, :, kw:2 * kw], qkv[:, :, 2 * kw:], heads, d, scale)


def attention_p32_lse(q: Tensor, k: Tensor, v: Tensor, heads: int, d: int, scale: float):
    """``attention_p32`` that also returns the fp32 row log-sum-exp [B*heads*n] (log2 domain) the backward needs, or
    None when the kernel that ran does not keep it (short contexts)."""
    B, n = q.shape[0], q.shape[1]
    m = k.shape[1]
    for t in (q, k, v):
        assert t.dtype == torch.bfloat16 and t.stride(2) == 1 and t.stride(0) == t.shape[1] * t.stride(1)
    out = torch.empty(B, n, heads * d, device=q.device, dtype=q.dtype)
    lse = torch.empty(B * heads * n, device=q.device, dtype=torch.float32)
    wrote = C.c_int(0)
    check(_lib.lib().sdb_attention_p32_lse(_ptr(q), q.stride(1), _ptr(k), k.stride(1), _ptr(v), v.stride(1), _ptr(out),
                                           out.stride(1), B, heads, n, m, d, scale, _ptr(lse), C.byref(wrote),
                                           _stream()), "attention_p32_lse")
    return out, (lse if wrote.value else None)


class _DropoutState:
    """Per-device RNG state of the training-mode dropout kernels: one int64 in device memory (advanced once per train
    step, also inside captured graphs) and a host-side counter that numbers the dropout calls of one forward."""

    def __init__(self, device):
        self.state = torch.zeros(1, device=device, dtype=torch.int64)
        self.site = 0


_dropout_states = {}


def dropout_state(device) -> _DropoutState:
    device = torch.device(device)
    if device.type == "cuda" and device.index is None:       # "cuda" and "cuda:0" must name the same state
        device = torch.device("cuda", torch.cuda.current_device())
    key = str(device)
    if key not in _dropout_states:
        _dropout_states[key] = _DropoutState(device)
    return _dropout_states[key]


def dropout_seed(seed: int, device) -> None:
    st = dropout_state(device)
    st.state.fill_(int(seed))
    st.site = 0


def dropout_begin_step(device) -> None:
    """Call at the start of every forward: dropout layers are numbered in call order from 0."""
    dropout_state(device).site = 0


def dropout_advance(device) -> None:
    """Next step draws new masks (a device-side increment: legal inside CUDA-graph capture)."""
    dropout_state(device).state.add_(1)


def dropout_raw(x: Tensor, p: float, site: int) -> Tensor:
    """The kernel itself: mask keyed by (device state, site); applying it to a gradient is the backward."""
    x = x.contiguous()
    out = torch.empty_like(x)
    st = dropout_state(x.device)
    check(_lib.lib().sdb_dropout(_ptr(x), _ptr(out), x.numel(), float(p), _ptr(st.state), int(site), _dt(x), _stream()),
          "dropout")
    return out


def dropout(x: Tensor, p: float, training: bool = True) -> Tensor:
    """nn.Dropout(p) (resnet2d.py:190, transformer.py:58,147): identity unless ``training`` and p > 0."""
    if not training or not p:
        return x
    st = dropout_state(x.device)
    site = st.site
    st.site += 1
    if _ag.wants(x):
        return _ag.dropout(x, p, site)
    return dropout_raw(x, p, site)


def silu(x: Tensor, out_dtype=None) -> Tensor:
    if _ag.wants(x):
        return _ag.silu(x, out_dtype)
    out = torch.empty(x.shape, device=x.device, dtype=out_dtype or x.dtype)
    check(_lib.lib().sdb_silu(_ptr(x.contiguous()), _ptr(out), x.numel(), _dt(x), _dt(out), _stream()), "silu")
    return out


def cast(x: Tensor, dtype) -> Tensor:
    if _ag.wants(x):
        return _ag.cast(x, dtype)
    out = torch.empty(x.shape, device=x.device, dtype=dtype)
    check(_lib.lib().sdb_cast(_ptr(x.contiguous()), _dt(x), _ptr(out), _dt(out), x.numel(), _stream()), "cast")
    return out


def to_uint8_image(x_nchw: Tensor) -> Tensor:
    _ag.refuse("to_uint8_image", x_nchw)
    B, Cc, H, W = x_nchw.shape
    out = torch.empty(B, H, W, Cc, device=x_nchw.device, dtype=torch.uint8)
    check(_lib.lib().sdb_to_uint8_image(_ptr(x_nchw.contiguous()), _dt(x_nchw), _ptr(out), B, Cc, H, W, _stream()),
          "to_uint8_image")
    return out


def debug_timeout_flag() -> int:
    return _lib.lib().sdb_debug_timeout_flag()


def launch_count() -> int:
    return _lib.lib().sdb_launch_count()


def mse_loss(pred: Tensor, target: Tensor) -> Tensor:
    """Mean squared error in fp32 (train_unet.py:517); the gradient is produced by the same kernel."""
    return _ag.mse_loss(pred, target)


def sumsq(x: Tensor, out: Optional[Tensor] = None) -> Tensor:
    """Adds sum(x^2) of a flat fp32 buffer into ``out`` (zeroed if not given): the grad-norm of the clip."""
    if out is None:
        out = torch.zeros(1, device=x.device, dtype=torch.float32)
    check(_lib.lib().sdb_sumsq(_ptr(x), x.numel(), _ptr(out), _stream()), "sumsq")
    return out


def wgrad_linear(dy: Tensor, x: Tensor, out: Optional[Tensor] = None) -> Optional[Tensor]:
    """dW [N, K] (fp32) = dy[M, N]^T @ x[M, K] on the tcgen05 weight-gradient kernel (row-strided bf16 views allowed;
    accumulates into ``out`` when given).  Returns None when the shape is outside the kernel (caller falls back to
    the transpose + GEMM route)."""
    M, N = dy.shape
    K = x.shape[1]
    assert x.shape[0] == M and dy.dtype == torch.bfloat16 and x.dtype == torch.bfloat16
    assert dy.stride(1) == 1 and x.stride(1) == 1
    if (K % 160 and K % 64) or dy.stride(0) % 8 or x.stride(0) % 8 or dy.data_ptr() % 16 or x.data_ptr() % 16:
        return None
    if out is None:
        out = torch.zeros(N, K, device=dy.device, dtype=torch.float32)
    assert out.dtype == torch.float32 and out.stride(1) == 1 and tuple(out.shape) == (N, K)
    check(_lib.lib().sdb_wgrad_linear(_ptr(dy), dy.stride(0), _ptr(x), x.stride(0), _ptr(out), out.stride(0), M, N, K,
                                      _stream()), "wgrad_linear")
    return out


def wgrad_conv3x3(dy: Tensor, x: Tensor) -> Optional[Tensor]:
    """dW [Cout, 3, 3, Cin] (fp32) of a 3x3 / pad 1 convolution from NHWC bf16 ``dy`` [B,H,W,Cout] and ``x``
    [B,H,W,Cin]; None when the geometry is outside the kernel."""
    B, H, W, Cin = x.shape
    Cout = dy.shape[-1]
    assert tuple(dy.shape[:3]) == (B, H, W) and dy.is_contiguous() and x.is_contiguous()
    assert dy.dtype == torch.bfloat16 and x.dtype == torch.bfloat16
    th = min(H, 64 // W) if W <= 64 and 64 % W == 0 else 0
    if (Cin % 160 and Cin % 64) or Cout % 8 or not th or H % th or 64 % (W * th):
        return None
    out = torch.zeros(Cout, 3, 3, Cin, device=x.device, dtype=torch.float32)
    check(_lib.lib().sdb_wgrad_conv3x3(_ptr(dy), _ptr(x), _ptr(out), B, H, W, Cin, Cout, _stream()), "wgrad_conv3x3")
    return out
