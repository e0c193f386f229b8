"""ResBlock / UpSample / DownSample with the reference's constructor signatures and state_dict
keys (stable_diffusion/modules/resnet2d.py:9-238), executed by libsdb200 kernels.

Public ``forward`` takes and returns NCHW tensors like the reference; inside the UNet / decoder
the channels-last ``forward_nhwc`` entry points are chained directly so no layout change happens
between blocks.  ``nn.GroupNorm`` / ``nn.Conv2d`` / ``nn.Linear`` children are parameter holders
only (identical names, shapes and default init); their torch ``forward`` is never called.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .. import _pack, ops
from .._lib import ACT_SILU

Tensor = torch.Tensor


class _NhwcModule(nn.Module):
    """NCHW facade over a channels-last implementation."""

    @property
    def act_dtype(self):  # activations between kernels: bf16, or fp32 in the parity mode
        return _pack.compute_dtype()

    def __init__(self):
        super().__init__()
        self._cache = _pack.PackCache()

    def _to_nhwc(self, x: Tensor) -> Tensor:
        _pack.require_cuda(x, type(self).__name__)
        return ops.nchw_to_nhwc(x, self.act_dtype)

    @staticmethod
    def _to_nchw(x: Tensor, dtype) -> Tensor:
        return ops.nhwc_to_nchw(x, dtype)


class UpSample(_NhwcModule):
    """nearest x2 then 3x3 conv (resnet2d.py:46-63)."""

    def __init__(self, in_channels: int, out_channels: Optional[int] = None, scale_factor: float = 2,
                 padding: int = 1):
        super().__init__()
        assert scale_factor == 2 and padding == 1, "only the reference defaults are built"
        self.in_channels = in_channels
        self.out_channels = out_channels or in_channels
        self.scale_factor = scale_factor
        self.conv = nn.Conv2d(self.in_channels, self.out_channels, kernel_size=3, padding=padding)

    def forward_nhwc(self, x: Tensor) -> Tensor:
        assert x.shape[-1] == self.in_channels, (
            f"input channel does not match: x.shape[1]({x.shape[-1]}) != self.in_channels({self.in_channels}))")
        b = self._cache.get("b", [self.conv.bias], lambda: _pack.f32(self.conv.bias))
        if x.dtype == torch.bfloat16 and not _pack.in_graph_mode() and not (torch.is_grad_enabled() and x.requires_grad):
            # inference: the upsampling is folded into the conv as four 2x2-tap phase convolutions (4/9 of the
            # flops, no 4x tensor written and re-read)
            wp = self._cache.get("wup", [self.conv.weight], lambda: _pack.conv3x3_up_w(self.conv.weight))
            return ops.conv3x3(x, wp, b, up2=True, sums=True)
        w = self._cache.get("w", [self.conv.weight], lambda: _pack.conv3x3_w(self.conv.weight))
        return ops.conv3x3(ops.upsample2x(x), w, b)

    def forward(self, x: Tensor) -> Tensor:
        assert x.shape[1] == self.in_channels, (
            f"input channel does not match: x.shape[1]({x.shape[1]}) != self.in_channels({self.in_channels}))")
        return self._to_nchw(self.forward_nhwc(self._to_nhwc(x)), x.dtype)


class DownSample(_NhwcModule):
    """3x3 conv then nearest x0.5 (resnet2d.py:108-124) == keep the even rows/cols of the conv."""

    def __init__(self, in_channels: int, out_channels: Optional[int] = None, scale_factor: float = 1 / 2,
                 padding: int = 1):
        super().__init__()
        assert scale_factor == 1 / 2 and padding == 1, "only the reference defaults are built"
        self.in_channels = in_channels
        self.scale_factor = scale_factor
        self.out_channels = out_channels or in_channels
        self.conv = nn.Conv2d(self.in_channels, self.out_channels, kernel_size=3, padding=padding)

    def forward_nhwc(self, x: Tensor) -> Tensor:
        assert x.shape[-1] == self.in_channels, (
            f"input channel does not match: x.shape[1]({x.shape[-1]}) != self.in_channels({self.in_channels}))")
        if x.shape[1] % 2 or x.shape[2] % 2:
            raise ValueError("DownSample: H and W must be even (conv + nearest x0.5 == stride-2 sampling)")
        w = self._cache.get("w", [self.conv.weight], lambda: _pack.conv3x3_w(self.conv.weight))
        b = self._cache.get("b", [self.conv.bias], lambda: _pack.f32(self.conv.bias))
        return ops.conv3x3(x, w, b, down2=True)

    def forward(self, x: Tensor) -> Tensor:
        assert x.shape[1] == self.in_channels, (
            f"input channel does not match: x.shape[1]({x.shape[1]}) != self.in_channels({self.in_channels}))")
        return self._to_nchw(self.forward_nhwc(self._to_nhwc(x)), x.dtype)


class ResBlock(_NhwcModule):
    """GN-SiLU-conv3x3 (+time emb) -> GN-SiLU-Dropout-conv3x3(zero-init) + skip (resnet2d.py:127-238).

    Kernel plan: GN statistics + normalise/SiLU (virtual concat of ``x`` and the UNet skip tensor),
    tcgen05 implicit-GEMM conv with bias + time-embedding epilogue, GN again, tcgen05 conv with
    bias + residual epilogue; the 1x1 skip projection is a tcgen05 GEMM over the two concat halves.
    """

    def __init__(self, in_channels: int, out_channels: Optional[int] = None, time_emb_dim: Optional[int] = None,
                 dropout: Optional[float] = 0, padding: Optional[int] = 1, groups: int = 2) -> None:
        super().__init__()
        assert in_channels % groups == 0, (
            f"in_channels({in_channels}) must be divisible by num_groups({groups})")
        assert padding == 1
        self.in_channels = in_channels
        self.out_channels = out_channels or in_channels
        self.time_emb_dim = time_emb_dim
        self.dropout = dropout
        self.groups = groups
        self.in_layers = nn.Sequential(
            nn.GroupNorm(num_groups=groups, num_channels=in_channels), nn.SiLU(),
            nn.Conv2d(self.in_channels, self.out_channels, kernel_size=3, padding=padding))
        if self.time_emb_dim:
            self.time_embedding = nn.Sequential(nn.SiLU(), nn.Linear(self.time_emb_dim, self.out_channels))
        else:
            self.time_embedding = nn.Identity()
        self.out_layers = nn.Sequential(
            nn.GroupNorm(num_groups=groups, num_channels=self.out_channels), nn.SiLU(), nn.Dropout(self.dropout),
            _pack.zero_module(nn.Conv2d(self.out_channels, self.out_channels, kernel_size=3, padding=padding)))
        if self.in_channels != self.out_channels:
            self.skip_connection = nn.Conv2d(self.in_channels, self.out_channels, kernel_size=1)
        else:
            self.skip_connection = nn.Identity()

    # -- packed operands ---------------------------------------------------------------------
    def _p(self):
        c = self._cache
        gn1, conv1 = self.in_layers[0], self.in_layers[2]
        gn2, conv2 = self.out_layers[0], self.out_layers[3]
        d = {
            "g1": c.get("g1", [gn1.weight], lambda: _pack.f32(gn1.weight)),
            "b1": c.get("b1", [gn1.bias], lambda: _pack.f32(gn1.bias)),
            "w1": c.get("w1", [conv1.weight], lambda: _pack.conv3x3_w(conv1.weight)),
            "cb1": c.get("cb1", [conv1.bias], lambda: _pack.f32(conv1.bias)),
            "g2": c.get("g2", [gn2.weight], lambda: _pack.f32(gn2.weight)),
            "b2": c.get("b2", [gn2.bias], lambda: _pack.f32(gn2.bias)),
            "w2": c.get("w2", [conv2.weight], lambda: _pack.conv3x3_w(conv2.weight)),
            "cb2": c.get("cb2", [conv2.bias], lambda: _pack.f32(conv2.bias)),
        }
        if isinstance(self.skip_connection, nn.Conv2d):
            sk = self.skip_connection
            d["ws"] = c.get("ws", [sk.weight], lambda: _pack.conv1x1_w(sk.weight))
            d["bs"] = c.get("bs", [sk.bias], lambda: _pack.f32(sk.bias))
        return d

    def time_proj_params(self):
        """(weight [Cout, T] fp32, bias [Cout] fp32) of ``time_embedding.1`` for the UNet's batched
        projection, or None."""
        if not self.time_emb_dim:
            return None
        lin = self.time_embedding[1]
        return lin.weight, lin.bias

    def forward_nhwc(self, x: Tensor, temb_proj: Optional[Tensor] = None, x2: Optional[Tensor] = None) -> Tensor:
        """x (+ x2 = UNet skip, virtual concat) NHWC; temb_proj = Linear(SiLU(time_emb)) fp32 [B, Cout]."""
        c_in = x.shape[-1] + (x2.shape[-1] if x2 is not None else 0)
        assert c_in == self.in_channels, f"input channels {c_in} != in_channels {self.in_channels}"
        p = self._p()
        B, H, W, _ = x.shape
        h = ops.group_norm(x, p["g1"], p["b1"], self.groups, ACT_SILU, x2=x2)
        h = ops.conv3x3(h, p["w1"], p["cb1"], temb=temb_proj, sums=True)     # GN2's statistics from this epilogue
        h = ops.group_norm(h, p["g2"], p["b2"], self.groups, ACT_SILU)
        h = ops.dropout(h, self.dropout or 0.0, self.training)             # out_layers[2] (resnet2d.py:190)
        if "ws" in p:
            c1 = x.shape[-1]
            skip = ops.gemm(x.view(B * H * W, c1), p["ws"], p["bs"],
                            a2=None if x2 is None else x2.view(B * H * W, x2.shape[-1]))
            skip = skip.view(B, H, W, self.out_channels)
        else:
            skip = x if x2 is None else torch.cat([x, x2], dim=-1)
        return ops.conv3x3(h, p["w2"], p["cb2"], residual=skip, sums=True)  # the next block's GroupNorm reads it

    def project_time(self, time_emb: Tensor) -> Tensor:
        """Linear(SiLU(time_emb)) (resnet2d.py:179-183) for a stand-alone call."""
        lin = self.time_embedding[1]
        w = self._cache.get("wt", [lin.weight], lambda: _pack.linear_w(lin.weight))
        b = self._cache.get("bt", [lin.bias], lambda: _pack.f32(lin.bias))
        return ops.gemm(ops.silu(time_emb, _pack.compute_dtype()), w, b, out_dtype=torch.float32)

    def forward(self, x: Tensor, time_emb: Optional[Tensor] = None) -> Tensor:
        tp = None
        if time_emb is not None:
            assert x.shape[0] == time_emb.shape[0], (
                f"batch size does not match: x.shape[0]({x.shape[0]}) != time_emb.shape[0]({time_emb.shape[0]})")
            assert time_emb.shape[1] == self.time_emb_dim, (
                f"time_emb_dim does not match: time_emb.shape[1]({time_emb.shape[1]}) != self.time_emb_dim({self.time_emb_dim})")
            tp = self.project_time(time_emb)
        return self._to_nchw(self.forward_nhwc(self._to_nhwc(x), tp), x.dtype)
