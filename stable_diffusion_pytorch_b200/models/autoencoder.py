"""AutoEncoderKL with the reference's config fields, state_dict keys and ``encode`` / ``decode``
signatures (stable_diffusion/models/autoencoder.py:22-246).

decode (the hot-path row V1): post_quant_conv (1x1, thin NCHW kernel) -> conv_in (CUDA-core, NCHW ->
NHWC) -> bottleneck [ResBlock(g=2), bare 1-head self-attention, ResBlock] -> up blocks (tcgen05
implicit-GEMM convs, nearest-x2 upsampling) -> GroupNorm+SiLU -> head conv (writes NCHW).
The returned tensor also answers ``.sample`` so the diffusers-style call in the reference trainer
(``autoencoder.decode(x).sample``, train_unet.py:562-563) works unchanged.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import torch
from torch import nn

from .. import _pack, ops
from .._lib import ACT_SILU
from ..modules.distributions import GaussianDistribution
from .utils import (build_bottleneck, build_conv_in, build_final_output, build_input_blocks,
                    build_output_blocks)

Tensor = torch.Tensor


@dataclass
class AutoencoderConfig:
    in_channels: int = 3
    latent_channels: int = 4
    out_channels: Optional[int] = 3
    autoencoder_channels_list: List[int] = field(default_factory=lambda: [64, 128])
    autoencoder_num_res_blocks: int = 2
    groups: int = 32
    kl_weight: float = 1.0


class DecodedImage(torch.Tensor):
    """A plain tensor that also exposes ``.sample`` (diffusers' DecoderOutput duck type)."""

    @property
    def sample(self):
        return self.as_subclass(torch.Tensor)


@dataclass
class AutoEncoderKLOutput:
    latent_dist: "GaussianDistribution"


class _ConvNet(nn.Module):
    """Shared machinery of Encoder / Decoder: thin conv_in, channels-last body, thin head."""

    def __init__(self, groups: int):
        super().__init__()
        self.groups = groups
        self._cache = _pack.PackCache()

    def _thin(self):
        c = self._cache
        gn, head = self.out[0], self.out[2]
        return {
            "wci": c.get("wci", [self.conv_in.weight], lambda: _pack.f32(self.conv_in.weight)),
            "bci": c.get("bci", [self.conv_in.bias], lambda: _pack.f32(self.conv_in.bias)),
            "go": c.get("go", [gn.weight], lambda: _pack.f32(gn.weight)),
            "bo": c.get("bo", [gn.bias], lambda: _pack.f32(gn.bias)),
            "wh": c.get("wh", [head.weight, head.bias], lambda: _pack.head_w16(head.weight, head.bias)),
        }

    def _head(self, h: Tensor, p, out_dtype, uint8: bool = False) -> Tensor:
        h = ops.group_norm(h, p["go"], p["bo"], self.groups, ACT_SILU)
        w16, b16 = p["wh"]
        if uint8 and h.dtype == torch.bfloat16:
            # detransform (utils/prepare_dataset.py:145-150) in the head conv's epilogue: uint8 NHWC straight out
            return ops.conv3x3_head_uint8(h, w16, b16, self.out[2].weight.shape[0])
        y = ops.conv3x3_head(h, w16, b16, self.out[2].weight.shape[0], out_dtype)
        return ops.to_uint8_image(y) if uint8 else y


class Encoder(_ConvNet):
    def __init__(self, in_channels: int, out_channels: int, channels_list: List[int], num_res_blocks: int,
                 groups: int = 4):
        super().__init__(groups)
        self.conv_in = build_conv_in(in_channels=in_channels, out_channels=channels_list[0])
        self.down, _, mid_ch, _, _ = build_input_blocks(
            in_channels=channels_list[0], num_res_blocks=num_res_blocks, levels=len(channels_list),
            channels_list=channels_list, groups=groups)
        self.bottleneck = build_bottleneck(mid_ch, d_head=mid_ch, use_attn_only=True, groups=groups)
        self.out = build_final_output(out_ch=mid_ch, out_channels=2 * out_channels, groups=groups)

    def forward(self, x: Tensor) -> Tensor:
        _pack.require_cuda(x, "Encoder")
        p = self._thin()
        with ops.gn_sums_scope(x.device):      # GroupNorm statistics come out of the producing conv epilogues
            h = ops.conv3x3_in(x, p["wci"], p["bci"], _pack.compute_dtype())
            for module in self.down:
                h = module.forward_nhwc(h)
            h = self.bottleneck.forward_nhwc(h)
            return self._head(h, p, x.dtype)


class Decoder(_ConvNet):
    def __init__(self, in_channels: int, out_channels: int, channels_list: List[int], num_res_blocks: int,
                 groups: int = 4):
        super().__init__(groups)
        self.conv_in = build_conv_in(in_channels=in_channels, out_channels=channels_list[0])
        self.bottleneck = build_bottleneck(in_ch=channels_list[0], d_head=channels_list[0], use_attn_only=True,
                                           groups=groups)
        self.up, mid_ch = build_output_blocks(num_res_blocks=num_res_blocks, in_ch=channels_list[0],
                                              levels=len(channels_list), channels_list=channels_list, groups=groups)
        self.out = build_final_output(out_ch=mid_ch, out_channels=out_channels, groups=groups)

    def forward(self, x: Tensor, uint8: bool = False) -> Tensor:
        """``uint8`` (extension): return the PNG-ready [B, H, W, C] uint8 image instead of the float NCHW one."""
        _pack.require_cuda(x, "Decoder")
        p = self._thin()
        with ops.gn_sums_scope(x.device):      # GroupNorm statistics come out of the producing conv epilogues
            h = ops.conv3x3_in(x, p["wci"], p["bci"], _pack.compute_dtype())
            h = self.bottleneck.forward_nhwc(h)
            for module in self.up:
                h = module.forward_nhwc(h)
            return self._head(h, p, x.dtype, uint8)


class AutoEncoderKL(nn.Module):
    def __init__(self, cfg):
        assert cfg.out_channels is None or cfg.out_channels == cfg.in_channels, (
            f"input channels({cfg.in_channels}) of image should be equal to output channels({cfg.out_channels})")
        super().__init__()
        self.latent_channels = latent_channels = cfg.latent_channels
        self.encoder = self.build_encoder(cfg)
        self.decoder = self.build_decoder(cfg)
        self.quant_conv = nn.Conv2d(2 * latent_channels, 2 * latent_channels, 1)
        self.post_quant_conv = nn.Conv2d(latent_channels, latent_channels, 1)
        self._cache = _pack.PackCache()

    @classmethod
    def build_encoder(cls, cfg):
        return Encoder(in_channels=cfg.in_channels, out_channels=cfg.latent_channels,
                       channels_list=cfg.autoencoder_channels_list, num_res_blocks=cfg.autoencoder_num_res_blocks,
                       groups=cfg.groups)

    @staticmethod
    def build_decoder(cfg):
        return Decoder(in_channels=cfg.latent_channels, out_channels=cfg.out_channels or cfg.in_channels,
                       channels_list=cfg.autoencoder_channels_list, num_res_blocks=cfg.autoencoder_num_res_blocks,
                       groups=cfg.groups)

    def _thin_conv(self, name: str, conv: nn.Conv2d):
        w = self._cache.get(name + "w", [conv.weight],
                            lambda: _pack.f32(conv.weight.reshape(conv.weight.shape[0], conv.weight.shape[1])))
        b = self._cache.get(name + "b", [conv.bias], lambda: _pack.f32(conv.bias))
        return w, b

    def encode(self, img: Tensor) -> AutoEncoderKLOutput:
        z = self.encoder(img)
        moments = ops.conv1x1_nchw_small(z, *self._thin_conv("q", self.quant_conv))
        return AutoEncoderKLOutput(GaussianDistribution(moments))

    def decode(self, latent: Tensor) -> Tensor:
        assert latent.shape[1] == self.latent_channels, (
            f"Expected latent representation to have {self.latent_channels} channels, got {latent.shape[1]}")
        _pack.require_cuda(latent, "AutoEncoderKL.decode")
        z = ops.conv1x1_nchw_small(latent, *self._thin_conv("pq", self.post_quant_conv))
        return self.decoder(z).as_subclass(DecodedImage)

    @torch.no_grad()
    def decode_uint8(self, latent: Tensor, chunk: int = 16) -> Tensor:
        """``decode`` followed by the reference's ``detransform`` (utils/prepare_dataset.py:145-150: (x + 1) / 2,
        clip, * 255, uint8) inside the head conv's epilogue: [B, H, W, C] uint8, PNG-ready, a quarter of the fp32 image bytes
        -- what the multi-GPU image gather moves; no float image is materialised.  Decodes ``chunk`` latents at a time (2 GB of activations each at 512 x 512)."""
        assert latent.shape[1] == self.latent_channels, (
            f"Expected latent representation to have {self.latent_channels} channels, got {latent.shape[1]}")
        _pack.require_cuda(latent, "AutoEncoderKL.decode_uint8")
        pq = self._thin_conv("pq", self.post_quant_conv)
        outs = [self.decoder(ops.conv1x1_nchw_small(z, *pq), uint8=True) for z in latent.split(chunk)]
        return outs[0] if len(outs) == 1 else torch.cat(outs)
