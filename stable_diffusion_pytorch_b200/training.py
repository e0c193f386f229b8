"""The UNet train step (row T1; train_unet.py:489-529 ``__one_step`` + :359-375 backward / clip / AdamW).

What the reference does with accelerate + DDP + torch.optim.AdamW, this module does with
  * the forward of ``LatentDiffusion.pred_noise`` (CFG batch doubling included, train_unet.py:524-526) running on the
    library's kernels through ``autograd_ops`` (weights re-packed differentiably each step, ``_pack.graph_mode``),
  * one flat fp32 parameter buffer, one flat fp32 gradient buffer (every ``p.grad`` is a view into it), so that
      - the data-parallel gradient exchange is ONE all-reduce of the flat buffer (NCCL over NVLink on the GPU box,
        gloo in the CPU tests) instead of DDP's bucket bookkeeping,
      - ``clip_grad_norm_`` is one sum-of-squares kernel whose result stays on the device,
      - AdamW (torch.optim.AdamW semantics: decoupled weight decay, bias correction) is one kernel over the whole
        model that also applies the 1/world mean and the clip factor -- no host synchronisation in the step.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterable, Optional

import torch
import torch.distributed as dist
from torch import nn

from . import _lib, _pack, ops
from ._lib import check

Tensor = torch.Tensor


@dataclass
class OptimConfig:
    """Field names and defaults of the reference's OptimConfig (trainer_args.py:69-90); torch.optim.AdamW defaults
    for the rest (train_unet.py:231-236 passes only lr and weight_decay)."""
    learning_rate: float = 4e-5
    adam_weight_decay: float = 0.1
    max_grad_norm: float = 0.1
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_eps: float = 1e-8


class FlatAdamW:
    """Flat-bucket AdamW over ``params`` (all fp32, on one CUDA device)."""

    def __init__(self, params: Iterable[nn.Parameter], cfg: Optional[OptimConfig] = None):
        self.cfg = cfg or OptimConfig()
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("FlatAdamW: no trainable parameters")
        dev = self.params[0].device
        for p in self.params:
            if p.dtype != torch.float32 or p.device != dev:
                raise TypeError("FlatAdamW: master parameters must be fp32 on one device")
        _pack.require_cuda(self.params[0], "FlatAdamW")
        n = sum(p.numel() for p in self.params)
        self.numel = n
        self.flat = torch.empty(n, device=dev, dtype=torch.float32)
        self.grad = torch.zeros(n, device=dev, dtype=torch.float32)
        self.m = torch.zeros(n, device=dev, dtype=torch.float32)
        self.v = torch.zeros(n, device=dev, dtype=torch.float32)
        self.gnorm_sq = torch.zeros(1, device=dev, dtype=torch.float32)
        self.step_count = 0
        o = 0
        with torch.no_grad():
            for p in self.params:
                k = p.numel()
                self.flat[o:o + k].copy_(p.reshape(-1))
                p.data = self.flat[o:o + k].view(p.shape)
                p.grad = self.grad[o:o + k].view(p.shape)
                o += k

    def zero_grad(self) -> None:
        self.grad.zero_()
        for p, g in zip(self.params, self._grad_views()):   # a backward may have re-pointed .grad; restore the views
            if p.grad is None or p.grad.data_ptr() != g.data_ptr():
                p.grad = g

    def _grad_views(self):
        o = 0
        for p in self.params:
            k = p.numel()
            yield self.grad[o:o + k].view(p.shape)
            o += k

    def _collect_strays(self) -> None:
        """autograd accumulates in place into the flat views; a parameter whose .grad was replaced (not expected)
        is copied back so the flat buffer is always the truth."""
        for p, g in zip(self.params, self._grad_views()):
            if p.grad is not None and p.grad.data_ptr() != g.data_ptr():
                g.copy_(p.grad)
                p.grad = g

    def all_reduce(self, group=None) -> int:
        """Sum the flat gradient over the data-parallel group; returns the world size (the mean is applied inside
        the AdamW kernel)."""
        self._collect_strays()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.grad, op=dist.ReduceOp.SUM, group=group)
            return dist.get_world_size(group)
        return 1

    def grad_norm(self, world: int = 1) -> Tensor:
        """Global L2 norm of the (mean) gradient, on the device."""
        self.gnorm_sq.zero_()
        ops.sumsq(self.grad, self.gnorm_sq)
        return self.gnorm_sq.sqrt() / world

    def step(self, world: int = 1, lr: Optional[float] = None) -> None:
        """clip_grad_norm_(max_grad_norm) + AdamW on the whole model, one kernel after the norm kernel."""
        c = self.cfg
        self.step_count += 1
        self.gnorm_sq.zero_()
        ops.sumsq(self.grad, self.gnorm_sq)
        check(_lib.lib().sdb_adamw(ops._ptr(self.flat), ops._ptr(self.grad), ops._ptr(self.m), ops._ptr(self.v), self.numel,
                                   c.learning_rate if lr is None else lr, c.adam_beta1, c.adam_beta2, c.adam_eps,
                                   c.adam_weight_decay, self.step_count, 1.0 / world, ops._ptr(self.gnorm_sq),
                                   c.max_grad_norm, ops._stream()), "adamw")


def diffusion_loss(model, x_t: Tensor, timesteps: Tensor, text_condition: Tensor, noise: Tensor,
                   guidance_scale: float = 7.5) -> Tensor:
    """MSE(pred_noise(x_t, t, ctx, guidance).float(), noise.float()) (train_unet.py:524-529) with the graph kept."""
    with _pack.graph_mode():
        pred = model.pred_noise(x_t, timesteps, text_condition, guidance_scale=guidance_scale)
        return ops.mse_loss(pred, noise)


def train_step(model, opt: FlatAdamW, x_t: Tensor, timesteps: Tensor, text_condition: Tensor, noise: Tensor,
               guidance_scale: float = 7.5, sync: bool = True, lr: Optional[float] = None, group=None) -> Tensor:
    """One micro-step: loss, backward into the flat gradient; with ``sync`` (the accumulation boundary,
    train_unet.py:359 ``accelerator.accumulate``) also the gradient all-reduce, clip and AdamW update, then the
    gradient buffer is cleared.  Returns the (device) loss of this rank's micro-batch."""
    loss = diffusion_loss(model, x_t, timesteps, text_condition, noise, guidance_scale)
    loss.backward()
    if sync:
        world = opt.all_reduce(group)
        opt.step(world, lr)
        opt.zero_grad()
    return loss.detach()
