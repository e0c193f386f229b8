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

    def __init__(self, params: Iterable[nn.Parameter], cfg: Optional[OptimConfig] = None, seed: Optional[int] = None):
        """``seed``: seeds the training-mode dropout stream of this device from (seed, data-parallel rank), so ranks
        draw independent masks like the reference's per-rank CUDA generators (``ops.dropout_seed`` does it by hand)."""
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
        self.micro_steps = 0        # micro-batches accumulated into `grad` since the last update
        if seed is not None:
            rank = dist.get_rank() if (dist.is_available() and dist.is_initialized()) else 0
            ops.dropout_seed((int(seed) * 0x9E3779B1 + rank * 0x85EBCA77 + 1) & 0x7FFFFFFFFFFF, dev)
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
        self.micro_steps = 0
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

    def grad_norm(self, world: int = 1, accum_steps: Optional[int] = None) -> Tensor:
        """Global L2 norm of the (mean) gradient, on the device."""
        self.gnorm_sq.zero_()
        ops.sumsq(self.grad, self.gnorm_sq)
        return self.gnorm_sq.sqrt() / (world * (accum_steps or max(self.micro_steps, 1)))

    def step(self, world: int = 1, lr: Optional[float] = None, accum_steps: Optional[int] = None) -> None:
        """clip_grad_norm_(max_grad_norm) + AdamW on the whole model, one kernel after the norm kernel.

        The flat buffer holds the SUM over ranks and over the micro-batches accumulated since the last update; the
        reference's ``accelerator.backward`` divides every micro-loss by gradient_accumulation_steps and DDP averages
        over ranks (train_unet.py:359-366), so the gradient that is clipped and applied is that sum times
        1 / (world * accum_steps).  ``accum_steps`` defaults to the micro-steps ``train_step`` counted."""
        c = self.cfg
        accum = accum_steps or max(self.micro_steps, 1)
        self.step_count += 1
        self.gnorm_sq.zero_()
        ops.sumsq(self.grad, self.gnorm_sq)
        check(_lib.lib().sdb_adamw(ops._ptr(self.flat), ops._ptr(self.grad), ops._ptr(self.m), ops._ptr(self.v), self.numel,
                                   c.learning_rate if lr is None else lr, c.adam_beta1, c.adam_beta2, c.adam_eps,
                                   c.adam_weight_decay, self.step_count, 1.0 / (world * accum), ops._ptr(self.gnorm_sq),
                                   c.max_grad_norm, ops._stream()), "adamw")
        # the kernel wrote the parameters through raw pointers (torch's version counters did not move): packed bf16
        # copies, the cached context K/V and captured sampling graphs are stale from here on
        _pack.bump_epoch()


def diffusion_loss(model, x_t: Tensor, timesteps: Tensor, text_condition: Tensor, noise: Tensor,
                   guidance_scale: float = 7.5) -> Tensor:
    """MSE(pred_noise(x_t, t, ctx, guidance).float(), noise.float()) (train_unet.py:524-529) with the graph kept."""
    ops.dropout_begin_step(x_t.device)
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
    opt.micro_steps += 1
    ops.dropout_advance(x_t.device)          # the next micro-step draws new masks
    if sync:
        world = opt.all_reduce(group)
        opt.step(world, lr)
        opt.zero_grad()
    return loss.detach()


class GraphedTrainStep:
    """The train step replayed from CUDA graphs (the launch-bound case: at the reference's default 8x8 training
    latents a step is ~1 100 kernel launches of a few microseconds each).

    Graph A = loss + backward into the flat gradient bucket (static input tensors, weights re-packed inside the graph
    so it always sees the current parameters); the data-parallel all-reduce runs between the graphs on the same
    stream (NCCL, not captured); graph B = gradient norm + clip + AdamW + clearing the bucket, with the learning rate
    and the bias corrections of the step count read from a 3-float device buffer that is refreshed before every
    replay.  ``step(x_t, t, ctx, noise)`` copies the batch into the static tensors and returns the (device) loss.
    """

    def __init__(self, model, opt: FlatAdamW, batch: int, latent_hw, ctx_len: int = 77, ctx_dim: int = 768,
                 latent_channels: int = 4, guidance_scale: float = 7.5, group=None, warmup: int = 3,
                 accum_steps: int = 1):
        """``accum_steps``: micro-batches per update (the reference's gradient_accumulation_steps, default 4 in its
        configs): call ``step(..., sync=False)`` for the first accum_steps - 1 of them.  Fixed at capture time -- the
        1 / (world * accum_steps) gradient scale is an argument of the captured AdamW launch."""
        dev = opt.flat.device
        self.accum_steps = int(accum_steps)
        H, W = latent_hw
        self.model, self.opt, self.group, self.g = model, opt, group, guidance_scale
        self.x_t = torch.zeros(batch, latent_channels, H, W, device=dev)
        self.noise = torch.zeros_like(self.x_t)
        self.t = torch.zeros(batch, dtype=torch.long, device=dev)
        self.ctx = torch.zeros(batch, ctx_len, ctx_dim, device=dev)
        self.hyper = torch.zeros(3, device=dev)
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.graph_a = self.graph_b = None
        self.loss = None
        self._warmup = warmup

    def _fwd_bwd(self):
        loss = diffusion_loss(self.model, self.x_t, self.t, self.ctx, self.noise, self.g)
        loss.backward()
        ops.dropout_advance(self.x_t.device)   # device-side: every replay of the graph draws new masks
        return loss.detach()

    def _update(self):
        o, c = self.opt, self.opt.cfg
        o.gnorm_sq.zero_()
        ops.sumsq(o.grad, o.gnorm_sq)
        check(_lib.lib().sdb_adamw_dev(ops._ptr(o.flat), ops._ptr(o.grad), ops._ptr(o.m), ops._ptr(o.v), o.numel,
                                       ops._ptr(self.hyper), c.adam_beta1, c.adam_beta2, c.adam_eps, c.adam_weight_decay,
                                       1.0 / (self.world * self.accum_steps), ops._ptr(o.gnorm_sq), c.max_grad_norm,
                                       ops._stream()), "adamw_dev")
        o.grad.zero_()

    def _set_hyper(self, lr: Optional[float]) -> None:
        o, c = self.opt, self.opt.cfg
        o.step_count += 1
        # a fresh pageable host tensor per step: the H2D copy is staged before copy_ returns, so the next step cannot
        # overwrite values a queued copy has not read yet
        self.hyper.copy_(torch.tensor([c.learning_rate if lr is None else lr, 1.0 - c.adam_beta1 ** o.step_count,
                                       1.0 - c.adam_beta2 ** o.step_count], dtype=torch.float32))

    def capture(self) -> None:
        """Warm-up on a side stream (allocator, lazy inits), then capture the two graphs.  The warm-up steps do not
        touch the parameters: their gradients are discarded."""
        o = self.opt
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(self._warmup):
                self._fwd_bwd()
                o.grad.zero_()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        o._collect_strays()
        self.graph_a = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_a):
            self.loss = self._fwd_bwd()
        o.grad.zero_()                      # the capture pass itself does not run; keep the bucket clean anyway
        self.graph_b = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_b):
            self._update()
        torch.cuda.synchronize()

    def step(self, x_t: Tensor, timesteps: Tensor, text_condition: Tensor, noise: Tensor, sync: bool = True,
             lr: Optional[float] = None) -> Tensor:
        if self.graph_a is None:
            self.capture()
        self.x_t.copy_(x_t, non_blocking=True)
        self.t.copy_(timesteps, non_blocking=True)
        self.ctx.copy_(text_condition, non_blocking=True)
        self.noise.copy_(noise, non_blocking=True)
        self.graph_a.replay()
        if sync:
            if self.world > 1:
                dist.all_reduce(self.opt.grad, op=dist.ReduceOp.SUM, group=self.group)
            self._set_hyper(lr)
            self.graph_b.replay()
            _pack.bump_epoch()      # parameters rewritten through raw pointers: see FlatAdamW.step
        return self.loss
