"""CUDA-graph replay of one latent-step (CFG pair of UNet evaluations + guidance combine + sampler
update) with every per-step scalar read from device memory, so ONE captured graph serves all steps
of a sampling run.

Static device state:
  x      [B, C, H, W]   latent, updated in place by the fused K7 kernel
  c_in   [2B, L, D]     contexts in the reference's order [uncond | cond] (latent_diffusion.py:61)
  t      [2B] int64     current timestep, rewritten by ``sdb_sampler_advance``
  slot   int32          index into the coefficient / timestep / noise tables
The timesteps visit ``steps`` in the given order (the reference's ascending 0..T-1 by default).
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from . import ops

Tensor = torch.Tensor


class GraphedLatentStep:
    def __init__(self, unet, scheduler, batch: int, height: int, width: int, ctx_len: int = 77,
                 steps: Optional[Sequence[int]] = None, time_steps: int = 50, guidance_scale: float = 7.5,
                 sampler: str = "ddim", dtype=torch.bfloat16, scale_factor: float = 1.0, seed: int = 0,
                 device="cuda", host_noise: Optional[Tensor] = None, ctx_dtype=None):
        self.unet, self.sch = unet, scheduler
        self.B, self.H, self.W = batch, height, width
        self.C = unet.latent_channels
        self.device = torch.device(device)
        self.dtype = dtype
        self.g = float(guidance_scale)
        self.cfg = self.g > 1
        self.kind = {"ddpm": 0, "ddim": 1}[sampler]
        self.scale_factor, self.seed = float(scale_factor), int(seed)
        self.steps = list(steps) if steps is not None else list(reversed(range(time_steps - 1, -1, -1)))
        n = len(self.steps)
        rep = 2 if self.cfg else 1
        dev = self.device
        self.x = torch.zeros(batch, self.C, height, width, device=dev, dtype=dtype)
        self.c_in = torch.zeros(rep * batch, ctx_len, unet.context_dim, device=dev, dtype=ctx_dtype or dtype)
        self.t_table = torch.tensor(self.steps, dtype=torch.int64, device=dev)
        self.t = torch.full((rep * batch,), self.steps[0], dtype=torch.int64, device=dev)
        self.slot = torch.zeros(1, dtype=torch.int32, device=dev)
        self.coef = scheduler.coef_rows(self.steps).to(dev)
        # optional pre-drawn host noise table [n_steps, B, C, H, W] (seeded parity with the reference)
        self.noise = None if host_noise is None else host_noise.to(dev, dtype).contiguous()
        if self.noise is not None:
            assert self.noise.shape[0] == n
        self.graph = None
        self.launches_per_step = 0

    def set_noise(self, host_noise: Tensor):
        """Refills the static noise table [n_steps, B, C, H, W] a captured graph reads (same storage, new draws)."""
        assert self.noise is not None and tuple(host_noise.shape) == tuple(self.noise.shape)
        self.noise.copy_(host_noise, non_blocking=True)

    # ------------------------------------------------------------------------------------------
    def _one_step(self):
        eps = self.unet(self.x, self.t, self.c_in, x_repeat=2 if self.cfg else 1)
        ops.cfg_sampler_step(eps, self.x, self.noise, self.coef, slot_dev=self.slot, kind=self.kind,
                             cfg_mode=int(self.cfg), guidance=self.g, scale_factor=self.scale_factor,
                             seed=self.seed, x_prev=self.x, want_x0=False)
        ops.sampler_advance(self.slot, self.t_table, self.t)

    def reset(self, x: Optional[Tensor] = None, context: Optional[Tensor] = None,
              uncond: Optional[Tensor] = None):
        """(Re)loads the static buffers and rewinds the step slot.  Tensors may live on the host."""
        if x is not None:
            self.x.copy_(x, non_blocking=True)
        if context is not None:
            if self.cfg:
                self.c_in[self.B:].copy_(context, non_blocking=True)
            else:
                self.c_in.copy_(context, non_blocking=True)
        if uncond is not None and self.cfg:
            self.c_in[: self.B].copy_(uncond.expand(self.B, -1, -1), non_blocking=True)
        self.slot.zero_()
        self.t.fill_(self.steps[0])

    @torch.no_grad()
    def capture(self):
        """Warm up eagerly on a side stream (lazy attribute setting / weight packing must not happen
        under capture), then capture one latent-step."""
        saved = (self.x.clone(), self.slot.clone(), self.t.clone())
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            for _ in range(2):
                self._one_step()
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        before = ops.launch_count()
        with torch.cuda.graph(self.graph):
            self._one_step()
        self.launches_per_step = ops.launch_count() - before
        self.x.copy_(saved[0]); self.slot.copy_(saved[1]); self.t.copy_(saved[2])
        torch.cuda.synchronize(self.device)
        return self

    def step(self):
        """Enqueues one latent-step on the current stream."""
        if self.graph is None:
            self._one_step()
        else:
            self.graph.replay()

    @torch.no_grad()
    def run(self, x: Tensor, context: Tensor, uncond: Optional[Tensor] = None) -> Tensor:
        """Full sampling run over ``steps``; returns the final latents (a clone of the static x)."""
        self.reset(x, context, uncond)
        for _ in self.steps:
            self.step()
        return self.x.clone()

    @torch.no_grad()
    def step_host(self, x_host: Tensor, ctx_host: Tensor, out_host: Tensor):
        """End-to-end single step with HOST (pinned) buffers: H2D latent + context, one latent-step,
        D2H of the updated latent.  Used by bench.py's e2e leg."""
        # The PCIe copies run on a side stream through device staging buffers, so the H2D of this call overlaps the
        # kernels of the previous one and the D2H of its result overlaps the next one's; the hand-overs to the buffers the
        # captured graph reads and writes are device-to-device copies on the main stream.  Every call still moves its own
        # inputs from, and its own result to, host memory.
        c_dst = self.c_in[self.B:] if self.cfg else self.c_in
        main = torch.cuda.current_stream(self.device)
        if getattr(self, "_io", None) is None:
            self._io = {
                "stream": torch.cuda.Stream(device=self.device),
                "x": torch.empty_like(self.x), "c": torch.empty_like(c_dst), "o": torch.empty_like(self.x),
                "in_free": torch.cuda.Event(), "in_ready": torch.cuda.Event(),
                "out_ready": torch.cuda.Event(), "out_free": torch.cuda.Event(),
            }
            self._io["in_free"].record(main)
            self._io["out_free"].record(main)
        io = self._io
        cs = io["stream"]
        cs.wait_event(io["in_free"])                     # the previous call has taken its inputs out of the staging
        with torch.cuda.stream(cs):
            io["x"].copy_(x_host, non_blocking=True)
            io["c"].copy_(ctx_host, non_blocking=True)
            io["in_ready"].record(cs)
        main.wait_event(io["in_ready"])
        self.x.copy_(io["x"])
        c_dst.copy_(io["c"])
        io["in_free"].record(main)
        self.step()
        main.wait_event(io["out_free"])                  # the previous result has left the output staging
        io["o"].copy_(self.x)
        io["out_ready"].record(main)
        cs.wait_event(io["out_ready"])
        with torch.cuda.stream(cs):
            out_host.copy_(io["o"], non_blocking=True)
            io["out_free"].record(cs)

    def join_host(self):
        """Orders the current stream behind every copy ``step_host`` has enqueued (no host wait): an event recorded
        afterwards covers the last result's trip to host memory."""
        if getattr(self, "_io", None) is not None:
            torch.cuda.current_stream(self.device).wait_stream(self._io["stream"])

    def sync_host(self):
        """Wait until the results of all ``step_host`` calls have landed in their host buffers."""
        self.join_host()
        torch.cuda.current_stream(self.device).synchronize()
