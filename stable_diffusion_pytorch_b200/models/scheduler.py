"""DDPMScheduler with the reference's attributes and ``add_noise`` / ``step`` signatures
(stable_diffusion/models/scheduler.py:9-219), plus a DDIM (eta = 0) step the reference lacks.

The coefficient tables are built on the HOST with the reference's own torch expressions, in the
same order, so every entry is bit-identical (scheduler.py:52-87); they stay CPU attributes like the
reference's.  The arithmetic of ``step`` / ``add_noise`` runs in libsdb200's K7 kernels.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Sequence

import torch

from .. import _pack, ops

Tensor = torch.Tensor


@dataclass
class DDPMConfig:
    noise_schedule: str = "linear"
    noise_steps: int = 1000
    beta_start: float = 1e-4
    beta_end: float = 0.02


class DDPMScheduler:
    def __init__(self, cfg):
        self.noise_steps: int = cfg.noise_steps
        self.noise_time_steps = torch.arange(self.noise_steps - 1, -1, -1)
        # The reference advertises noise_schedule in {linear, cosine, cubic} (scheduler.py:11-17) but always builds the
        # linear one (:89-93).  Here the field is honoured; "linear" (the default) is the reference's table bit for bit.
        kind = getattr(cfg, "noise_schedule", "linear")
        if kind == "linear":
            self.betas = self.prepare_linear_noise_schedule(cfg.beta_start, cfg.beta_end)
        elif kind == "cosine":
            self.betas = self.prepare_cosine_noise_schedule()
        elif kind == "cubic":
            self.betas = self.prepare_cubic_noise_schedule(cfg.beta_start, cfg.beta_end)
        else:
            raise ValueError(f"noise_schedule must be linear, cosine or cubic, got {kind!r}")
        self.alpha = 1.0 - self.betas
        self.alphas_cumprod = torch.cumprod(self.alpha, dim=0)
        alpha_bar_prev = torch.cat([self.alphas_cumprod.new_tensor([1.0]), self.alphas_cumprod[:-1]])
        self.sqrt_alpha_bar = self.alphas_cumprod ** 0.5
        self.sqrt_1m_alpha_bar = (1.0 - self.alphas_cumprod) ** 0.5
        self.sqrt_recip_alpha_bar = self.alphas_cumprod ** -0.5
        self.sqrt_recip_m1_alpha_bar = (1 / self.alphas_cumprod - 1) ** 0.5
        variance = self.betas * (1.0 - alpha_bar_prev) / (1.0 - self.alphas_cumprod)
        self.log_var = torch.log(torch.clamp(variance, min=1e-20))
        self.mean_x0_coef = self.betas * (alpha_bar_prev ** 0.5) / (1.0 - self.alphas_cumprod)
        self.mean_xt_coef = (1.0 - alpha_bar_prev) * ((1 - self.betas) ** 0.5) / (1.0 - self.alphas_cumprod)
        # host-built derived tables for the kernels
        self._sigma = (0.5 * self.log_var).exp()                      # scheduler.py:213
        self._ddim_a = alpha_bar_prev ** 0.5                           # sqrt(abar_{t-1}), 1 at t = 0
        self._ddim_b = (1.0 - alpha_bar_prev) ** 0.5
        self._dev: Dict[str, Dict[str, Tensor]] = {}

    def prepare_linear_noise_schedule(self, beta_start, beta_end):
        return torch.linspace(beta_start, beta_end, self.noise_steps)

    def prepare_cosine_noise_schedule(self, s: float = 0.008, max_beta: float = 0.999):
        """Nichol & Dhariwal 2021: abar(t) = cos^2(((t / T + s) / (1 + s)) pi / 2), beta_t = min(1 - abar(t+1) / abar(t),
        max_beta).  Not in the reference (advertised only): parity unpinned, held to the oracle's restatement."""
        T = self.noise_steps
        t = torch.arange(T + 1, dtype=torch.float64) / T
        abar = torch.cos((t + s) / (1 + s) * torch.pi / 2) ** 2
        return torch.clamp(1 - abar[1:] / abar[:-1], max=max_beta).to(torch.float32)

    def prepare_cubic_noise_schedule(self, beta_start, beta_end):
        """beta_t = linspace(beta_start^(1/3), beta_end^(1/3), T)^3 (the cube-root analogue of the "scaled linear"
        schedule).  Not in the reference (advertised only): parity unpinned."""
        return torch.linspace(beta_start ** (1 / 3), beta_end ** (1 / 3), self.noise_steps) ** 3

    # ------------------------------------------------------------------------------------------
    def coef_rows(self, steps: Sequence[int], add_noise_at_zero: bool = False) -> Tensor:
        """fp32 [len(steps), 8] rows in the layout ``sdb_cfg_sampler_step`` documents.  sigma is 0
        for time_step == 0 (the reference adds zeros there, scheduler.py:200-201)."""
        idx = torch.as_tensor(list(steps), dtype=torch.long)
        sigma = self._sigma[idx].clone()
        if not add_noise_at_zero:
            sigma[idx == 0] = 0.0
        rows = torch.stack([self.sqrt_recip_alpha_bar[idx], self.sqrt_recip_m1_alpha_bar[idx],
                            self.mean_x0_coef[idx], self.mean_xt_coef[idx], sigma,
                            self._ddim_a[idx], self._ddim_b[idx], torch.zeros(len(idx))], dim=1)
        return rows.to(torch.float32).contiguous()

    def _device_tables(self, device) -> Dict[str, Tensor]:
        key = str(device)
        if key not in self._dev:
            self._dev[key] = {"coef": self.coef_rows(range(self.noise_steps)).to(device),
                              "ac": self.alphas_cumprod.to(device=device, dtype=torch.float32).contiguous()}
        return self._dev[key]

    def add_noise(self, original_samples: Tensor, noise: Tensor, timesteps: Tensor):
        _pack.require_cuda(original_samples, "DDPMScheduler.add_noise")
        assert timesteps.max().item() < self.noise_steps, (
            f"timesteps({timesteps}) should be less than {self.noise_steps}")
        tabs = self._device_tables(original_samples.device)
        t = timesteps.to(device=original_samples.device, dtype=torch.int64).contiguous()
        return ops.add_noise(original_samples, noise.to(original_samples.device, original_samples.dtype), t,
                             tabs["ac"])

    @torch.no_grad()
    def step(self, pred_noise: Tensor, x_t: Tensor, time_step: int, repeat_noise: bool = False,
             scale_factor: float = 1.0, noise: Optional[Tensor] = None):
        """One DDPM posterior sample (scheduler.py:141-219) -> (x_prev, x0).

        Like the reference, the noise is drawn on the HOST from torch's global CPU generator
        (``torch.randn(x_t.shape)``), so a seeded run consumes the same stream; pass ``noise`` to
        supply it explicitly."""
        assert time_step < self.noise_steps, f"timesteps({time_step}) should be less than {self.noise_steps}"
        _pack.require_cuda(x_t, "DDPMScheduler.step")
        tabs = self._device_tables(x_t.device)
        if time_step == 0:
            noise = None
        elif noise is None:
            if repeat_noise:
                noise = torch.randn((1, *x_t.shape[1:])).expand(x_t.shape)
            else:
                noise = torch.randn(x_t.shape)
        if noise is not None:
            noise = noise.to(x_t.device, dtype=x_t.dtype).contiguous()
        x_prev, x0 = ops.cfg_sampler_step(pred_noise.contiguous(), x_t.contiguous(), noise, tabs["coef"],
                                          slot=int(time_step), kind=0, cfg_mode=0, scale_factor=scale_factor)
        return x_prev, x0

    @torch.no_grad()
    def ddim_step(self, pred_noise: Tensor, x_t: Tensor, time_step: int):
        """eta = 0 DDIM update on the same tables (not in the reference; parity unpinned):
        x0 as scheduler.py:184, x_prev = sqrt(abar_{t-1}) x0 + sqrt(1 - abar_{t-1}) eps."""
        assert time_step < self.noise_steps
        tabs = self._device_tables(x_t.device)
        return ops.cfg_sampler_step(pred_noise.contiguous(), x_t.contiguous(), None, tabs["coef"],
                                    slot=int(time_step), kind=1, cfg_mode=0)
