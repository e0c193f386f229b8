"""LatentDiffusion pipeline with the reference's ``pred_noise`` / ``sample`` signatures
(stable_diffusion/models/latent_diffusion.py:12-126).

Reproduced on purpose: the sampling loop visits timesteps ASCENDING 0..T-1 (:103-108) and the
guidance combine treats the first batch half (the unconditional one, c_in = [uncond, ctx], :61) as
"cond": eps = e0 + g * (e0 - e1) (:62-65).  ``cfg_order="fixed"`` opts into the textbook combine.

Differences that do not change results: the empty-prompt embedding is computed once and cached
instead of every step (:55-60); ``torch.cat([x] * 2)`` is virtual (the UNet's first conv reads each
latent twice); inside ``sample`` the guidance combine and the scheduler update are one kernel.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .. import _pack, ops

Tensor = torch.Tensor


class LatentDiffusion(nn.Module):
    def __init__(self, unet, autoencoder, text_encoder, noise_scheduler):
        super().__init__()
        self.unet = unet
        self.autoencoder = autoencoder
        self.text_encoder = text_encoder
        self.noise_scheduler = noise_scheduler
        self._uncond = {}

    # ------------------------------------------------------------------------------------------
    def _uncond_emb(self, bsz: int, like: Tensor) -> Tensor:
        key = (bsz, str(like.device), like.dtype)
        emb = self._uncond.get(key)
        if emb is None:
            ids = self.text_encoder.tokenize([""] * bsz).input_ids.to(like.device)
            emb = self.text_encoder.encode_text(ids)[0].to(like.device, like.dtype).contiguous()
            self._uncond[key] = emb
        return emb

    def _c_in(self, context_emb: Tensor) -> Tensor:
        return torch.cat([self._uncond_emb(context_emb.shape[0], context_emb), context_emb])

    def pred_noise(self, noised_sample: Tensor, time_step: Tensor, context_emb: Tensor,
                   guidance_scale: float = 1.0, cfg_order: str = "reference"):
        do_cfg = guidance_scale > 1
        if not do_cfg:
            return self.unet(noised_sample, time_step, context_emb)
        eps2 = self._eps_pair(noised_sample, time_step, self._c_in(context_emb))
        if cfg_order == "fixed":  # textbook: uncond + g * (cond - uncond)
            e_unc, e_cond = eps2[: eps2.shape[0] // 2], eps2[eps2.shape[0] // 2:]
            eps2 = torch.cat([e_cond, e_unc])
            return ops.cfg_combine(eps2.contiguous(), guidance_scale - 1.0)
        return ops.cfg_combine(eps2, guidance_scale)

    def _eps_pair(self, x: Tensor, t: Tensor, c_in: Tensor) -> Tensor:
        """UNet on the doubled batch [x | x] with contexts [uncond | cond] -> [2B, C, H, W]."""
        t_in = torch.cat([t] * 2)
        return self.unet(x, t_in, c_in, x_repeat=2)

    @torch.no_grad()
    def sample(self, noised_sample: Tensor, context_emb: Tensor, guidance_scale: float = 7.5,
               repeat_noise: bool = False, scale_factor: float = 1.0, time_steps: Optional[int] = None,
               sampler: str = "ddpm", device_noise: bool = False, seed: int = 0,
               host_noise: Optional[Tensor] = None):
        """Reverse process from x_T.  ``sampler`` = "ddpm" (the reference's) or "ddim" (eta = 0).
        ``device_noise=False`` draws every step's noise on the host from torch's global CPU generator
        in the reference's order (seeded parity); True uses the in-kernel Philox stream.  ``host_noise``
        ([n_steps, B, C, H, W], e.g. ``sharding.host_noise_table``) replaces the per-step draw: a rank that owns
        a slice of a larger batch passes its rows of the full-batch noise."""
        _pack.require_cuda(noised_sample, "LatentDiffusion.sample")
        sch = self.noise_scheduler
        x = noised_sample.contiguous()
        bsz = x.shape[0]
        if time_steps is not None:
            steps = list(reversed(range(time_steps - 1, -1, -1)))      # ascending, like the reference
        else:
            steps = [int(s) for s in reversed(sch.noise_time_steps)]
        tabs = sch._device_tables(x.device)
        do_cfg = guidance_scale > 1
        c_in = self._c_in(context_emb) if do_cfg else context_emb
        kind = {"ddpm": 0, "ddim": 1}[sampler]
        for i, step in enumerate(steps):
            t = torch.full((bsz * (2 if do_cfg else 1),), step, dtype=torch.long, device=x.device)
            eps = self.unet(x, t, c_in, x_repeat=2 if do_cfg else 1)
            noise = None
            if kind == 0 and step != 0 and host_noise is not None:
                noise = host_noise[i].to(x.device, dtype=x.dtype).contiguous()
            elif kind == 0 and step != 0 and not device_noise:
                shape = (1, *x.shape[1:]) if repeat_noise else x.shape
                noise = torch.randn(shape).expand(x.shape).to(x.device, dtype=x.dtype).contiguous()
            # (the sigma entry of row 0 is 0: no noise is added at step 0, scheduler.py:200-201)
            x, _ = ops.cfg_sampler_step(eps, x, noise, tabs["coef"], slot=step, kind=kind, cfg_mode=int(do_cfg),
                                        guidance=guidance_scale, scale_factor=scale_factor, seed=seed,
                                        want_x0=False)
        return x
