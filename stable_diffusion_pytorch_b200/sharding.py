"""Latent-batch sharding for multi-GPU sampling: every latent's trajectory is independent,
the weights are replicated, so rank r simply owns latents [lo, hi) of the global batch.  There is no
collective in the sampling loop; the only exchange is the final gather of the decoded images.

Seeded parity with a single-device run: the reference draws every step's noise for the WHOLE batch from one
CPU generator (scheduler.py:203-208).  ``host_noise_table`` therefore draws the full-batch noise with the
same seed on every rank and keeps only the rank's rows, so the union of the shards equals the single-device
result bit for bit.

Works on any ``torch.distributed`` backend: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

Tensor = torch.Tensor


def shard_range(total: int, world: int, rank: int) -> Tuple[int, int]:
    """Rows [lo, hi) of a batch of ``total`` owned by ``rank``: contiguous, sizes differ by at most one
    (the first ``total % world`` ranks take the extra row)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(total: int, world: int) -> List[int]:
    return [hi - lo for lo, hi in (shard_range(total, world, r) for r in range(world))]


def shard_batch(x: Tensor, world: int, rank: int) -> Tensor:
    lo, hi = shard_range(x.shape[0], world, rank)
    return x[lo:hi]


def host_noise_table(shape: Sequence[int], steps: Sequence[int], seed: int, world: int = 1, rank: int = 0,
                     repeat_noise: bool = False, dtype=torch.float32) -> Tensor:
    """Noise of a whole sampling run, [len(steps), hi - lo, C, H, W]: drawn for the FULL batch ``shape`` in
    the reference's order (one ``randn`` per step with ``step != 0``, nothing at step 0, one shared sample
    when ``repeat_noise``) and sliced to the rank's rows."""
    g = torch.Generator().manual_seed(seed)
    lo, hi = shard_range(shape[0], world, rank)
    out = torch.zeros(len(steps), hi - lo, *shape[1:], dtype=dtype)
    for i, step in enumerate(steps):
        if step == 0:
            continue
        if repeat_noise:
            z = torch.randn(1, *shape[1:], generator=g).expand(*shape)
        else:
            z = torch.randn(*shape, generator=g)
        out[i] = z[lo:hi].to(dtype)
    return out


def gather_batch(local: Tensor, total: int, group=None) -> Tensor:
    """All ranks end up with the full batch (rows in global order).  Shards may differ in size by one row:
    they are padded to the largest shard for the collective and trimmed afterwards."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = shard_sizes(total, world)
    if local.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} rows, expected {sizes[rank]} of {total}")
    biggest = max(sizes)
    if min(sizes) == 0:
        # more ranks than rows: a rank with an empty shard may not know the row shape (it decoded nothing); every
        # rank agrees on it first so the collective below sees equal buffers everywhere
        shapes = [None] * world
        dist.all_gather_object(shapes, tuple(local.shape[1:]) if local.shape[0] else None, group=group)
        row = next(s for s in shapes if s is not None)
        if local.shape[0] == 0:
            local = local.new_zeros(0, *row)
    padded = local
    if local.shape[0] < biggest:
        padded = torch.cat([local, local.new_zeros(biggest - local.shape[0], *local.shape[1:])])
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded.contiguous(), group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)])


def sharded_sample(latent_diffusion, x_full: Tensor, context_full: Tensor, world: Optional[int] = None,
                   rank: Optional[int] = None, gather: bool = True, **sample_kwargs) -> Tensor:
    """``LatentDiffusion.sample`` on this rank's rows of the global batch (inputs are the FULL batch, on any
    device); returns the gathered final latents (or the local shard with ``gather=False``)."""
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    dev = next(latent_diffusion.unet.parameters()).device
    x = shard_batch(x_full, world, rank).to(dev)
    ctx = shard_batch(context_full, world, rank).to(dev)
    if sample_kwargs.get("device_noise"):
        # the in-kernel Philox stream is keyed by (seed, step, LOCAL element index): give every shard its own key so
        # row i of different ranks does not see the same noise (statistical parity only, like device_noise itself)
        lo, _ = shard_range(x_full.shape[0], world, rank)
        sample_kwargs = dict(sample_kwargs, seed=(int(sample_kwargs.get("seed", 0)) + 0x9E3779B97F4A7C15 * (lo + 1))
                             & 0x7FFFFFFFFFFFFFFF)
    if x.shape[0] == 0:     # more ranks than latents: nothing to sample here, but the gather below is collective
        out = x.clone()
    else:
        out = latent_diffusion.sample(x, ctx, **sample_kwargs)
    return gather_batch(out, x_full.shape[0]) if gather else out


def sharded_images(latent_diffusion, x_full: Tensor, context_full: Tensor, world: Optional[int] = None,
                   rank: Optional[int] = None, gather: bool = True, **sample_kwargs) -> Tensor:
    """txt2img over the data-parallel group (scripts/txt2img.py:26-46 for a batch): every rank samples its rows of the
    global latent batch, decodes them with ``latent_diffusion.autoencoder`` and converts to uint8 on the device; the
    ONE collective of the whole job is the final all-gather of the [B, H, W, 3] uint8 images."""
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    latents = sharded_sample(latent_diffusion, x_full, context_full, world, rank, gather=False, **sample_kwargs)
    if latents.shape[0] == 0:   # the trailing shape is learnt from the other ranks inside gather_batch
        images = torch.zeros(0, dtype=torch.uint8, device=latents.device)
    else:
        images = latent_diffusion.autoencoder.decode_uint8(latents)
    return gather_batch(images, x_full.shape[0]) if gather else images
