"""Host-side logic of multi-GPU sampling (latent batch sharded over ranks, no collective in the loop, one
final gather): checked here on CPU with a world_size-2 gloo group."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stable_diffusion_pytorch_b200 import sharding


@pytest.mark.parametrize("total,world", [(16, 1), (16, 2), (17, 2), (5, 8), (512, 8), (3, 4)])
def test_shard_range_partitions_the_batch(total, world):
    ranges = [sharding.shard_range(total, world, r) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == total
    for (lo0, hi0), (lo1, hi1) in zip(ranges, ranges[1:]):
        assert hi0 == lo1
    sizes = [hi - lo for lo, hi in ranges]
    assert max(sizes) - min(sizes) <= 1 and sizes == sharding.shard_sizes(total, world)
    with pytest.raises(ValueError):
        sharding.shard_range(total, world, world)


@pytest.mark.parametrize("repeat", [False, True])
def test_host_noise_shards_reassemble_the_single_device_stream(repeat):
    shape, steps, seed = (5, 4, 8, 8), [0, 1, 2, 7], 1234
    full = sharding.host_noise_table(shape, steps, seed, repeat_noise=repeat)
    # the reference's draw order: one randn per step != 0 from one generator (scheduler.py:203-208)
    g = torch.Generator().manual_seed(seed)
    for i, s in enumerate(steps):
        if s == 0:
            assert full[i].abs().sum() == 0
        elif repeat:
            assert torch.equal(full[i], torch.randn(1, *shape[1:], generator=g).expand(*shape))
        else:
            assert torch.equal(full[i], torch.randn(*shape, generator=g))
    parts = [sharding.host_noise_table(shape, steps, seed, world=3, rank=r, repeat_noise=repeat) for r in range(3)]
    assert torch.equal(torch.cat(parts, dim=1), full)


class _FakeUnet(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.w = torch.nn.Parameter(torch.ones(1))


class _FakeLatentDiffusion:
    """Row-independent stand-in for LatentDiffusion.sample: x <- 0.9 x + 0.1 mean(ctx) + 0.01 noise per step."""

    class _FakeVae:
        @staticmethod
        def decode_uint8(latent):       # detransform of a 2x nearest "decode": [B, H, W, C] uint8
            img = torch.nn.functional.interpolate(latent[:, :3], scale_factor=2, mode="nearest")
            return (((img + 1) / 2).clamp(0, 1) * 255).to(torch.uint8).permute(0, 2, 3, 1).contiguous()

    def __init__(self):
        self.unet = _FakeUnet()
        self.autoencoder = self._FakeVae()

    def sample(self, x, ctx, host_noise=None, time_steps=4, **kw):
        for i in range(time_steps):
            x = 0.9 * x + 0.1 * ctx.mean(dim=(1, 2))[:, None, None, None] + 0.01 * host_noise[i]
        return x


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(7)
        x = torch.rand(total, 4, 8, 8, generator=g)
        ctx = torch.randn(total, 77, 16, generator=g)
        steps = list(range(4))
        noise = sharding.host_noise_table(tuple(x.shape), steps, 99, world, rank)
        out = sharding.sharded_sample(_FakeLatentDiffusion(), x, ctx, host_noise=noise, time_steps=4)
        ref = _FakeLatentDiffusion().sample(x, ctx, host_noise=sharding.host_noise_table(tuple(x.shape), steps, 99),
                                            time_steps=4)
        # uneven uint8 image shards through the same gather
        lo, hi = sharding.shard_range(total, world, rank)
        imgs = (torch.arange(total * 3 * 4 * 4) % 251).to(torch.uint8).reshape(total, 3, 4, 4)
        got = sharding.gather_batch(imgs[lo:hi].clone(), total)
        # the whole txt2img job: sample the shard, decode + uint8 on the rank, ONE final image gather
        pics = sharding.sharded_images(_FakeLatentDiffusion(), x, ctx, host_noise=noise, time_steps=4)
        pics_ok = pics.dtype == torch.uint8 and torch.equal(pics, _FakeLatentDiffusion._FakeVae.decode_uint8(ref))
        q.put((rank, bool(torch.equal(out, ref)), bool(torch.equal(got, imgs)) and bool(pics_ok), tuple(out.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [6, 5, 1])   # 1: rank 1 owns an empty shard and still joins the gathers
def test_sharded_sample_equals_single_process_gloo_world2(total):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, same, imgs_ok, shape in res:
        assert same, f"rank {rank}: sharded result differs from the single-process one"
        assert imgs_ok, f"rank {rank}: gathered images out of order"
        assert shape == (total, 4, 8, 8)


def test_gather_batch_is_identity_without_a_group():
    x = torch.arange(12.0).reshape(3, 4)
    assert sharding.gather_batch(x, 3) is x
