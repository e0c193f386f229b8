"""Multi-GPU check of the sampling path's one collective: `sharding.sharded_images` -- every rank samples
its rows of the global latent batch with the real LatentDiffusion (DDPM + CFG, full-batch host noise table), decodes
them to uint8 on the device and all-gathers the images over NCCL -- against the single-GPU run of the whole batch.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
        tools/check_sharded_nccl.py

Prints one JSON line on rank 0; exits non-zero on a mismatch."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import sd_oracle as O                                     # noqa: E402  (seeded weights only)
from stable_diffusion_pytorch_b200 import sharding                    # noqa: E402
from stable_diffusion_pytorch_b200.models import (AutoEncoderKL, AutoencoderConfig, DDPMConfig, DDPMScheduler,  # noqa: E402
                                                  LatentDiffusion, UNetModel, UnetConfig)
from tests.golden import cases                                        # noqa: E402


class StubText:
    def __init__(self, uncond):
        self.uncond = uncond

    def tokenize(self, prompts):
        class T:
            pass
        t = T()
        t.input_ids = torch.zeros(len(prompts), self.uncond.shape[1], dtype=torch.long)
        return t

    def encode_text(self, ids):
        return (self.uncond.to(ids.device).expand(ids.shape[0], -1, -1).clone(),)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cfg, seed = cases.TINY, 11
    total, H, W, steps, g = 5, 16, 16, 6, 7.5          # 5 latents over `world` ranks: uneven shards
    unet = UNetModel(cfg.latent_channels, cfg.groups, UnetConfig(
        num_res_blocks=cfg.num_res_blocks, n_heads=cfg.n_heads, attention_resolutions=list(cfg.attention_resolutions),
        channels_list=list(cfg.channels_list), time_emb_dim=cfg.time_emb_dim, dropout=0.0, n_layers=cfg.n_layers,
        context_dim=cfg.context_dim)).eval()
    unet.load_state_dict(O.seeded_state_dict(O.unet_param_shapes(cfg), seed), strict=True)
    vcfg = cases.VAE_DEFAULT
    vae = AutoEncoderKL(AutoencoderConfig(autoencoder_channels_list=list(vcfg.autoencoder_channels_list),
                                          groups=vcfg.groups)).eval()
    vae.load_state_dict(O.seeded_state_dict(O.vae_decoder_param_shapes(vcfg), 21), strict=False)
    x, ctx, uncond = cases.sample_inputs(cfg, total, H, W, seed)
    ld = LatentDiffusion(unet.to(dev), vae.to(dev), StubText(uncond), DDPMScheduler(DDPMConfig()))
    step_list = list(range(steps))
    noise = sharding.host_noise_table(tuple(x.shape), step_list, 99, world, rank)
    imgs = sharding.sharded_images(ld, x, ctx, host_noise=noise, time_steps=steps, guidance_scale=g)
    lat = sharding.sharded_sample(ld, x, ctx, host_noise=noise, time_steps=steps, guidance_scale=g)
    # the same job on this GPU alone
    full_noise = sharding.host_noise_table(tuple(x.shape), step_list, 99)
    ref_lat = ld.sample(x.to(dev), ctx.to(dev), guidance_scale=g, time_steps=steps, host_noise=full_noise)
    ref_img = vae.decode_uint8(ref_lat)
    ok_lat = bool(torch.equal(lat, ref_lat))
    ok_img = bool(torch.equal(imgs, ref_img))
    diff = int((imgs.int() - ref_img.int()).abs().max())
    flags = torch.tensor([int(ok_lat), int(ok_img)], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"check": "sharded_images over NCCL == single-GPU run", "world": world, "latents": total,
                          "shards": sharding.shard_sizes(total, world), "steps": steps, "sampler": "ddpm + CFG 7.5",
                          "images": list(imgs.shape), "latents_bit_equal_all_ranks": bool(flags[0]),
                          "images_bit_equal_all_ranks": bool(flags[1]), "max_abs_uint8_diff_rank0": diff,
                          "backend": dist.get_backend()}), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flags.min()) == 1 else 1)


if __name__ == "__main__":
    main()
