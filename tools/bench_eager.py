"""The baseline SURVEY 2b / BASELINE.md name as "the number to beat": the reference's algorithm executed by torch eager
(ATen / cuDNN / cuBLAS) in bf16 on the same B200 -- here through the oracle restatement (oracle/sd_oracle.py, plain
functional PyTorch; /root/reference does not exist on the GPU box), next to this repository's kernels on the same
inputs.  Three legs: config 2 latent-step (also in bench.py's line), config 3 VAE decode, config 4 train step.

    python tools/bench_eager.py > profiles/r02_eager_baseline.json

The eager train step cannot hold config 4's batch (the materialised `sim` tensors of 64 samples x 10 attention layers,
kept for the backward, exceed 180 GB), so it runs a smaller batch and is compared per sample."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                          # noqa: E402  (build_unet, helpers)
from oracle import sd_oracle as O                                     # noqa: E402

DEV = torch.device("cuda", 0)


def events(fn, iters, warm):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def leg_latent_step():
    from stable_diffusion_pytorch_b200.graph_sampler import GraphedLatentStep
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler
    unet = bench.build_unet(DEV)
    eager = bench.gpu_eager_baseline(unet, 16, 64, steps=5, warm=2)
    gs = GraphedLatentStep(unet, DDPMScheduler(DDPMConfig()), 16, 64, 64, time_steps=50, guidance_scale=7.5,
                           sampler="ddim", dtype=torch.bfloat16, device=DEV)
    g = torch.Generator().manual_seed(5)
    gs.reset(torch.rand(16, 4, 64, 64, generator=g), torch.randn(16, 77, 768, generator=g),
             torch.randn(1, 77, 768, generator=g))
    gs.capture()
    ms = events(gs.step, 20, 5)
    return {"config": "2: 16 latents 64x64, CFG 7.5 (UNet batch 32), DDIM, bf16", "eager_ms_per_step": eager["ms_per_step"],
            "eager_latent_steps_per_s": eager["value"], "eager_peak_mem_gb": eager["peak_mem_gb"],
            "this_repo_ms_per_step": ms, "this_repo_latent_steps_per_s": 16 / (ms * 1e-3),
            "speedup": eager["ms_per_step"] / ms}


def leg_vae_decode():
    from stable_diffusion_pytorch_b200.models import AutoEncoderKL, AutoencoderConfig
    cfg = O.VaeCfg(autoencoder_channels_list=(128, 256, 512, 512), groups=32)
    sd = O.seeded_state_dict(O.vae_decoder_param_shapes(cfg), 23)
    vae = AutoEncoderKL(AutoencoderConfig(autoencoder_channels_list=[128, 256, 512, 512], groups=32)).eval()
    vae.load_state_dict(sd, strict=False)
    vae = vae.to(DEV)
    z = torch.randn(16, 4, 64, 64, generator=torch.Generator().manual_seed(3)).to(DEV, torch.bfloat16)
    sdb = {k: v.to(DEV, torch.bfloat16) for k, v in sd.items()}
    with torch.no_grad():
        ms_eager = events(lambda: O.vae_decode(sdb, cfg, z), 3, 1)
        ms_ours = events(lambda: vae.decode(z), 5, 2)
        ref = O.vae_decode({k: v.to(DEV) for k, v in sd.items()}, cfg, z[:2].float())
        e_ours = float((vae.decode(z[:2]).float() - ref).norm() / ref.norm())
        e_eager = float((O.vae_decode(sdb, cfg, z[:2]).float() - ref).norm() / ref.norm())
    return {"config": "3: VAE decode, 16 latents 64x64 -> 512x512, [128,256,512,512], bf16", "eager_ms": ms_eager,
            "this_repo_ms": ms_ours, "speedup": ms_eager / ms_ours,
            "rel_l2_vs_fp32_eager": {"this_repo": e_ours, "eager_bf16": e_eager}}


def leg_train_step(eager_batch=2):
    cfg = O.UnetCfg()
    sd = O.seeded_state_dict(O.unet_param_shapes(cfg), 13)
    params = {k: v.to(DEV).requires_grad_(True) for k, v in sd.items()}
    opt = torch.optim.AdamW(list(params.values()), lr=4e-5, weight_decay=0.1)
    g = torch.Generator().manual_seed(9)
    B = eager_batch
    x_t = torch.randn(B, 4, 64, 64, generator=g).to(DEV)
    noise = torch.randn(B, 4, 64, 64, generator=g).to(DEV)
    t = torch.randint(0, 1000, (B,), generator=g).to(DEV)
    ctx = torch.randn(B, 77, 768, generator=g).to(DEV)
    unc = torch.randn(1, 77, 768, generator=g).to(DEV).expand(B, -1, -1)

    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16):       # the reference trains under accelerate mixed precision
            loss = O.train_loss(params, cfg, x_t, t, ctx, unc, noise, 7.5)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(list(params.values()), 0.1)
        opt.step()
        opt.zero_grad(set_to_none=True)

    torch.cuda.reset_peak_memory_stats()
    ms = events(step, 3, 1)
    peak = torch.cuda.max_memory_allocated() / 2**30
    del params, opt
    torch.cuda.empty_cache()
    unet = bench.build_unet(DEV, dropout=0.1)
    ms_ours, info = bench.time_train_step(unet, DEV, 0, 32, 64)
    return {"config": "4: train step (loss, backward, clip 0.1, AdamW), 64x64 latents, CFG doubling, bf16 compute",
            "eager_batch": B, "eager_ms_per_step": ms, "eager_samples_per_s": B / (ms * 1e-3), "eager_peak_mem_gb": round(peak, 1),
            "eager_note": "autocast bf16, dropout off (the oracle restates eval-mode modules); batch 32 does not fit: "
                          "the materialised attention matrices kept for the backward exceed 180 GB",
            "this_repo_batch": 32, "this_repo_ms_per_step": ms_ours, "this_repo_samples_per_s": 32 / (ms_ours * 1e-3),
            "speedup_per_sample": (32 / ms_ours) / (B / ms)}


if __name__ == "__main__":
    torch.cuda.set_device(0)
    out = {"what": "torch-eager execution of the reference's algorithm (oracle restatement) vs this repository, one B200",
           "torch": torch.__version__, "gpu": torch.cuda.get_device_name(0)}
    for name, fn in (("latent_step", leg_latent_step), ("vae_decode", leg_vae_decode), ("train_step", leg_train_step)):
        try:
            out[name] = fn()
        except Exception as ex:      # one leg running out of memory must not lose the others
            out[name] = {"error": repr(ex)[:300]}
        torch.cuda.empty_cache()
    print(json.dumps(out, indent=1))
