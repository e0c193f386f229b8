"""Row T1 measurement: UNet train step (loss + backward + grad all-reduce + clip + AdamW) on synthetic data.

    python tools/bench_train.py --batch 64 --size 8            # the reference's default training latent
    torchrun --nproc-per-node 2 tools/bench_train.py ...        # data parallel, flat-bucket NCCL all-reduce

Prints one JSON line (samples/s over all ranks, device-timed, max over ranks).  PROFILE=1 adds the top kernels of
one step (torch profiler, CUDA activities only).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class StubText:
    """Stand-in for the CLIP wrapper (out of scope): a fixed 'empty prompt' embedding."""

    def __init__(self, ctx_len, dim, device):
        g = torch.Generator().manual_seed(7)
        self.uncond = torch.randn(1, ctx_len, dim, generator=g).to(device)

    def tokenize(self, prompts):
        class T:
            pass
        t = T()
        t.input_ids = torch.zeros(len(prompts), self.uncond.shape[1], dtype=torch.long)
        return t

    def encode_text(self, ids):
        return (self.uncond.expand(ids.shape[0], -1, -1).clone(),)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64, help="per-GPU micro-batch (C4: 64)")
    ap.add_argument("--size", type=int, default=8, help="latent height = width")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--guidance", type=float, default=7.5)
    ap.add_argument("--graph", action="store_true", help="replay the step from CUDA graphs (GraphedTrainStep)")
    args = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl")
    dev = torch.device("cuda", local)

    from stable_diffusion_pytorch_b200 import ops
    from stable_diffusion_pytorch_b200.models import (DDPMConfig, DDPMScheduler, LatentDiffusion, UNetModel, UnetConfig)
    from stable_diffusion_pytorch_b200.training import FlatAdamW, GraphedTrainStep, train_step

    torch.manual_seed(0)
    unet = UNetModel(4, 32, UnetConfig(dropout=0.0)).train()
    # the reference zero-initialises half the convs (zero_module): give them values so no gradient is trivially zero
    with torch.no_grad():
        for p in unet.parameters():
            if float(p.abs().sum()) == 0.0:
                p.normal_(0.0, 0.02)
    unet.to(dev)
    ld = LatentDiffusion(unet, None, StubText(77, 768, dev), DDPMScheduler(DDPMConfig()))
    opt = FlatAdamW(unet.parameters())
    B, S = args.batch, args.size
    g = torch.Generator().manual_seed(100 + rank)
    x_t = torch.randn(B, 4, S, S, generator=g).to(dev)
    noise = torch.randn(B, 4, S, S, generator=g).to(dev)
    t = torch.randint(0, 1000, (B,), generator=g).to(dev)
    ctx = torch.randn(B, 77, 768, generator=g).to(dev)

    gts = None
    if args.graph:
        gts = GraphedTrainStep(ld, opt, B, (S, S), guidance_scale=args.guidance)
        gts.capture()

    def step():
        if gts is not None:
            return gts.step(x_t, t, ctx, noise)
        return train_step(ld, opt, x_t, t, ctx, noise, guidance_scale=args.guidance)

    for _ in range(args.warmup):
        loss = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    launches = (ops.launch_count() - l0) // args.steps
    if world > 1:
        tt = torch.tensor([ms], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt)
    if rank == 0:
        print(json.dumps({
            "metric": "unet_train_samples_per_s", "value": round(B * world / (ms * 1e-3), 2), "unit": "samples/s",
            "n_gpus": world, "ms_per_step": round(ms, 3), "steps": args.steps, "warmup": args.warmup,
            "dtype": "bf16 compute, fp32 master weights / grads / AdamW", "data": "synthetic",
            "config": {"workload": f"T1 train step: repo-default UNet, per-GPU batch {B}, latent {S}x{S}, CFG doubling "
                                   f"(guidance {args.guidance}), MSE, clip 0.1, AdamW", "params": opt.numel},
            "loss": round(float(loss), 5), "sdb_launches_per_step": int(launches), "cuda_graph": bool(args.graph),
            "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2)}))
    if os.environ.get("PROFILE") and rank == 0:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            step()
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=70))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
