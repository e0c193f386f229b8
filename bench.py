#!/usr/bin/env python
"""Benchmark of the latent-diffusion hot path (BASELINE.json config 2 at N = 1).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one latent-step of the workload batch: the CFG pair of UNet evaluations (UNet batch
2*B), the guidance combine and the sampler update, i.e. one iteration of the reference's sampling
loop (models/latent_diffusion.py:109-123).  Metric: latent-steps/s = B * n_gpus * K / time.
Prints ONE JSON line on rank 0 (see the contract in the task statement / DESIGN.md section 6).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = "unet-default bs16 64x64x4 latents, CFG g=7.5 (UNet batch 32), DDIM eta=0, bf16"
GF_PER_LATENT_STEP = 453.46  # 2 UNet evals @64x64, reference formulation
MUFU_EX2_PER_CLK_PER_SM = 15.84  # measured ex2.approx rate (tools/ubench.cu, profiles/r01_ubench_mufu_fma.txt)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=16, help="latents per GPU (weak scaling)")
    ap.add_argument("--size", type=int, default=64, help="latent height = width")
    ap.add_argument("--sampler", default="ddim", choices=["ddim", "ddpm"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-decode", action="store_true", help="skip the VAE-decode (images/s) leg")
    ap.add_argument("--no-train", action="store_true", help="skip the config-4 train-step leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the 50-step parity check and the torch-eager GPU baseline")
    ap.add_argument("--train-batch", type=int, default=32, help="per-GPU batch of the train-step leg (config 4: 32)")
    ap.add_argument("--profile-step", action="store_true",
                    help="after warm-up run ONE eager step between cudaProfilerStart/Stop and exit "
                         "(for `ncu --profile-from-start off`; prints no bench line)")
    return ap.parse_args()


def peaks():
    p = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            d = json.load(open(path))
            p.update({k: d[k] for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained") if k in d})
            p["source"] = "measured"
        except Exception:
            pass
    return p


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "samples": len(sm),
                "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------
# CPU legs: the oracle port of the reference's algorithm, timed on the host cores (the only place
# bench.py executes oracle/).
# ----------------------------------------------------------------------------------------------
def cpu_latent_steps(size: int, steps: int, warmup: int, budget_s: float = 25.0):
    import torch
    from oracle import sd_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = O.UnetCfg()
    sd = O.seeded_state_dict(O.unet_param_shapes(cfg), 13)
    tab = O.ddpm_tables(O.DdpmCfg())
    g = torch.Generator().manual_seed(0)
    x = torch.rand(1, 4, size, size, generator=g)
    ctx = torch.randn(1, 77, 768, generator=g)
    unc = torch.randn(1, 77, 768, generator=g)
    fn = lambda a, b, c: O.unet_forward(sd, cfg, a, b, c)  # noqa: E731
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            t = torch.full((1,), i, dtype=torch.long)
            eps = O.pred_noise(fn, x, t, ctx, unc, 7.5)
            x, _ = O.ddim_step(tab, eps, x, i)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
            if i >= warmup and sum(times) > budget_s:
                break
    n = len(times)
    total = sum(times)
    return {"value": n / total, "unit": "latent-steps/s", "cores": cores, "kind": "port",
            "sample": f"{n} CFG latent-steps of ONE {size}x{size} latent (UNet batch 2, fp32, torch CPU "
                      f"{torch.get_num_threads()} threads) = 1/16 of the workload batch",
            "ms_per_step": 1e3 * total / n, "steps": n}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    res = cpu_latent_steps(args.size, max(1, args.steps), max(0, min(args.warmup, 1)))
    line = {"impl": "reference", "metric": "latent_steps_per_sec", "value": res["value"],
            "unit": "latent-steps/s", "n_gpus": args.gpus, "steps": res["steps"], "warmup": min(args.warmup, 1),
            "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "note": "oracle port of the reference's CPU path; each step is a "
                                                      "bounded sample (1 latent) of the workload batch"},
            "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": res["value"], "unit": "latent-steps/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------------------------
def build_unet(device, dropout=0.0):
    """Repo-default UNet, random init (reference init + every all-zero tensor re-randomised so the
    zero_module'd convs do real work; weights differ from any checkpoint: data = synthetic).  ``dropout`` only
    matters in train mode: the sampling legs run eval(), the train leg uses the reference default 0.1."""
    import torch
    from stable_diffusion_pytorch_b200.models import UNetModel, UnetConfig
    torch.manual_seed(0)
    m = UNetModel(4, 32, UnetConfig(dropout=dropout)).eval()
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for p in m.parameters():
            if float(p.abs().sum()) == 0.0:
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    return m.to(device)


def time_kernel(fn, iters=10, warm=3):
    import torch
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters  # ms


def time_vae_decode(latents, dev, iters=3):
    """ms per decode of `latents` [B,4,64,64] with the AutoencoderKL-shape decoder (random init)."""
    import torch
    from stable_diffusion_pytorch_b200.models import AutoEncoderKL, AutoencoderConfig
    g = torch.Generator().manual_seed(11)
    vae = AutoEncoderKL(AutoencoderConfig(autoencoder_channels_list=[128, 256, 512, 512], groups=32)).eval()
    with torch.no_grad():
        for p in vae.parameters():
            if p.abs().sum() == 0:
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    vae = vae.to(dev)
    z = latents.detach().clone()
    chunks = list(z.split(16))      # config 3 is a batch of 16; larger shards decode 16 at a time (2 GB activations each)
    with torch.no_grad():
        for _ in range(2):
            vae.decode_uint8(chunks[0])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            u8 = torch.cat([vae.decode_uint8(c) for c in chunks]) if len(chunks) > 1 else vae.decode_uint8(chunks[0])
        e1.record()
        torch.cuda.synchronize()
    del vae
    torch.cuda.empty_cache()
    return e0.elapsed_time(e1) / iters, u8


def time_image_gather(u8, world, iters=3):
    """The ONE collective of the sampling job: all-gather of the decoded [B, 512, 512, 3] uint8 images
    over NCCL; ms per gather (CUDA events on this rank; the caller takes the max over ranks)."""
    import torch
    from stable_diffusion_pytorch_b200 import sharding
    total = u8.shape[0] * world
    out = sharding.gather_batch(u8, total)      # warm-up (communicator set-up)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        out = sharding.gather_batch(u8, total)
    e1.record()
    torch.cuda.synchronize()
    assert out.shape[0] == total and out.dtype == torch.uint8
    return e0.elapsed_time(e1) / iters


def time_train_step(unet, dev, rank, batch, size, iters=3, warm=2):
    """ms per train step (tools/bench_train.py's loop) on this rank; all ranks call it together (the step holds the
    gradient all-reduce)."""
    import torch
    from stable_diffusion_pytorch_b200 import ops
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler, LatentDiffusion
    from stable_diffusion_pytorch_b200.training import FlatAdamW, GraphedTrainStep

    class StubText:   # the empty-prompt embedding pred_noise asks for (CLIP weights are a download)
        def __init__(self):
            self.uncond = torch.randn(1, 77, 768, generator=torch.Generator().manual_seed(7)).to(dev)

        def tokenize(self, prompts):
            class T:
                pass
            t = T()
            t.input_ids = torch.zeros(len(prompts), 77, dtype=torch.long)
            return t

        def encode_text(self, ids):
            return (self.uncond.expand(ids.shape[0], -1, -1).clone(),)

    unet.train()
    ld = LatentDiffusion(unet, None, StubText(), DDPMScheduler(DDPMConfig()))
    opt = FlatAdamW(unet.parameters())
    g = torch.Generator().manual_seed(200 + rank)
    x_t = torch.randn(batch, 4, size, size, generator=g).to(dev)
    noise = torch.randn(batch, 4, size, size, generator=g).to(dev)
    t = torch.randint(0, 1000, (batch,), generator=g).to(dev)
    ctx = torch.randn(batch, 77, 768, generator=g).to(dev)
    gts = GraphedTrainStep(ld, opt, batch, (size, size), guidance_scale=7.5)   # fwd+bwd graph, NCCL, update graph
    gts.capture()
    for _ in range(warm):
        loss = gts.step(x_t, t, ctx, noise)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        loss = gts.step(x_t, t, ctx, noise)
    e1.record()
    torch.cuda.synchronize()
    info = {"steps": iters, "warmup": warm, "loss": float(loss), "cuda_graph": True,
            "params": opt.numel, "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2)}
    return e0.elapsed_time(e1) / iters, info


def oracle_state_dict(unet, device, dtype):
    return {k: v.detach().to(device=device, dtype=dtype) for k, v in unet.state_dict().items()}


def gpu_eager_baseline(unet, B, size, steps=3, warm=1):
    """The number to beat ("the unmodified reference running torch-eager in bf16 on the same B200"):
    the oracle restatement of the reference's algorithm -- pred_noise with CFG + DDIM update, every op an ATen /
    cuDNN / cuBLAS call, `sim` materialised -- on the same GPU, same batch, bf16, CUDA events.  A baseline leg:
    it imports oracle/, outside the timed b200 region."""
    import torch
    from oracle import sd_oracle as O
    dev = next(unet.parameters()).device
    cfg = O.UnetCfg()
    sd = oracle_state_dict(unet, dev, torch.bfloat16)
    tab = O.ddpm_tables(O.DdpmCfg())
    g = torch.Generator().manual_seed(5)
    x = torch.rand(B, 4, size, size, generator=g).to(dev, torch.bfloat16)
    ctx = torch.randn(B, 77, 768, generator=g).to(dev, torch.bfloat16)
    unc = torch.randn(1, 77, 768, generator=g).to(dev, torch.bfloat16).expand(B, -1, -1)
    fn = lambda a, b, c: O.unet_forward(sd, cfg, a, b, c)  # noqa: E731
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.no_grad():
        for i in range(warm + steps):
            if i == warm:
                torch.cuda.synchronize()
                e0.record()
            t = torch.full((B,), i, dtype=torch.long, device=dev)
            eps = O.pred_noise(fn, x, t, ctx, unc, 7.5)
            x, _ = O.ddim_step(tab, eps, x, i)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    peak = torch.cuda.max_memory_allocated() / 2**30
    del sd, x, ctx, unc, eps
    torch.cuda.empty_cache()
    return {"value": B / (ms * 1e-3), "unit": "latent-steps/s", "ms_per_step": ms, "steps": steps, "warmup": warm,
            "kind": "oracle restatement of the reference on torch eager (ATen/cuDNN/cuBLAS), bf16, same B200, "
                    "same batch; `sim` materialised like the reference", "peak_mem_gb": round(peak, 1)}


def sample_parity(gs, unet, x_host, ctx_host, unc, n_check=2):
    """rel-L2 of the final latents of the benchmarked path (the captured graph, 50 DDIM steps with CFG, bf16) against
    the oracle restatement run by torch eager in FP32 on the same GPU, for the first ``n_check`` latents of the batch
    (trajectories are independent per latent).  tests/test_graphed_sampler_gpu.py ties that fp32-on-GPU run to the CPU
    oracle (< 1e-4)."""
    import torch
    from oracle import sd_oracle as O
    dev = gs.device
    tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        got = gs.run(x_host, ctx_host, unc)[:n_check].float()
        cfg = O.UnetCfg()
        sd = oracle_state_dict(unet, dev, torch.float32)
        tab = O.ddpm_tables(O.DdpmCfg())
        fn = lambda a, b, c: O.unet_forward(sd, cfg, a, b, c)  # noqa: E731
        with torch.no_grad():
            ref = O.sample(fn, tab, x_host[:n_check].float().to(dev), ctx_host[:n_check].float().to(dev),
                           unc.float().to(dev).expand(n_check, -1, -1), 7.5, time_steps=len(gs.steps), sampler="ddim")
        err = float((got.double() - ref.double()).norm() / ref.double().norm())
        # context for the tolerance: the same restatement run by torch eager in bf16 (what the reference does with
        # --dtype bfloat16) drifts from its own fp32 run by this much over the same 50 steps
        sdb = oracle_state_dict(unet, dev, torch.bfloat16)
        fnb = lambda a, b, c: O.unet_forward(sdb, cfg, a, b, c)  # noqa: E731
        with torch.no_grad():
            refb = O.sample(fnb, tab, x_host[:n_check].to(dev), ctx_host[:n_check].to(dev),
                            unc.to(dev).expand(n_check, -1, -1), 7.5, time_steps=len(gs.steps), sampler="ddim")
        err_eager = float((refb.double() - ref.double()).norm() / ref.double().norm())
        del sd, sdb, ref, refb
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.cuda.empty_cache()
    return {"rel_l2_final_latents": err, "steps": len(gs.steps), "latents_checked": n_check,
            "against": "oracle restatement, fp32 torch eager on the same GPU (TF32 off)", "tolerance": 1e-1,
            "torch_eager_bf16_rel_l2": err_eager,
            "note": "tolerance for 50 bf16 CFG steps at 64x64 (guidance 7.5 amplifies e0 - e1 every step): 1e-1; the "
                    "oracle run in bf16 by torch eager is listed beside it for scale"}


def hbm_kernels(B2, size, pk):
    """The HBM-bound kernels of the step timed alone at the workload's shapes: achieved GB/s of algorithmic bytes
    against the measured copy peak (north_star: norm and sampler kernels against the HBM roofline)."""
    import torch
    from stable_diffusion_pytorch_b200 import ops
    from stable_diffusion_pytorch_b200._lib import ACT_SILU
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler
    dev = "cuda"
    out = []
    x = torch.randn(B2, size, size, 160, device=dev).to(torch.bfloat16)
    gam, bet = torch.ones(160, device=dev), torch.zeros(160, device=dev)
    scr = ops.GroupNormScratch()
    nbytes = x.numel() * 2.0
    ms = time_kernel(lambda: ops.group_norm(x, gam, bet, 32, ACT_SILU, scratch=scr))
    out.append({"kernel": f"GroupNorm(+SiLU) [{B2},{size},{size},160] bf16, one cooperative launch (gn_coop_kernel: statistics pass, "
                          f"grid barrier, normalise pass out of L2)",
                "ms": ms, "bytes": 2 * nbytes, "note": "algorithmic bytes = one read + one write of the tensor"})
    tok = x.view(B2, size * size, 160)
    res = torch.randn_like(tok)
    ms = time_kernel(lambda: ops.add_layernorm(tok, res, gam, bet))
    out.append({"kernel": f"add_layernorm [{B2 * size * size},160] bf16 (stand-alone kernel; the UNet path runs LN in the "
                          f"GEMM epilogue)", "ms": ms, "bytes": 3 * nbytes})
    sch = DDPMScheduler(DDPMConfig())
    coef = sch.coef_rows(range(50)).to(dev)
    for Bs in (B2 // 2, 512):
        eps = torch.randn(2 * Bs, 4, size, size, device=dev).to(torch.bfloat16)
        xt = torch.randn(Bs, 4, size, size, device=dev).to(torch.bfloat16)
        ms = time_kernel(lambda: ops.cfg_sampler_step(eps, xt, None, coef, slot=7, kind=1, cfg_mode=1, guidance=7.5,
                                                      x_prev=xt, want_x0=False))
        out.append({"kernel": f"cfg_sampler_step DDIM+CFG in place, {Bs} latents bf16", "ms": ms,
                    "bytes": xt.numel() * 2.0 * 4})      # eps x2 + x_t read, x_prev written
    for o in out:
        o["gbs"] = o["bytes"] / (o["ms"] * 1e-3) / 1e9
        o["frac_of_hbm_peak"] = o["gbs"] / pk["hbm_gbs"]
    return out


def kernel_rooflines(B2, size, pk):
    """Isolated timings (CUDA events, current stream) of the three dominant kernel families at the
    workload's shapes; algorithmic flops in the reference's formulation."""
    import math
    import torch
    from stable_diffusion_pytorch_b200 import ops
    dev = "cuda"
    n = size * size
    out = []
    # K4 self-attention, level 0: [B2*8 heads, n, n, d=20] on head-padded operands (32 columns per head:
    # q/k zero beyond d, v carries the ones column that yields the softmax denominator)
    def padded(ones):
        t = torch.zeros(B2, n, 8, 32, device=dev)
        t[..., :20] = torch.randn(B2, n, 8, 20, device=dev)
        if ones:
            t[..., 20] = 1.0
        return t.reshape(B2, n, 256).to(torch.bfloat16)
    q, k, v = padded(False), padded(False), padded(True)
    ms = time_kernel(lambda: ops.attention_p32(q, k, v, 8, 20, 1 / math.sqrt(20)))
    fl = 4.0 * B2 * 8 * n * n * 20
    out.append({"kernel": "attention_umma_wg1_kernel<2,4,spread> self-attn d=20 (tcgen05, head-padded)", "ms": ms, "flops": fl,
                "exp_pairs_per_clk_per_sm": B2 * 8.0 * n * n / (ms * 1e-3) / 148 / 1.965e9})
    del q, k, v
    # K1 conv3x3 160->160 @ size^2
    x = torch.randn(B2, size, size, 160, device=dev).to(torch.bfloat16)
    w = (torch.randn(160, 3, 3, 160, device=dev) * 0.03).to(torch.bfloat16)
    b = torch.zeros(160, device=dev)
    ms = time_kernel(lambda: ops.conv3x3(x, w, b))
    out.append({"kernel": "umma_gemm_kernel conv3x3 160->160", "ms": ms, "flops": 2.0 * B2 * n * 9 * 160 * 160})
    # K1 conv3x3 320->320 @ (size/2)^2
    x2 = torch.randn(B2, size // 2, size // 2, 320, device=dev).to(torch.bfloat16)
    w2 = (torch.randn(320, 3, 3, 320, device=dev) * 0.02).to(torch.bfloat16)
    b2 = torch.zeros(320, device=dev)
    ms = time_kernel(lambda: ops.conv3x3(x2, w2, b2))
    out.append({"kernel": "umma_gemm_kernel conv3x3 320->320", "ms": ms, "flops": 2.0 * B2 * (n // 4) * 9 * 320 * 320})
    # K3 GEGLU GEMM [B2*n,160]x[160,1280]
    a = torch.randn(B2 * n, 160, device=dev).to(torch.bfloat16)
    wg = (torch.randn(1280, 160, device=dev) * 0.05).to(torch.bfloat16)
    bg = torch.zeros(1280, device=dev)
    ms = time_kernel(lambda: ops.gemm(a, wg, bg, act=2))
    out.append({"kernel": "umma_gemm_kernel GEGLU 160->1280", "ms": ms, "flops": 2.0 * B2 * n * 160 * 1280,
                "bytes": B2 * n * (160 + 640) * 2.0})
    for o in out:
        o["tflops"] = o["flops"] / (o["ms"] * 1e-3) / 1e12
        o["frac_of_bf16_peak"] = o["tflops"] / pk["bf16_tflops"]
        if "bytes" in o:
            o["gbs"] = o["bytes"] / (o["ms"] * 1e-3) / 1e9
            o["frac_of_hbm_peak"] = o["gbs"] / pk["hbm_gbs"]
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from stable_diffusion_pytorch_b200 import ops
    from stable_diffusion_pytorch_b200.graph_sampler import GraphedLatentStep
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler
    pk = peaks()
    B, S = args.batch, args.size
    unet = build_unet(dev)
    sch = DDPMScheduler(DDPMConfig())
    gs = GraphedLatentStep(unet, sch, B, S, S, time_steps=50, guidance_scale=7.5, sampler=args.sampler,
                           dtype=torch.bfloat16, device=dev, seed=1234 + rank)
    g = torch.Generator().manual_seed(100 + rank)   # each rank owns its shard of the latent batch
    x_host = torch.rand(B, 4, S, S, generator=g).to(torch.bfloat16).pin_memory()
    ctx_host = torch.randn(B, 77, 768, generator=g).to(torch.bfloat16).pin_memory()
    unc = torch.randn(1, 77, 768, generator=torch.Generator().manual_seed(7)).to(torch.bfloat16)
    out_host = torch.empty_like(x_host).pin_memory()
    gs.reset(x_host, ctx_host, unc)
    if not args.no_graph:
        gs.capture()
    gs.reset(x_host, ctx_host, unc)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing --------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        gs.step()
    barrier()
    if args.profile_step:
        torch.cuda.profiler.start()
        gs.step()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    clk = ClockSampler(local)
    if rank == 0:
        clk.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches0 = ops.launch_count()
    e0.record()
    for _ in range(args.steps):
        gs.step()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    clocks = clk.stop() if rank == 0 else None
    eager_launches = ops.launch_count() - launches0
    # ---- end-to-end timing: host buffers in, host result out, every step ------------------------
    gs.reset(x_host, ctx_host, unc)
    for _ in range(3):
        gs.step_host(x_host, ctx_host, out_host)
    gs.join_host()
    barrier()
    e0.record()
    for _ in range(args.steps):
        gs.step_host(x_host, ctx_host, out_host)
    gs.join_host()        # the last result's D2H (copy stream) is inside the timed region
    e1.record()
    barrier()
    ms_e2e = torch.tensor([e0.elapsed_time(e1)], device=dev)
    # ---- images/s: 50 latent-steps + one AutoencoderKL-shape decode (config 3) of this rank's latents ----
    ms_dec = torch.tensor([float("nan")], device=dev)
    ms_gather = torch.tensor([0.0], device=dev)
    if not args.no_decode and S == 64:
        ms_d, u8 = time_vae_decode(gs.x, dev)
        ms_dec = torch.tensor([ms_d], device=dev)
        if world > 1:
            ms_gather = torch.tensor([time_image_gather(u8, world)], device=dev)
        del u8
    finite = bool(torch.isfinite(gs.x.float()).all())
    # ---- config 4: UNet train step (loss + backward + gradient all-reduce + clip + AdamW), batch 32 per GPU ----
    ms_train = torch.tensor([float("nan")], device=dev)
    train_info = None
    graph_on, graph_lps = gs.graph is not None, gs.launches_per_step
    parity = eager = None
    if rank == 0 and (B, S) == (16, 64) and not args.no_parity:
        parity = sample_parity(gs, unet, x_host, ctx_host, unc)
        eager = gpu_eager_baseline(unet, B, S)
    if not args.no_train and S == 64:
        del gs
        torch.cuda.empty_cache()
        # config 4 trains the reference's default model: dropout 0.1 (models/unet.py:29-57), active in train mode
        train_unet = build_unet(dev, dropout=0.1)
        ms_t, train_info = time_train_step(train_unet, dev, rank, args.train_batch, S)
        del train_unet
        ms_train = torch.tensor([ms_t], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(ms_e2e, op=dist.ReduceOp.MAX)
        dist.all_reduce(ms_dec, op=dist.ReduceOp.MAX)
        dist.all_reduce(ms_gather, op=dist.ReduceOp.MAX)
        dist.all_reduce(ms_train, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_dec, ms_train, ms_gather = float(ms), float(ms_e2e), float(ms_dec), float(ms_train), float(ms_gather)
    flag = ops.debug_timeout_flag()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    n_lat = B * world
    value = n_lat * args.steps / (ms * 1e-3)
    e2e = n_lat * args.steps / (ms_e2e * 1e-3)
    kr = kernel_rooflines(2 * B, S, pk)
    dom = kr[0]
    kr = kr + hbm_kernels(2 * B, S, pk)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("attention_umma_dram_bytes_per_launch")
        except Exception:
            traffic = None
    step_tf = GF_PER_LATENT_STEP * 1e9 * B * (S * S / 4096.0) / (ms / args.steps * 1e-3) / 1e12
    line = {
        "metric": "latent_steps_per_sec", "value": value, "unit": "latent-steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD if (B, S) == (16, 64) else f"unet-default bs{B} {S}x{S}x4, CFG, {args.sampler}",
                   "batch_per_gpu": B, "latent": [4, S, S], "sampler": args.sampler, "guidance_scale": 7.5,
                   "parallelism": f"latent-batch sharded x{world}, no collective in the step",
                   "cuda_graph": not args.no_graph,
                   "cfg_prefix_shared": os.environ.get("SDB_CFG_SHARE", "1") != "0",
                   "cfg_note": "the two CFG copies of a latent differ only in their context: the layers ahead of the first "
                               "cross-attention (conv_in, ResBlock 1, GroupNorm / proj_in / self-attention of transformer 1) are "
                               "evaluated once per latent and copied -- same outputs; SDB_CFG_SHARE=0 evaluates both copies",
                   "l2": "no flush: per-step activation working set (~3 GB) is far larger than the 126 MB L2"},
        "e2e": {"value": e2e, "unit": "latent-steps/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": x_host.numel() * 2 + ctx_host.numel() * 2,
                "d2h_bytes_per_step": out_host.numel() * 2},
        "gpu_launches": (graph_lps * args.steps) if graph_on else eager_launches,
        "launches_per_step": graph_lps if graph_on else eager_launches // max(args.steps, 1),
        "clocks": clocks,
        "roofline": {"bound": "tensor", "kernel": dom["kernel"], "achieved": dom["tflops"], "peak": pk["bf16_tflops"],
                     "unit": "TFLOP/s", "frac": dom["tflops"] / pk["bf16_tflops"], "traffic": traffic,
                     "peak_source": pk["source"] + " (burst: kernel timed alone)", "ms_per_launch": dom["ms"],
                     "mufu_frac": dom["exp_pairs_per_clk_per_sm"] / MUFU_EX2_PER_CLK_PER_SM,
                     "exp_pairs_per_clk_per_sm": dom["exp_pairs_per_clk_per_sm"],
                     "note": "d_head=20 attention: 80 flops and ONE exp per (query, key) pair, so the MUFU ex2 pipe "
                             f"({MUFU_EX2_PER_CLK_PER_SM}/clk/SM measured, profiles/r01_ubench_mufu_fma.txt) binds long "
                             "before the tensor pipe; `frac` is the tensor-roofline fraction the contract asks for, "
                             "`mufu_frac` = (query, key) pairs/clk/SM over the MUFU-only rate (> 1 is possible: part "
                             "of the exponentials runs as a polynomial on the FMA pipe)"},
        "kernels": kr,
        "step_roofline": {"algorithmic_tflops": step_tf, "peak_sustained": pk["bf16_tflops_sustained"],
                          "frac": step_tf / pk["bf16_tflops_sustained"]},
        "images": None if ms_dec != ms_dec else {
            "images_per_sec": n_lat / ((50 * ms / args.steps + ms_dec + ms_gather) * 1e-3), "ms_decode": ms_dec,
            "ms_gather": ms_gather,
            "decoder": "AutoencoderKL-shape [128,256,512,512], 64x64x4 -> 512x512x3 uint8 (decode + detransform), bf16",
            "decode_alg_tflops": 38306.0 * (B / 16.0) / ms_dec,
            "gather": "none (1 GPU)" if world == 1 else
                      f"NCCL all-gather of the [{n_lat},512,512,3] uint8 images, the job's one collective",
            "formula": "latents / (50 x ms_per_step + ms_decode + ms_gather), max over ranks"},
        "train": None if ms_train != ms_train else {
            "samples_per_sec": args.train_batch * world / (ms_train * 1e-3), "ms_per_step": ms_train,
            "workload": f"BASELINE config 4: train_unet.py step, batch {args.train_batch}/GPU (CFG doubling -> UNet "
                        f"batch {2 * args.train_batch}), {S}x{S} latents, random CLIP-shaped context, dropout 0.1 (the reference "
                        f"default, 36 mask sites forward and backward), MSE, clip 0.1, AdamW; fp32 master weights / "
                        f"gradients, bf16 compute",
            "gradient_exchange": "one NCCL all-reduce of the flat fp32 gradient bucket" if world > 1 else "none (1 GPU)",
            **(train_info or {})},
        "parity": parity, "gpu_eager_baseline": eager,
        "finite": finite, "timeout_flag": flag,
    }
    if eager:
        line["gpu_eager_baseline"]["speedup_of_this_repo"] = value / world / eager["value"]
    if world == 1 and not args.no_cpu_baseline:
        try:
            cb = cpu_latent_steps(S, 3, 1)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as ex:  # pragma: no cover
            line["cpu_baseline"] = {"error": repr(ex)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
