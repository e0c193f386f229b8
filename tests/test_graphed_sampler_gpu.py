"""GPU parity of the path bench.py times: ``GraphedLatentStep`` (CFG pair of UNet evaluations + fused guidance
combine + DDIM / DDPM update IN PLACE, step row selected by a device counter, replayed from a CUDA graph) against
the oracle's ``pred_noise`` + ``ddim_step`` / ``ddpm_step`` looped over the reference's ascending timesteps
(models/latent_diffusion.py:27-126, models/scheduler.py:173-213).

Tolerances: final latents of a bf16 run vs the fp32 oracle, rel-L2 <= 5e-2; graph replay vs
eager: bit for bit."""
import pytest
import torch

from oracle import sd_oracle as O
from tests.golden import cases

pytestmark = pytest.mark.gpu
DEV = "cuda"
SAMPLE_TOL = 5e-2


@pytest.fixture(scope="module", autouse=True)
def _gpu():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from stable_diffusion_pytorch_b200 import ops
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    assert ops.debug_timeout_flag() == 0


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm())


def build_unet(cfg, seed):
    from stable_diffusion_pytorch_b200.models import UNetModel, UnetConfig
    m = UNetModel(cfg.latent_channels, cfg.groups, UnetConfig(
        num_res_blocks=cfg.num_res_blocks, n_heads=cfg.n_heads, attention_resolutions=list(cfg.attention_resolutions),
        channels_list=list(cfg.channels_list), time_emb_dim=cfg.time_emb_dim, dropout=0.0, n_layers=cfg.n_layers,
        context_dim=cfg.context_dim)).eval()
    sd = O.seeded_state_dict(O.unet_param_shapes(cfg), seed)
    m.load_state_dict(sd, strict=True)
    return m.to(DEV), sd


def oracle_run(sd, cfg, x, ctx, uncond, steps, sampler, g, noises=None, device="cpu"):
    """O.sample on `device` (cpu = the oracle proper; cuda = the same restatement executed by torch eager in fp32)."""
    sdd = {k: v.to(device) for k, v in sd.items()}
    tab = O.ddpm_tables(O.DdpmCfg())
    fn = lambda a, b, c: O.unet_forward(sdd, cfg, a, b, c)  # noqa: E731
    B = x.shape[0]
    nz = None if noises is None else [n.to(device) for n in noises]
    with torch.no_grad():
        return O.sample(fn, tab, x.to(device), ctx.to(device), uncond.expand(B, -1, -1).to(device), g,
                        time_steps=steps, noises=nz, sampler=sampler).cpu()


def graphed(m, B, H, W, steps, sampler, g, host_noise=None, ctx_len=77):
    from stable_diffusion_pytorch_b200.graph_sampler import GraphedLatentStep
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler
    return GraphedLatentStep(m, DDPMScheduler(DDPMConfig()), B, H, W, ctx_len=ctx_len, time_steps=steps,
                             guidance_scale=g, sampler=sampler, dtype=torch.bfloat16, device=DEV,
                             host_noise=host_noise)


@pytest.mark.parametrize("sampler", ["ddim", "ddpm"])
def test_graphed_step_50_steps_vs_oracle_small(sampler):
    """TINY UNet, 2 x 4 x 16 x 16, CFG 7.5, the full 50 ascending steps: eager, then captured; replay == eager."""
    cfg, seed, B, H, W, steps, g = cases.TINY, 11, 2, 16, 16, 50, 7.5
    m, sd = build_unet(cfg, seed)
    x, ctx, uncond = cases.sample_inputs(cfg, B, H, W, seed)
    noises = None
    if sampler == "ddpm":   # the reference draws torch.randn(x.shape) on the host every step but step 0
        gen = torch.Generator().manual_seed(cases.SAMPLE_NOISE_SEED)
        noises = [torch.randn(x.shape, generator=gen) for _ in range(steps)]
    # bf16-rounded inputs for both sides (the graphed path holds x / ctx in bf16)
    xb, cb, ub = (t.to(torch.bfloat16).float() for t in (x, ctx, uncond))
    ref = oracle_run(sd, cfg, xb, cb, ub, steps, sampler, g, noises)
    gs = graphed(m, B, H, W, steps, sampler, g, None if noises is None else torch.stack(noises))
    eager = gs.run(xb, cb, ub)
    assert gs.graph is None
    gs.capture()
    assert gs.graph is not None and gs.launches_per_step > 0
    replay = gs.run(xb, cb, ub)
    assert torch.equal(eager, replay), "CUDA-graph replay differs from the eager run"
    assert int(gs.slot) == 0 and eager.dtype == torch.bfloat16      # the slot wraps after the last step: ready for the next run
    e = rel_l2(replay.float(), ref)
    print(f"graphed {sampler} tiny 50 steps: rel-L2 vs oracle = {e:.3e}")
    assert e < SAMPLE_TOL
    # a second run on the same captured graph starts from slot 0 again and reproduces the first
    assert torch.equal(gs.run(xb, cb, ub), replay)


def test_graphed_step_config2_geometry_vs_oracle():
    """BASELINE config 2 geometry: 16 x 4 x 64 x 64 latents, UNet batch 32 (grid (32, 8, 32) of the attention kernel,
    the 148-CTA persistent tile walk at M = 131 072), default UNet, DDIM + CFG 7.5, 3 steps from a captured graph.
    Every latent is held to the oracle restatement executed by torch eager in fp32 on the GPU; latents 0 and 15 also to
    the CPU oracle proper (trajectories are independent per latent, so a 2-latent CPU run is the same arithmetic)."""
    cfg, seed, B, H, W, steps, g = cases.DEFAULT, 13, 16, 64, 64, 3, 7.5
    m, sd = build_unet(cfg, seed)
    x, ctx, uncond = cases.sample_inputs(cfg, B, H, W, seed)
    xb, cb, ub = (t.to(torch.bfloat16).float() for t in (x, ctx, uncond))
    gs = graphed(m, B, H, W, 50, "ddim", g)
    gs.steps = gs.steps[:steps]          # only the loop length changes: tables / slots still cover 50 steps
    gs.capture()
    got = gs.run(xb, cb, ub).float().cpu()
    assert int(gs.slot) == steps
    # all 16 rows vs the oracle on the GPU (fp32 eager, TF32 off), 4 latents at a time to bound `sim` (4.3 GB each)
    ref = torch.cat([oracle_run(sd, cfg, xb[i:i + 4], cb[i:i + 4], ub, steps, "ddim", g, device=DEV)
                     for i in range(0, B, 4)])
    per_row = [rel_l2(got[i], ref[i]) for i in range(B)]
    print("graphed ddim config-2 geometry, 3 steps, per-latent rel-L2 vs fp32 oracle (cuda):",
          " ".join(f"{e:.2e}" for e in per_row))
    assert max(per_row) < SAMPLE_TOL
    idx = [0, 15]
    ref_cpu = oracle_run(sd, cfg, xb[idx], cb[idx], ub, steps, "ddim", g)
    e_cpu = rel_l2(got[idx], ref_cpu)
    e_oo = rel_l2(ref[idx], ref_cpu)
    print(f"  latents 0/15 vs CPU oracle: {e_cpu:.3e}; cuda-fp32 oracle vs CPU oracle: {e_oo:.3e}")
    assert e_cpu < SAMPLE_TOL and e_oo < 1e-4


def test_latent_diffusion_sample_ddim_and_graph_cache_vs_oracle():
    """The drop-in call: LatentDiffusion.sample(sampler='ddim') (and the reference's own 'ddpm') replays a cached
    GraphedLatentStep; same result as the eager loop bit for bit, and within tolerance of the oracle."""
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler, LatentDiffusion
    from tests.test_parity_gpu import StubText
    cfg, seed, B, H, W, steps, g = cases.TINY, 11, 2, 16, 16, 6, 7.5
    m, sd = build_unet(cfg, seed)
    x, ctx, uncond = cases.sample_inputs(cfg, B, H, W, seed)
    ld = LatentDiffusion(m, None, StubText(uncond), DDPMScheduler(DDPMConfig()))
    for sampler in ("ddim", "ddpm"):
        noises = None
        if sampler == "ddpm":
            gen = torch.Generator().manual_seed(cases.SAMPLE_NOISE_SEED)
            noises = [torch.randn(x.shape, generator=gen) for _ in range(steps)]
        ref = oracle_run(sd, cfg, x, ctx, uncond, steps, sampler, g, noises)
        outs = {}
        for use_graph in (False, True):
            torch.manual_seed(cases.SAMPLE_NOISE_SEED)     # sample() draws the host noise like the reference
            outs[use_graph] = ld.sample(x.to(DEV), ctx.to(DEV), guidance_scale=g, time_steps=steps, sampler=sampler,
                                        use_graph=use_graph)
        assert outs[True].dtype == torch.float32 and tuple(outs[True].shape) == tuple(x.shape)
        assert torch.equal(outs[True], outs[False]), f"{sampler}: graphed sample() differs from the eager loop"
        e = rel_l2(outs[True], ref)
        print(f"LatentDiffusion.sample {sampler}: rel-L2 vs oracle = {e:.3e}")
        assert e < SAMPLE_TOL
    assert len(ld._graphs) == 2      # one cached graph per (shape, sampler, guidance) key
    # second call with the same key: cache hit, same answer
    torch.manual_seed(cases.SAMPLE_NOISE_SEED)
    again = ld.sample(x.to(DEV), ctx.to(DEV), guidance_scale=g, time_steps=steps, sampler="ddpm")
    assert torch.equal(again, outs[True]) and len(ld._graphs) == 2


def test_step_host_pipelined_copies_match_device_resident_steps():
    """``step_host`` (the bench's end-to-end call: host buffers in, host result out, PCIe copies on a side stream through
    staging buffers) must return, for every call, exactly what a device-resident step on the same inputs returns --
    also when calls are enqueued back to back without a host wait in between."""
    cfg, seed, B, H, W = cases.TINY, 5, 2, 16, 16
    m, _ = build_unet(cfg, seed)
    gs = graphed(m, B, H, W, 50, "ddim", 7.5)
    _, ctx0, uncond = cases.sample_inputs(cfg, B, H, W, seed)
    gs.reset(torch.zeros(B, cfg.latent_channels, H, W), ctx0, uncond)
    gs.capture()
    g = torch.Generator().manual_seed(3)
    xs = [torch.randn(B, cfg.latent_channels, H, W, generator=g).to(torch.bfloat16).pin_memory() for _ in range(4)]
    cs = [torch.randn(B, 77, cfg.context_dim, generator=g).to(torch.bfloat16).pin_memory() for _ in range(4)]
    outs = [torch.empty_like(xs[0]).pin_memory() for _ in range(4)]
    want = []
    for x, c in zip(xs, cs):                       # device-resident reference: same slot (0) every time
        gs.reset(x, c)
        gs.step()
        want.append(gs.x.clone())
    for i in range(4):                             # back to back, no host synchronisation
        gs.slot.zero_(); gs.t.fill_(gs.steps[0])
        gs.step_host(xs[i], cs[i], outs[i])
    gs.sync_host()
    for i in range(4):
        assert torch.equal(outs[i], want[i].cpu()), f"call {i}"
