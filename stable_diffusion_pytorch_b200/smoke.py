"""One tiny pass of the hot path on a CUDA device, checked against the CPU oracle
(used by ``__graft_entry__.smoke``; the oracle is only the checker here)."""
import torch


def run(device) -> None:
    from oracle import sd_oracle as O
    from . import ops
    from .models import DDPMConfig, DDPMScheduler, UNetModel, UnetConfig

    cfg = O.UnetCfg(groups=32, n_heads=4, channels_list=(64, 128), time_emb_dim=128, n_layers=1, context_dim=64)
    m = UNetModel(cfg.latent_channels, cfg.groups, UnetConfig(
        n_heads=4, channels_list=[64, 128], time_emb_dim=128, dropout=0.0, n_layers=1, context_dim=64)).eval()
    sd = O.seeded_state_dict(O.unet_param_shapes(cfg), 11)
    m.load_state_dict(sd, strict=True)
    m = m.to(device)
    g = torch.Generator().manual_seed(0)
    B = 2
    x = torch.rand(B, 4, 16, 16, generator=g)
    ctx = torch.randn(2 * B, 77, 64, generator=g)
    t = torch.full((2 * B,), 3, dtype=torch.long)
    noise = torch.randn(B, 4, 16, 16, generator=g)
    sch = DDPMScheduler(DDPMConfig())
    with torch.no_grad():
        eps = m(x.to(device), t.to(device), ctx.to(device), x_repeat=2)             # CFG pair of UNet evals
        xp, _ = ops.cfg_sampler_step(eps, x.to(device), noise.to(device), sch._device_tables(device)["coef"],
                                     slot=3, kind=0, cfg_mode=1, guidance=7.5)
        ref_eps = O.unet_forward(sd, cfg, torch.cat([x, x]), t, ctx)
        ref_xp, _ = O.ddpm_step(O.ddpm_tables(O.DdpmCfg()), O.cfg_combine(ref_eps, 7.5), x, 3, noise)
    torch.cuda.synchronize(device)
    e1 = float((eps.cpu().double() - ref_eps.double()).norm() / ref_eps.double().norm())
    e2 = float((xp.cpu().double() - ref_xp.double()).norm() / ref_xp.double().norm())
    assert ops.debug_timeout_flag() == 0
    assert e1 < 2e-2 and e2 < 5e-2, (e1, e2)
    print(f"smoke ok: unet rel-L2 {e1:.2e}, sampler step rel-L2 {e2:.2e}, launches {ops.launch_count()}")
