"""Training-mode dropout (resnet2d.py:190, transformer.py:58,147).  The masks are device-RNG ("not covered by any
oracle"), so the tests are the properties nn.Dropout guarantees: keep rate, 1/(1-p) scaling, identity in
eval mode, the backward using the forward's mask, fresh masks per step -- also under CUDA-graph replay."""
import pytest
import torch

from oracle import sd_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module", autouse=True)
def _gpu():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("p", [0.1, 0.5])
def test_dropout_kernel_statistics_and_mask_reuse(dtype, p):
    from stable_diffusion_pytorch_b200 import ops
    ops.dropout_seed(123, DEV)
    n = 1 << 20
    x = torch.ones(n + 3, device=DEV, dtype=dtype)          # n % 4 != 0: ragged tail quad
    y = ops.dropout_raw(x, p, site=5)
    kept = y != 0
    rate = float(kept.float().mean())
    assert abs(rate - (1 - p)) < 4e-3                          # 1M Bernoulli draws: sigma <= 5e-4
    vals = y[kept].float()
    assert float((vals - vals[0]).abs().max()) == 0 and abs(float(vals[0]) - 1 / (1 - p)) < 1e-2
    assert torch.equal(ops.dropout_raw(x, p, site=5), y)       # same (state, site): same mask = the backward's mask
    assert not torch.equal(ops.dropout_raw(x, p, site=6) != 0, kept)      # another layer: another mask
    ops.dropout_advance(DEV)
    y2 = ops.dropout_raw(x, p, site=5)
    assert not torch.equal(y2 != 0, kept)                      # next step: fresh mask
    both = float(((y2 != 0) & kept).float().mean())
    assert abs(both - (1 - p) ** 2) < 6e-3                     # and independent of the previous one


def test_dropout_autograd_uses_the_forward_mask():
    from stable_diffusion_pytorch_b200 import ops
    ops.dropout_seed(7, DEV)
    ops.dropout_begin_step(DEV)
    x = torch.randn(64, 160, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    y = ops.dropout(x, 0.25, training=True)
    g = torch.randn_like(y)
    y.backward(g)
    mask = (y != 0) | (x == 0)
    scale = 1 / (1 - 0.25)
    want = torch.where(mask, g.float() * scale, torch.zeros_like(g, dtype=torch.float32))
    assert float((x.grad.float() - want).abs().max()) < 2e-2 * float(want.abs().max())
    assert ops.dropout(x, 0.25, training=False) is x and ops.dropout(x, 0.0, training=True) is x


def _tiny_ld(dropout):
    from stable_diffusion_pytorch_b200.models import DDPMConfig, DDPMScheduler, LatentDiffusion, UNetModel, UnetConfig
    from tests.test_train_gpu import StubText
    cfg = O.UnetCfg(groups=32, n_heads=4, channels_list=(64, 128), time_emb_dim=128, n_layers=1, context_dim=64)
    m = UNetModel(4, 32, UnetConfig(n_heads=4, channels_list=[64, 128], time_emb_dim=128, n_layers=1, context_dim=64,
                                    dropout=dropout))
    m.load_state_dict(O.seeded_state_dict(O.unet_param_shapes(cfg), 3))
    unc = torch.randn(1, 77, 64, generator=torch.Generator().manual_seed(1))
    return LatentDiffusion(m.to(DEV), None, StubText(unc), DDPMScheduler(DDPMConfig()))


def _batch(seed=0, B=2):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(B, 4, 16, 16, generator=g).to(DEV), torch.randint(0, 1000, (B,), generator=g).to(DEV),
            torch.randn(B, 77, 64, generator=g).to(DEV), torch.randn(B, 4, 16, 16, generator=g).to(DEV))


def test_unet_with_reference_default_dropout_trains_and_eval_is_deterministic():
    """UnetConfig's default dropout = 0.1 (unet.py:29-57): train mode perturbs the output with fresh masks each step,
    eval mode is the dropout-free forward, and a train step runs end to end (loss finite, every gradient finite)."""
    from stable_diffusion_pytorch_b200 import ops
    from stable_diffusion_pytorch_b200.training import FlatAdamW, train_step
    ld, ld0 = _tiny_ld(0.1), _tiny_ld(0.0)
    x, t, ctx, noise = _batch()
    ops.dropout_seed(11, DEV)
    with torch.no_grad():
        ld.unet.eval(); ld0.unet.eval()
        e1, e0 = ld.unet(x, t, ctx), ld0.unet(x, t, ctx)
        assert torch.equal(e1, e0)                              # eval: dropout is the identity
        ld.unet.train()
        ops.dropout_begin_step(DEV)
        a = ld.unet(x, t, ctx)
        ops.dropout_begin_step(DEV)
        b = ld.unet(x, t, ctx)                                  # same state, same sites: same masks
        ops.dropout_advance(DEV)
        ops.dropout_begin_step(DEV)
        c = ld.unet(x, t, ctx)
    rel = lambda u, v: float((u - v).norm() / v.norm())       # noqa: E731
    assert torch.equal(a, b) and rel(a, e0) > 1e-2 and rel(c, a) > 1e-2 and rel(a, e0) < 1.0
    opt = FlatAdamW(ld.unet.parameters())
    before = opt.flat.clone()
    losses = [float(train_step(ld, opt, x, t, ctx, noise)) for _ in range(3)]
    assert all(l == l and l < 1e3 for l in losses) and len(set(losses)) == 3      # finite, and the masks change
    assert bool(torch.isfinite(opt.flat).all()) and float((opt.flat - before).abs().max()) > 0


def test_graphed_train_step_draws_fresh_dropout_masks_per_replay():
    from stable_diffusion_pytorch_b200 import ops
    from stable_diffusion_pytorch_b200.training import FlatAdamW, GraphedTrainStep, OptimConfig
    ld = _tiny_ld(0.1)
    ld.unet.train()
    opt = FlatAdamW(ld.unet.parameters(), OptimConfig(learning_rate=0.0, adam_weight_decay=0.0))   # weights frozen
    x, t, ctx, noise = _batch()
    ops.dropout_seed(5, DEV)
    gts = GraphedTrainStep(ld, opt, 2, (16, 16), ctx_len=77, ctx_dim=64)
    gts.capture()
    losses = [float(gts.step(x, t, ctx, noise)) for _ in range(4)]
    assert len(set(losses)) == 4, losses            # same batch, same weights: only the dropout masks differ
    assert max(losses) - min(losses) < 0.5 * max(losses)
