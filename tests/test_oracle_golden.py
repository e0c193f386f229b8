"""CPU: the oracle restatement vs golden vectors produced by the unmodified reference
(tests/golden/make_golden.py).  Pins the oracle."""
import numpy as np
import pytest
import torch

from oracle import sd_oracle as O
from tests.golden import cases


def rel_l2(a, b):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def test_tables_bit_exact(golden):
    tab = O.ddpm_tables(O.DdpmCfg())
    for name in O.TABLE_NAMES + ("noise_time_steps",):
        assert np.array_equal(tab[name].numpy(), golden[f"table_{name}"]), name


def test_sample_timesteps_ascending():
    tab = O.ddpm_tables(O.DdpmCfg())
    assert O.sample_timesteps(tab, 50) == list(range(50))
    assert O.sample_timesteps(tab, None) == list(range(1000))


@pytest.mark.parametrize("dim", [64, 160])
def test_timestep_embedding_bit_exact(golden, dim):
    t = torch.tensor(cases.TIMESTEP_CASES, dtype=torch.long)
    assert np.array_equal(O.sinusoidal_time_proj(t, dim).numpy(), golden[f"temb_{dim}"])


@pytest.mark.parametrize("step", cases.STEP_CASES)
def test_ddpm_step_bit_exact(golden, step):
    tab = O.ddpm_tables(O.DdpmCfg())
    eps, x_t = cases.step_inputs(step)
    torch.manual_seed(cases.STEP_NOISE_SEED + step)
    noise = None if step == 0 else torch.randn(x_t.shape)
    x_prev, x0 = O.ddpm_step(tab, eps, x_t, step, noise, cases.STEP_SCALE)
    assert np.array_equal(x_prev.numpy(), golden[f"step_{step}_xprev"])
    assert np.array_equal(x0.numpy(), golden[f"step_{step}_x0"])


def test_add_noise_bit_exact(golden):
    tab = O.ddpm_tables(O.DdpmCfg())
    x0, nz, ts = cases.add_noise_inputs()
    assert np.array_equal(O.add_noise(tab, x0, nz, ts).numpy(), golden["add_noise"])


@pytest.mark.parametrize("name", list(cases.UNET_CASES))
def test_unet_forward(golden, name):
    cfg, seed, B, H, W = cases.UNET_CASES[name]
    sd = O.seeded_state_dict(O.unet_param_shapes(cfg), seed)
    x, t, ctx = cases.unet_inputs(cfg, B, H, W, seed)
    with torch.no_grad():
        y = O.unet_forward(sd, cfg, x, t, ctx)
    assert rel_l2(y, golden[f"unet_{name}"]) < 1e-5


@pytest.mark.parametrize("name", list(cases.SAMPLE_CASES))
def test_pred_noise_and_sample(golden, name):
    cfg, seed, B, H, W, steps, g = cases.SAMPLE_CASES[name]
    sd = O.seeded_state_dict(O.unet_param_shapes(cfg), seed)
    x, ctx, uncond = cases.sample_inputs(cfg, B, H, W, seed)
    tab = O.ddpm_tables(O.DdpmCfg())
    fn = lambda a, b, c: O.unet_forward(sd, cfg, a, b, c)
    un = uncond.expand(B, -1, -1)
    with torch.no_grad():
        t = torch.full((B,), 7, dtype=torch.long)
        pn = O.pred_noise(fn, x, t, ctx, un, g)
        assert rel_l2(pn, golden[f"prednoise_{name}"]) < 1e-5
        torch.manual_seed(cases.SAMPLE_NOISE_SEED)
        xs = O.sample(fn, tab, x, ctx, un, g, time_steps=steps)
    assert rel_l2(xs, golden[f"sample_{name}"]) < 1e-4


@pytest.mark.parametrize("name", list(cases.VAE_CASES))
def test_vae_decode(golden, name):
    cfg, seed, B, H, W = cases.VAE_CASES[name]
    sd = O.seeded_state_dict(O.vae_decoder_param_shapes(cfg), seed)
    z = cases.vae_inputs(cfg, B, H, W, seed)
    with torch.no_grad():
        y = O.vae_decode(sd, cfg, z)
    assert rel_l2(y, golden[f"vae_{name}"]) < 1e-5


@pytest.mark.parametrize("name", list(cases.VAE_ENC_CASES))
def test_vae_encode_and_gaussian_distribution(golden_encode, name):
    """encoder + quant_conv moments, GaussianDistribution.sample (same global-generator draw) and .kl
    against the unmodified reference (make_golden.py encode)."""
    cfg, seed, B, H, W = cases.VAE_ENC_CASES[name]
    sd = O.seeded_state_dict(O.vae_encoder_param_shapes(cfg), seed)
    img = cases.vae_encode_inputs(cfg, B, H, W, seed)
    with torch.no_grad():
        mom = O.vae_encode(sd, cfg, img)
        mean, log_var = O.gaussian_moments(mom)
        smp = O.gaussian_sample(mom, torch.from_numpy(golden_encode[f"enc_{name}_eps"]))
        kl = O.gaussian_kl(mom)
    G = golden_encode
    assert tuple(mean.shape) == G[f"enc_{name}_mean"].shape
    assert rel_l2(mean, G[f"enc_{name}_mean"]) < 1e-5
    assert rel_l2(log_var, G[f"enc_{name}_logvar"]) < 1e-5
    assert rel_l2(smp, G[f"enc_{name}_sample"]) < 1e-5
    np.testing.assert_allclose(kl.numpy(), G[f"enc_{name}_kl"], rtol=1e-4)


def test_cfg_combine_is_mislabelled_like_reference():
    # latent_diffusion.py:61-65: first chunk is the UNCOND half
    e_unc, e_cond = torch.full((1, 1), 1.0), torch.full((1, 1), 3.0)
    out = O.cfg_combine(torch.cat([e_unc, e_cond]), 2.0)
    assert out.item() == 1.0 + 2.0 * (1.0 - 3.0)


@pytest.mark.parametrize("name", list(cases.TRAIN_CASES))
def test_train_steps_vs_reference(golden_train, name):
    """Row T1: oracle loss / autograd gradients / clip / AdamW restatement vs the reference's modules run with
    F.mse_loss, clip_grad_norm_ and torch.optim.AdamW (make_golden.py train)."""
    cfg, seed, B, H, W, n_steps, g = cases.TRAIN_CASES[name]
    sd = O.seeded_state_dict(O.unet_param_shapes(cfg), seed)
    batches, uncond = cases.train_batches(cfg, B, H, W, seed, n_steps)
    torch.set_num_threads(8)
    losses, norms, grads, params = O.train_steps(sd, cfg, batches, uncond, g, cases.TRAIN_CLIP, cases.TRAIN_LR,
                                                 cases.TRAIN_WD)
    G = golden_train
    np.testing.assert_allclose(losses, G[f"train_{name}_loss"], rtol=2e-5)
    np.testing.assert_allclose(norms, G[f"train_{name}_gnorm"], rtol=2e-4)
    names = [str(k) for k in G[f"train_{name}_names"]]
    assert sorted(names) == sorted(sd)                         # same parameter set as the reference's module tree
    gn = np.array([float(grads[0][k].norm()) for k in names])
    np.testing.assert_allclose(gn, G[f"train_{name}_gradnorms"], rtol=2e-3, atol=1e-6)
    dn = np.array([float((params[k] - sd[k]).norm()) for k in names])
    np.testing.assert_allclose(dn, G[f"train_{name}_deltanorms"], rtol=2e-3, atol=1e-7)
    for k in cases.TRAIN_PROBE:
        assert rel_l2(grads[0][k], G[f"train_{name}_grad_{k}"]) < 1e-4, k
        assert rel_l2(params[k], G[f"train_{name}_param_{k}"]) < 1e-6, k


@pytest.mark.parametrize("name", list(cases.CLIP_CASES))
def test_clip_text_oracle_vs_transformers_golden(name):
    """The CLIP text tower restatement against the real transformers.CLIPTextModel (tests/golden/make_golden_clip.py)."""
    import os
    gold = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_clip.npz")))
    cfg, seed, B = cases.CLIP_CASES[name]
    sd = O.seeded_state_dict(O.clip_text_param_shapes(cfg), seed)
    with torch.no_grad():
        y = O.clip_text_forward(sd, cfg, cases.clip_inputs(cfg, B, seed))
    ref = torch.from_numpy(gold[f"clip_{name}"])
    assert tuple(y.shape) == tuple(ref.shape)
    assert float((y - ref).norm() / ref.norm()) < 1e-5
