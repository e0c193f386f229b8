"""Generate golden vectors by running the UNMODIFIED reference on seeded inputs.

Runs only in the build container (needs /root/reference); the GPU box has no
reference, so the outputs are committed as ``tests/golden/*.npz`` and this script
is the record of how they were made:

    python tests/golden/make_golden.py

Weights come from ``oracle.sd_oracle.seeded_state_dict`` (every tensor non-zero,
so zero_module'd layers -- models/utils.py:22-28 -- cannot hide errors) and are
loaded with ``load_state_dict(strict=True)``, which also proves that the oracle's
key/shape enumeration equals the reference's module tree.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

# import order matters: models before modules (circular import in the reference)
from stable_diffusion.models.unet import UNetModel, UnetConfig  # noqa: E402
from stable_diffusion.models.autoencoder import AutoEncoderKL, AutoencoderConfig  # noqa: E402
from stable_diffusion.models.scheduler import DDPMScheduler, DDPMConfig  # noqa: E402
from stable_diffusion.models.latent_diffusion import LatentDiffusion  # noqa: E402
from stable_diffusion.modules.timestep_embedding import sinusoidal_time_proj  # noqa: E402

from oracle import sd_oracle as O  # noqa: E402
from tests.golden import cases  # noqa: E402


class StubTextEncoder:
    """Duck type of models/clip_model.py:72-96 returning a fixed 'empty prompt' embedding."""

    def __init__(self, uncond):
        self.uncond = uncond  # [1, L, D]

    def tokenize(self, prompts):
        class _T:
            pass
        t = _T()
        t.input_ids = torch.zeros(len(prompts), self.uncond.shape[1], dtype=torch.long)
        return t

    def encode_text(self, ids):
        return (self.uncond.expand(ids.shape[0], -1, -1).clone(),)


def ref_unet(ucfg: O.UnetCfg, seed):
    cfg = UnetConfig(num_res_blocks=ucfg.num_res_blocks, n_heads=ucfg.n_heads,
                     attention_resolutions=list(ucfg.attention_resolutions),
                     channels_list=list(ucfg.channels_list), time_emb_dim=ucfg.time_emb_dim,
                     dropout=0.0, n_layers=ucfg.n_layers, context_dim=ucfg.context_dim)
    m = UNetModel(ucfg.latent_channels, ucfg.groups, cfg).eval()
    sd = O.seeded_state_dict(O.unet_param_shapes(ucfg), seed)
    m.load_state_dict(sd, strict=True)
    return m


def ref_vae(vcfg: O.VaeCfg, seed):
    cfg = AutoencoderConfig(in_channels=vcfg.in_channels, latent_channels=vcfg.latent_channels,
                            out_channels=vcfg.out_channels,
                            autoencoder_channels_list=list(vcfg.autoencoder_channels_list),
                            autoencoder_num_res_blocks=vcfg.autoencoder_num_res_blocks,
                            groups=vcfg.groups)
    m = AutoEncoderKL(cfg).eval()
    sd = O.seeded_state_dict(O.vae_decoder_param_shapes(vcfg), seed)
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all(k.startswith(("encoder.", "quant_conv.")) for k in missing), missing
    return m


def main():
    torch.set_num_threads(8)
    out = {}
    with torch.no_grad():
        # ---- timestep embedding -------------------------------------------------
        for dim in (64, 160):
            t = torch.tensor(cases.TIMESTEP_CASES, dtype=torch.long)
            out[f"temb_{dim}"] = sinusoidal_time_proj(t, dim).numpy()

        # ---- scheduler tables / step / add_noise ---------------------------------
        sch = DDPMScheduler(DDPMConfig())
        for name in O.TABLE_NAMES:
            out[f"table_{name}"] = getattr(sch, name).numpy()
        out["table_noise_time_steps"] = sch.noise_time_steps.numpy()
        for step in cases.STEP_CASES:
            eps, x_t = cases.step_inputs(step)
            torch.manual_seed(cases.STEP_NOISE_SEED + step)  # reference draws from the global CPU generator
            x_prev, x0 = sch.step(eps, x_t, step, scale_factor=cases.STEP_SCALE)
            out[f"step_{step}_xprev"] = x_prev.numpy()
            out[f"step_{step}_x0"] = x0.numpy()
        x0s, nz, ts = cases.add_noise_inputs()
        out["add_noise"] = sch.add_noise(x0s, nz, ts).numpy()

        # ---- UNet forwards -------------------------------------------------------
        for name, (ucfg, seed, B, H, W) in cases.UNET_CASES.items():
            m = ref_unet(ucfg, seed)
            x, t, ctx = cases.unet_inputs(ucfg, B, H, W, seed)
            out[f"unet_{name}"] = m(x, t, ctx).numpy()
            print("unet", name, out[f"unet_{name}"].shape, float(np.abs(out[f"unet_{name}"]).mean()))

        # ---- CFG + sampling loop (reference LatentDiffusion with a stub text encoder) --
        for name, (ucfg, seed, B, H, W, steps, g) in cases.SAMPLE_CASES.items():
            m = ref_unet(ucfg, seed)
            x, ctx, uncond = cases.sample_inputs(ucfg, B, H, W, seed)
            ld = LatentDiffusion(m, None, StubTextEncoder(uncond), sch)
            t = torch.full((B,), 7, dtype=torch.long)
            out[f"prednoise_{name}"] = ld.pred_noise(x, t, ctx, guidance_scale=g).numpy()
            torch.manual_seed(cases.SAMPLE_NOISE_SEED)
            out[f"sample_{name}"] = ld.sample(x, ctx, guidance_scale=g, time_steps=steps).numpy()
            print("sample", name, float(np.abs(out[f"sample_{name}"]).mean()))

        # ---- VAE decode ----------------------------------------------------------
        for name, (vcfg, seed, B, H, W) in cases.VAE_CASES.items():
            m = ref_vae(vcfg, seed)
            z = cases.vae_inputs(vcfg, B, H, W, seed)
            out[f"vae_{name}"] = m.decode(z).numpy()
            print("vae", name, out[f"vae_{name}"].shape, float(np.abs(out[f"vae_{name}"]).mean()))

    path = os.path.join(HERE, "reference_outputs.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


def main_encode():
    """AutoEncoderKL.encode (models/autoencoder.py:132-146) and GaussianDistribution.sample / .kl
    (modules/distributions.py:5-19) of the unmodified reference -> ``reference_vae_encode.npz``."""
    torch.set_num_threads(8)
    out = {}
    with torch.no_grad():
        for name, (vcfg, seed, B, H, W) in cases.VAE_ENC_CASES.items():
            cfg = AutoencoderConfig(in_channels=vcfg.in_channels, latent_channels=vcfg.latent_channels,
                                    out_channels=vcfg.out_channels,
                                    autoencoder_channels_list=list(vcfg.autoencoder_channels_list),
                                    autoencoder_num_res_blocks=vcfg.autoencoder_num_res_blocks, groups=vcfg.groups)
            m = AutoEncoderKL(cfg).eval()
            sd = O.seeded_state_dict(O.vae_encoder_param_shapes(vcfg), seed)
            missing, unexpected = m.load_state_dict(sd, strict=False)
            assert not unexpected, unexpected          # the oracle's key / shape enumeration is the reference's
            assert all(k.startswith(("decoder.", "post_quant_conv.")) for k in missing), missing
            img = cases.vae_encode_inputs(vcfg, B, H, W, seed)
            dist = m.encode(img).latent_dist
            out[f"enc_{name}_mean"] = dist.mean.numpy().copy()
            out[f"enc_{name}_logvar"] = dist.log_var.numpy().copy()
            torch.manual_seed(cases.VAE_ENC_EPS_SEED)       # sample() draws randn_like(std) from the global generator
            out[f"enc_{name}_sample"] = dist.sample().numpy().copy()
            # the draw itself, for the tests: randn_like follows the memory format of the moments (torch's CPU conv
            # hands back channels-last strides here), so it is recorded rather than re-derived from a shape
            torch.manual_seed(cases.VAE_ENC_EPS_SEED)
            eps = torch.randn_like(torch.exp(0.5 * dist.log_var))
            assert torch.equal(dist.mean + eps * torch.exp(0.5 * dist.log_var), torch.from_numpy(out[f"enc_{name}_sample"]))
            out[f"enc_{name}_eps"] = eps.contiguous().numpy().copy()
            out[f"enc_{name}_kl"] = dist.kl().numpy().copy()
            print("encode", name, out[f"enc_{name}_mean"].shape, float(np.abs(out[f"enc_{name}_mean"]).mean()),
                  out[f"enc_{name}_kl"])
    path = os.path.join(HERE, "reference_vae_encode.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


def main_train():
    """Row T1: the reference's own modules, F.mse_loss, clip_grad_norm_ and torch.optim.AdamW
    (train_unet.py:489-529, :359-375, :231-236) on seeded batches -> ``reference_train.npz``."""
    import torch.nn.functional as F
    torch.set_num_threads(8)
    out = {}
    sch = DDPMScheduler(DDPMConfig())
    for name, (ucfg, seed, B, H, W, n_steps, g) in cases.TRAIN_CASES.items():
        m = ref_unet(ucfg, seed).train()
        batches, uncond = cases.train_batches(ucfg, B, H, W, seed, n_steps)
        ld = LatentDiffusion(m, None, StubTextEncoder(uncond), sch)
        opt = torch.optim.AdamW(m.parameters(), lr=cases.TRAIN_LR, weight_decay=cases.TRAIN_WD)
        names = [k for k, _ in m.named_parameters()]
        losses, norms = [], []
        for i, (x_t, t, ctx, noise) in enumerate(batches):
            pred = ld.pred_noise(x_t, t, ctx, guidance_scale=g)
            loss = F.mse_loss(pred.float(), noise.float(), reduction="mean")
            loss.backward()
            if i == 0:
                out[f"train_{name}_gradnorms"] = np.array([float(p.grad.norm()) for p in m.parameters()], np.float32)
                for k in cases.TRAIN_PROBE:
                    out[f"train_{name}_grad_{k}"] = dict(m.named_parameters())[k].grad.numpy().copy()
            norms.append(float(torch.nn.utils.clip_grad_norm_(m.parameters(), cases.TRAIN_CLIP)))
            opt.step()
            opt.zero_grad()
            losses.append(float(loss))
        out[f"train_{name}_loss"] = np.array(losses, np.float32)
        out[f"train_{name}_gnorm"] = np.array(norms, np.float32)
        sd0 = O.seeded_state_dict(O.unet_param_shapes(ucfg), seed)
        pd = dict(m.named_parameters())
        out[f"train_{name}_deltanorms"] = np.array([float((pd[k].detach() - sd0[k]).norm()) for k in names], np.float32)
        for k in cases.TRAIN_PROBE:
            out[f"train_{name}_param_{k}"] = pd[k].detach().numpy().copy()
        out[f"train_{name}_names"] = np.array(names)
        print("train", name, losses, norms)
    path = os.path.join(HERE, "reference_train.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "train":
        main_train()
    elif len(sys.argv) > 1 and sys.argv[1] == "encode":
        main_encode()
    else:
        main()
