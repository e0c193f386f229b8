"""Golden vectors of the CLIP text encoder: runs the real ``transformers.CLIPTextModel`` (the third-party module
the reference wraps, models/clip_model.py:62-64; transformers 5.5.0 in the build container) on seeded weights.

    python tests/golden/make_golden_clip.py      ->  tests/golden/reference_clip.npz

No pretrained checkpoint is involved (there is no network): the weights are ``oracle.seeded_state_dict`` over the
model's own key/shape enumeration, loaded with strict=True, which also pins the oracle's key names to HF's.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from transformers import CLIPTextConfig, CLIPTextModel  # noqa: E402

from oracle import sd_oracle as O  # noqa: E402
from tests.golden import cases  # noqa: E402


def main():
    out = {}
    torch.set_num_threads(8)
    for name, (cfg, seed, B) in cases.CLIP_CASES.items():
        hf = CLIPTextModel(CLIPTextConfig(
            vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
            num_hidden_layers=cfg.num_hidden_layers, num_attention_heads=cfg.num_attention_heads,
            max_position_embeddings=cfg.max_position_embeddings, layer_norm_eps=cfg.layer_norm_eps,
            hidden_act="quick_gelu")).eval()
        sd = O.seeded_state_dict(O.clip_text_param_shapes(cfg), seed)
        missing = hf.load_state_dict(sd, strict=False)
        assert not missing.unexpected_keys, missing
        assert all("position_ids" in k for k in missing.missing_keys), missing   # buffer, not a parameter
        ids = cases.clip_inputs(cfg, B, seed)
        with torch.no_grad():
            y = hf(ids)[0]              # what latent_diffusion.py:55-60 / train_unet.py:506-508 consume
        out[f"clip_{name}"] = y.numpy().astype(np.float32)
        print(name, tuple(y.shape), float(y.abs().mean()))
    np.savez_compressed(os.path.join(HERE, "reference_clip.npz"), **out)


if __name__ == "__main__":
    main()
