"""CLIP text encoder behind the reference's ``CLIPModel`` duck type (stable_diffusion/models/clip_model.py:56-96:
``tokenize(prompts).input_ids`` and ``encode_text(ids)[0]``), SURVEY 8f-3.

The reference wraps the third-party ``transformers.CLIPTextModel``; here the text tower itself runs on libsdb200:
token + position embedding gather, twelve pre-LayerNorm blocks (fused-QKV tcgen05 GEMM, causal attention over the 77
tokens, out-projection GEMM with the residual in its epilogue, quick-GELU MLP on the GEMM kernel's SiLU epilogue),
final LayerNorm.  The parameter tree carries the ``transformers`` names
(``text_model.encoder.layers.3.self_attn.q_proj.weight`` ...), so ``load_state_dict(CLIPTextModel.state_dict())`` --
the ``text_encoder`` sub-folder of runwayml/stable-diffusion-v1-5 -- works unchanged.  The tokenizer (BPE vocabulary
files, pure host string work) is not part of the path: pass any callable with the reference's ``tokenize`` contract.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Optional

import torch
from torch import nn

from .. import _lib, _pack, ops
from .._lib import ACT_SILU, check

Tensor = torch.Tensor
QUICK_GELU = 1.702      # x * sigmoid(1.702 x) == silu(1.702 x) / 1.702


@dataclass
class ClipTextConfig:
    """Shape of the text tower; defaults = runwayml/stable-diffusion-v1-5 ``text_encoder/config.json``."""
    vocab_size: int = 49408
    hidden_size: int = 768
    intermediate_size: int = 3072
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    max_position_embeddings: int = 77
    layer_norm_eps: float = 1e-5


class _Attn(nn.Module):
    def __init__(self, C: int):
        super().__init__()
        self.k_proj, self.v_proj = nn.Linear(C, C), nn.Linear(C, C)
        self.q_proj, self.out_proj = nn.Linear(C, C), nn.Linear(C, C)


class _Mlp(nn.Module):
    def __init__(self, C: int, F: int):
        super().__init__()
        self.fc1, self.fc2 = nn.Linear(C, F), nn.Linear(F, C)


class _Layer(nn.Module):
    def __init__(self, cfg: ClipTextConfig):
        super().__init__()
        C = cfg.hidden_size
        self.self_attn = _Attn(C)
        self.layer_norm1 = nn.LayerNorm(C, eps=cfg.layer_norm_eps)
        self.mlp = _Mlp(C, cfg.intermediate_size)
        self.layer_norm2 = nn.LayerNorm(C, eps=cfg.layer_norm_eps)
        self._cache = _pack.PackCache()

    def packed(self):
        c, a, m = self._cache, self.self_attn, self.mlp
        qkv_w = [a.q_proj.weight, a.k_proj.weight, a.v_proj.weight]
        qkv_b = [a.q_proj.bias, a.k_proj.bias, a.v_proj.bias]
        return {
            "wqkv": c.get("wqkv", qkv_w, lambda: _pack.linear_w(torch.cat([_pack.d(w) for w in qkv_w], 0))),
            "bqkv": c.get("bqkv", qkv_b, lambda: _pack.f32(torch.cat([_pack.d(b) for b in qkv_b], 0))),
            "wo": c.get("wo", [a.out_proj.weight], lambda: _pack.linear_w(a.out_proj.weight)),
            "bo": c.get("bo", [a.out_proj.bias], lambda: _pack.f32(a.out_proj.bias)),
            # quick-GELU through the SiLU epilogue: fc1 scaled by 1.702, the 1/1.702 folded into fc2
            "w1": c.get("w1", [m.fc1.weight], lambda: _pack.linear_w(_pack.d(m.fc1.weight) * QUICK_GELU)),
            "b1": c.get("b1", [m.fc1.bias], lambda: _pack.f32(_pack.d(m.fc1.bias) * QUICK_GELU)),
            "w2": c.get("w2", [m.fc2.weight], lambda: _pack.linear_w(_pack.d(m.fc2.weight) / QUICK_GELU)),
            "b2": c.get("b2", [m.fc2.bias], lambda: _pack.f32(m.fc2.bias)),
            "g1": c.get("g1", [self.layer_norm1.weight], lambda: _pack.f32(self.layer_norm1.weight)),
            "e1": c.get("e1", [self.layer_norm1.bias], lambda: _pack.f32(self.layer_norm1.bias)),
            "g2": c.get("g2", [self.layer_norm2.weight], lambda: _pack.f32(self.layer_norm2.weight)),
            "e2": c.get("e2", [self.layer_norm2.bias], lambda: _pack.f32(self.layer_norm2.bias)),
        }


class _Embeddings(nn.Module):
    def __init__(self, cfg: ClipTextConfig):
        super().__init__()
        self.token_embedding = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        self.position_embedding = nn.Embedding(cfg.max_position_embeddings, cfg.hidden_size)


class _Encoder(nn.Module):
    def __init__(self, cfg: ClipTextConfig):
        super().__init__()
        self.layers = nn.ModuleList([_Layer(cfg) for _ in range(cfg.num_hidden_layers)])


class _TextTransformer(nn.Module):
    def __init__(self, cfg: ClipTextConfig):
        super().__init__()
        self.embeddings = _Embeddings(cfg)
        self.encoder = _Encoder(cfg)
        self.final_layer_norm = nn.LayerNorm(cfg.hidden_size, eps=cfg.layer_norm_eps)


class TextEncoderOutput(tuple):
    """``encode_text(ids)[0]`` / ``.last_hidden_state`` like transformers' BaseModelOutputWithPooling."""

    @property
    def last_hidden_state(self) -> Tensor:
        return self[0]


class CLIPTextEncoder(nn.Module):
    """``transformers.CLIPTextModel`` forward (last_hidden_state) on libsdb200 kernels; same state_dict keys."""

    def __init__(self, cfg: Optional[ClipTextConfig] = None):
        super().__init__()
        self.cfg = cfg or ClipTextConfig()
        assert self.cfg.hidden_size % self.cfg.num_attention_heads == 0
        self.text_model = _TextTransformer(self.cfg)
        self._cache = _pack.PackCache()

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        state_dict.pop(prefix + "text_model.embeddings.position_ids", None)   # a buffer in older transformers
        return super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    def forward(self, input_ids: Tensor) -> TextEncoderOutput:
        cfg, tm = self.cfg, self.text_model
        _pack.require_cuda(tm.final_layer_norm.weight, "CLIPTextEncoder (call .cuda() first)")
        if input_ids.dim() != 2 or input_ids.shape[1] > cfg.max_position_embeddings:
            raise ValueError(f"input_ids must be [batch, <= {cfg.max_position_embeddings}], got {tuple(input_ids.shape)}")
        dev = tm.final_layer_norm.weight.device
        ids = input_ids.to(device=dev, dtype=torch.int64).contiguous()
        B, L = ids.shape
        C, H = cfg.hidden_size, cfg.num_attention_heads
        d = C // H
        cd = _pack.compute_dtype()
        c = self._cache
        emb = tm.embeddings
        tok = c.get("tok", [emb.token_embedding.weight], lambda: _pack.f32(emb.token_embedding.weight))
        pos = c.get("pos", [emb.position_embedding.weight], lambda: _pack.f32(emb.position_embedding.weight))
        x = torch.empty(B * L, C, device=dev, dtype=cd)
        check(_lib.lib().sdb_embed_tokens(ops._ptr(ids), ops._ptr(tok), ops._ptr(pos), ops._ptr(x), ops._dt(x), B, L, C,
                                          cfg.vocab_size, ops._stream()), "embed_tokens")
        eps = cfg.layer_norm_eps
        for layer in tm.encoder.layers:
            p = layer.packed()
            h = ops.add_layernorm(x, None, p["g1"], p["e1"], eps)
            qkv = ops.gemm(h, p["wqkv"], p["bqkv"]).view(B, L, 3 * C)
            a = torch.empty(B, L, C, device=dev, dtype=cd)
            check(_lib.lib().sdb_attention_causal_small(
                ops._ptr(qkv[:, :, :C]), 3 * C, ops._ptr(qkv[:, :, C:2 * C]), 3 * C, ops._ptr(qkv[:, :, 2 * C:]), 3 * C,
                ops._ptr(a), C, B, H, L, d, d ** -0.5, ops._dt(a), ops._stream()), "attention_causal_small")
            x = ops.gemm(a.view(B * L, C), p["wo"], p["bo"], residual=x)
            h = ops.add_layernorm(x, None, p["g2"], p["e2"], eps)
            f = ops.gemm(h, p["w1"], p["b1"], act=ACT_SILU)
            x = ops.gemm(f, p["w2"], p["b2"], residual=x)
        g = c.get("gf", [tm.final_layer_norm.weight], lambda: _pack.f32(tm.final_layer_norm.weight))
        e = c.get("ef", [tm.final_layer_norm.bias], lambda: _pack.f32(tm.final_layer_norm.bias))
        y = ops.add_layernorm(x, None, g, e, eps).view(B, L, C)
        return TextEncoderOutput((y if y.dtype == torch.float32 else ops.cast(y, torch.float32),))


class CLIPModel(nn.Module):
    """Duck type of the reference's ``CLIPModel`` (clip_model.py:56-96): ``.tokenize(prompts).input_ids`` and
    ``.encode_text(ids)[0]``.  ``tokenizer`` is any callable ``(list[str], max_length) -> LongTensor [B, max_length]``
    (e.g. a ``transformers.CLIPTokenizer`` wrapped in a lambda); without one ``tokenize`` raises -- the BPE vocabulary
    is a download, not part of this library."""

    def __init__(self, cfg: Optional[ClipTextConfig] = None, max_seq_len: int = 77,
                 tokenizer: Optional[Callable] = None):
        super().__init__()
        self.max_seq_len = max_seq_len
        self.text_encoder = CLIPTextEncoder(cfg)
        self.tokenizer = tokenizer

    def tokenize(self, prompt="", max_length: Optional[int] = None, padding: str = "max_length",
                 truncation: bool = True):
        if self.tokenizer is None:
            raise RuntimeError("CLIPModel.tokenize: no tokenizer was given (CLIPTokenizer needs its vocabulary files)")
        ids = self.tokenizer([prompt] if isinstance(prompt, str) else list(prompt),
                             self.max_seq_len if max_length is None else max_length)

        class _Tokens:
            pass
        t = _Tokens()
        t.input_ids = torch.as_tensor(ids, dtype=torch.long)
        return t

    @torch.no_grad()
    def encode_text(self, text: Tensor) -> TextEncoderOutput:
        return self.text_encoder(text)
