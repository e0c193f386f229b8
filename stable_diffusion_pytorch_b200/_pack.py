"""Device-side weight packing shared by the modules.

Parameters keep the reference's names, shapes and fp32 OIHW / [out, in] layouts (so a reference
``state_dict`` loads unchanged); the kernels want bf16, K-major, channels-last operands.  Packed
copies are built lazily on first use and rebuilt whenever a source parameter changes (version
counter / storage / device), e.g. after ``load_state_dict`` or an optimizer step.
"""
from __future__ import annotations

from typing import Callable, Dict, Sequence, Tuple

import torch
from torch import nn

Tensor = torch.Tensor


# Bumped whenever parameters are rewritten behind torch's back (the fused AdamW kernel updates them through raw
# pointers, so ``_version`` does not move): every packed copy and every captured graph keyed on it is then rebuilt.
_EPOCH = 0


def epoch() -> int:
    return _EPOCH


def bump_epoch() -> int:
    global _EPOCH
    _EPOCH += 1
    return _EPOCH


class PackCache:
    def __init__(self):
        self._store: Dict[str, Tuple[tuple, object]] = {}

    @staticmethod
    def _sig(params: Sequence[Tensor]) -> tuple:
        return (_COMPUTE, _EPOCH) + tuple((p.data_ptr(), p._version, str(p.device), p.dtype) for p in params)

    def get(self, key: str, params: Sequence[Tensor], builder: Callable[[], object]):
        if _GRAPH:          # training graph: fresh, differentiable packing every forward
            return builder()
        sig = self._sig(params)
        hit = self._store.get(key)
        if hit is not None and hit[0] == sig:
            return hit[1]
        with torch.no_grad():
            val = builder()
        self._store[key] = (sig, val)
        return val

    def clear(self):
        self._store.clear()


# Training (T1): the packing below must stay inside torch's autograd graph -- the gradient of a packed
# operand (bf16, permuted / padded / interleaved) flows back to the fp32 parameter through torch's own
# permute / cat / cast backward (weight-layout plumbing on 33 M parameters, not activation compute).
# ``with graph_mode():`` switches ``d()`` from detach to identity and makes PackCache build fresh.
_GRAPH = False


class graph_mode:
    def __enter__(self):
        global _GRAPH
        self._prev, _GRAPH = _GRAPH, True
        return self

    def __exit__(self, *exc):
        global _GRAPH
        _GRAPH = self._prev
        return False


def in_graph_mode() -> bool:
    return _GRAPH


# Compute precision of the hot path.  bf16 (default) = the tcgen05 tensor-core kernels, what bench.py measures.
# fp32 = the parity mode north_star asks for ("a single UNet forward within 1e-3 relative L2 in fp32"; the
# reference run with `--dtype float32`, scripts/txt2img.py:23): activations and packed weights stay fp32 and
# GEMM / conv / attention run as fp32 FMAs on the CUDA cores (csrc/precise_f32.cu).  Inference only.
_COMPUTE = torch.bfloat16


def compute_dtype() -> torch.dtype:
    return _COMPUTE


def set_precision(dtype) -> torch.dtype:
    """``"bf16"`` / ``torch.bfloat16`` or ``"fp32"`` / ``torch.float32``; returns the previous setting."""
    global _COMPUTE
    table = {"bf16": torch.bfloat16, "bfloat16": torch.bfloat16, "fp32": torch.float32, "float32": torch.float32,
             torch.bfloat16: torch.bfloat16, torch.float32: torch.float32}
    if dtype not in table:
        raise ValueError(f"precision must be bf16 or fp32, got {dtype!r}")
    prev, _COMPUTE = _COMPUTE, table[dtype]
    return prev


class precision:
    """``with precision("fp32"): unet(x, t, ctx)`` -- scoped ``set_precision``."""

    def __init__(self, dtype):
        self._dtype = dtype

    def __enter__(self):
        self._prev = set_precision(self._dtype)
        return self

    def __exit__(self, *exc):
        set_precision(self._prev)
        return False


def d(p: Tensor) -> Tensor:
    return p if _GRAPH else p.detach()


def f32(p: Tensor) -> Tensor:
    return d(p).to(torch.float32).contiguous()


def linear_w(p: Tensor) -> Tensor:
    """[out, in] -> bf16 [N, K] (already K-major)."""
    return d(p).to(_COMPUTE).contiguous()


def conv1x1_w(p: Tensor) -> Tensor:
    """[Cout, Cin, 1, 1] -> bf16 [Cout, Cin]."""
    return d(p).reshape(p.shape[0], p.shape[1]).to(_COMPUTE).contiguous()


def conv3x3_w(p: Tensor) -> Tensor:
    """OIHW [Cout, Cin, 3, 3] -> bf16 [Cout, 3, 3, Cin] (tap-major, channels innermost)."""
    return d(p).permute(0, 2, 3, 1).to(_COMPUTE).contiguous()


def conv3x3_up_w(p: Tensor) -> Tensor:
    """OIHW [Cout, Cin, 3, 3] -> phase-collapsed weights of ``UpSample`` (nearest x2 then conv3x3, resnet2d.py:61-62),
    [4, Cout, 2, 2, Cin] in the compute dtype.  Output pixel (2y+a, 2x+b) of the upsampled conv reads source rows
    {y-1, y} when a = 0 (3x3 rows {0}, {1, 2}) and {y, y+1} when a = 1 (rows {0, 1}, {2}); columns likewise.  The taps
    that fall on the same source pixel are summed in fp32, then rounded once."""
    w = d(p).to(torch.float32)
    sets = (([0], [1, 2]), ([0, 1], [2]))
    phases = []
    for a in range(2):
        for b in range(2):
            taps = [[sum(w[:, :, ky, kx] for ky in sets[a][i] for kx in sets[b][j]) for j in range(2)] for i in range(2)]
            phases.append(torch.stack([torch.stack(r, dim=1) for r in taps], dim=1))      # [Cout, 2, 2, Cin]
    return torch.stack(phases).to(_COMPUTE).contiguous()


def geglu_w(w: Tensor, b: Tensor) -> Tuple[Tensor, Tensor]:
    """GEGLU proj [2*F, C] with halves (a | gate) -> rows interleaved in pairs (a0, a1, g0, g1, a2, a3, g2, g3,
    ...) so one accumulator tile holds both operands of a_j * gelu(g_j) and the epilogue's packed fp32x2
    arithmetic works on adjacent registers."""
    F2 = w.shape[0]
    half = F2 // 2
    assert half % 2 == 0
    wd, bd = d(w), d(b)
    wi = torch.stack([wd[:half].reshape(half // 2, 2, -1), wd[half:].reshape(half // 2, 2, -1)], dim=1)
    bi = torch.stack([bd[:half].reshape(half // 2, 2), bd[half:].reshape(half // 2, 2)], dim=1)
    return wi.reshape(F2, w.shape[1]).to(_COMPUTE).contiguous(), bi.reshape(F2).to(torch.float32).contiguous()


HEAD_PAD = 32  # padded head width of the tcgen05 attention operands (sdb_attention_p32)


def head_pad_w(w: Tensor, heads: int, d: int) -> Tensor:
    """Projection weight [heads*d, K] -> [heads*32, K]: head h's rows move to [32h, 32h+d), the rest are
    zero, so the GEMM writes zero-padded 64-byte head rows (what the attention TMA boxes load)."""
    K = w.shape[1]
    pad = torch.zeros(heads, HEAD_PAD - d, K, dtype=w.dtype, device=w.device)
    return torch.cat([globals()["d"](w).reshape(heads, d, K), pad], dim=1).reshape(heads * HEAD_PAD, K)


def head_pad_ones_bias(heads: int, d: int, device) -> Tensor:
    """fp32 bias [heads*32], 1.0 at column d of every head: the V projection then carries the
    column of ones that makes P.V also produce the softmax denominator."""
    b = torch.zeros(heads, HEAD_PAD, dtype=torch.float32, device=device)
    b[:, d] = 1.0
    return b.reshape(heads * HEAD_PAD)


def require_cuda(x: Tensor, what: str) -> None:
    if not x.is_cuda:
        raise RuntimeError(f"{what}: stable_diffusion_pytorch_b200 has no CPU fallback; got a {x.device} tensor")


def zero_module(module: nn.Module) -> nn.Module:
    """Same contract as the reference's models/utils.py:22-28."""
    for p in module.parameters():
        p.detach().zero_()
    return module


def head_w16(w: Tensor, b: Tensor) -> Tuple[Tensor, Tensor]:
    """Output-head conv OIHW [Cout <= 16, Cin, 3, 3] -> bf16 [16, 3, 3, Cin] with zero rows for the padding
    channels (the tensor-core conv needs Cout % 16 == 0), bias -> fp32 [16]."""
    cout = w.shape[0]
    assert cout <= 16
    wp = torch.cat([conv3x3_w(w), torch.zeros(16 - cout, 3, 3, w.shape[1], device=w.device, dtype=_COMPUTE)])
    bp = torch.cat([d(b).float(), torch.zeros(16 - cout, device=w.device, dtype=torch.float32)])
    return wp.contiguous(), bp
