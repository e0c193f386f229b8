"""ctypes binding of ``libsdb200.so`` (the C ABI declared in ``include/sdb200.h``).

There is no CPU fallback: if the shared library has not been built the import of any op fails
loudly (``python -c 'import __graft_entry__ as g; g.build()'`` builds it in-tree).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsdb200.so")

SDB_F32, SDB_BF16, SDB_U8 = 0, 1, 2
ACT_NONE, ACT_SILU, ACT_GEGLU = 0, 1, 2

_p, _i, _i64, _f, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint64


class SdbGemmArgs(C.Structure):
    _fields_ = [("a", _p), ("lda", _i64), ("K1", _i),
                ("a2", _p), ("lda2", _i64), ("K2", _i),
                ("w", _p), ("bias", _p),
                ("residual", _p), ("ldr", _i64), ("res_dtype", _i),
                ("out", _p), ("ldo", _i64), ("out_dtype", _i),
                ("M", _i), ("N", _i), ("act", _i),
                ("ln_gamma", _p), ("ln_beta", _p), ("ln_eps", _f),
                ("chan_sums", _p), ("rows_per_image", _i64)]


class SdbConvArgs(C.Structure):
    _fields_ = [("x", _p), ("w", _p), ("bias", _p),
                ("temb", _p), ("ld_temb", _i64),
                ("residual", _p), ("res_dtype", _i),
                ("out", _p), ("out_dtype", _i),
                ("B", _i), ("H", _i), ("W", _i), ("Cin", _i), ("Cout", _i),
                ("down2", _i), ("up2", _i), ("out_channels", _i),
                ("chan_sums", _p)]


# name -> (restype, argtypes); must list every symbol include/sdb200.h declares
SIGNATURES = {
    "sdb_version": (_i, []),
    "sdb_last_error": (C.c_char_p, []),
    "sdb_debug_timeout_flag": (_i, []),
    "sdb_launch_count": (_i64, []),
    "sdb_timestep_embed": (_i, [_p, _p, _p, _i, _i, _i, _p]),
    "sdb_cfg_sampler_step": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _f, _f, _u64, _p, _p, _i, _i64, _i, _p]),
    "sdb_sampler_advance": (_i, [_p, _p, _i, _p, _i, _p]),
    "sdb_add_noise": (_i, [_p, _p, _p, _p, _i, _p, _i, _i64, _i, _p]),
    "sdb_nchw_to_nhwc": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_nhwc_to_nchw": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_groupnorm_chunks": (_i, [_i64]),
    "sdb_groupnorm_stats": (_i, [_p, _i, _p, _i, _i, _i, _i64, _i, _f, _p, _p, _p]),
    "sdb_groupnorm_stats_static": (_i, [_p, _i, _p, _i, _i, _i, _i64, _i, _f, _p, _p]),
    "sdb_groupnorm_fused": (_i, [_p, _i, _p, _i, _p, _p, _i, _p, _i, _i64, _i, _f, _p]),
    "sdb_groupnorm_coop_scratch": (_i64, [_i, _i64, _i]),
    "sdb_groupnorm_coop": (_i, [_p, _i, _p, _i, _p, _p, _i, _p, _i, _i64, _i, _f, _p, _p, _p]),
    "sdb_groupnorm_apply": (_i, [_p, _i, _p, _i, _i, _p, _p, _p, _i, _p, _i, _i, _i64, _i, _p]),
    "sdb_groupnorm_apply_sums": (_i, [_p, _i, _p, _p, _i, _p, _i, _p, _p, _f, _i, _p, _i, _i, _i64, _i, _p]),
    "sdb_add_layernorm": (_i, [_p, _p, _p, _p, _p, _i64, _i, _f, _i, _p]),
    "sdb_gemm": (_i, [C.POINTER(SdbGemmArgs), _p]),
    "sdb_conv3x3": (_i, [C.POINTER(SdbConvArgs), _p]),
    "sdb_gemm_f32": (_i, [C.POINTER(SdbGemmArgs), _p]),
    "sdb_conv3x3_f32": (_i, [C.POINTER(SdbConvArgs), _p]),
    "sdb_conv3x3_in": (_i, [_p, _i, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _p]),
    "sdb_cfg_combine": (_i, [_p, _f, _p, _i64, _i, _p]),
    "sdb_conv3x3_out": (_i, [_p, _i, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "sdb_conv1x1_nchw_small": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "sdb_nhwc_slice_to_nchw": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_upsample2x": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_attention": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _i, _i, _i, _i, _i, _f, _i, _p]),
    "sdb_softmax_rows": (_i, [_p, _i64, _p, _i64, _i, _i64, _i, _i, _f, _p]),
    "sdb_attention_p32": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _i, _i, _i, _i, _i, _f, _p]),
    "sdb_colsum": (_i, [_p, _i, _i64, _i64, _i64, _i, _p, _p]),
    "sdb_geglu_fwd": (_i, [_p, _p, _i64, _i, _p]),
    "sdb_geglu_bwd": (_i, [_p, _p, _p, _i64, _i, _p]),
    "sdb_silu_bwd": (_i, [_p, _i, _p, _p, _i, _i64, _p]),
    "sdb_groupnorm_bwd": (_i, [_p, _i, _p, _i, _i, _p, _p, _p, _i, _p, _p, _p, _p, _p, _p, _i, _i64, _i, _p]),
    "sdb_groupnorm_bwd_vec": (_i, [_p, _i, _p, _i, _i, _p, _p, _p, _i, _p, _p, _p, _p, _p, _p, _p, _i, _i64, _i, _p]),
    "sdb_layernorm_bwd": (_i, [_p, _p, _p, _p, _p, _p, _i64, _i, _f, _i, _p]),
    "sdb_attention_bwd": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _p,
                                _i, _i, _i, _i, _i, _i, _i, _f, _p]),
    "sdb_attention_bwd_dq_p32": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _p, _p, _i64, _i, _i, _i, _i, _i, _f, _p]),
    "sdb_attention_ctx_bwd_dq": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _p, _i, _i, _i, _i, _i, _f, _p]),
    "sdb_attention_bwd_prep": (_i, [_p, _i64, _p, _i64, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p]),
    "sdb_attention_bwd_dkv_p32": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _p, _p, _i64, _p, _i64, _i, _i, _i, _i, _i,
                                        _f, _p]),
    "sdb_attention_bwd_dkv": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _p, _i, _i, _i, _i, _i,
                                    _i, _f, _p]),
    "sdb_attention_p32_lse": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _i, _i, _i, _i, _i, _f, _p, _p, _p]),
    "sdb_upsample2x_bwd": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_scatter_even": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_wgrad_conv3x3": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_wgrad_linear": (_i, [_p, _i64, _p, _i64, _p, _i64, _i, _i, _i, _p]),
    "sdb_im2col3x3": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_nchw_to_nhwc_pad": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _i, _p]),
    "sdb_mse": (_i, [_p, _p, _i, _i64, _f, _p, _p, _p]),
    "sdb_cfg_combine_bwd": (_i, [_p, _p, _i, _i64, _f, _p]),
    "sdb_sumsq": (_i, [_p, _i64, _p, _p]),
    "sdb_adamw": (_i, [_p, _p, _p, _p, _i64, _f, _f, _f, _f, _f, _i, _f, _p, _f, _p]),
    "sdb_embed_tokens": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "sdb_attention_causal_small": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _i, _i, _i, _i, _f, _i, _p]),
    "sdb_adamw_dev": (_i, [_p, _p, _p, _p, _i64, _p, _f, _f, _f, _f, _f, _p, _f, _p]),
    "sdb_dropout": (_i, [_p, _p, _i64, _f, _p, _u64, _i, _p]),
    "sdb_silu": (_i, [_p, _p, _i64, _i, _i, _p]),
    "sdb_cast": (_i, [_p, _i, _p, _i, _i64, _p]),
    "sdb_to_uint8_image": (_i, [_p, _i, _p, _i, _i, _i, _i, _p]),
}

_lib = None


class SdbError(RuntimeError):
    pass


def lib() -> C.CDLL:
    """Loads the library once.  Raises if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SdbError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "stable_diffusion_pytorch_b200 has no CPU / PyTorch fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().sdb_last_error().decode("utf-8", "replace")
        if rc == -1:
            raise ValueError(f"{what}: {msg}")
        raise SdbError(f"{what}: rc={rc}: {msg}")
