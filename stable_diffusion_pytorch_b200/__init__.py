"""B200-native (sm_100a) latent-diffusion hot path behind the reference's
``stable_diffusion.models`` / ``stable_diffusion.modules`` API.

PyTorch provides device memory, streams and ``torch.distributed``; all arithmetic runs in the
hand-written CUDA kernels of ``libsdb200.so`` (C ABI: ``include/sdb200.h``).  No CPU fallback.
"""
__version__ = "0.1.0"

from ._pack import compute_dtype, precision, set_precision  # noqa: E402,F401  (bf16 product path / fp32 parity mode)
