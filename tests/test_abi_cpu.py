"""CPU: the C-ABI library builds, loads, and exports every symbol include/sdb200.h declares
(no compute calls -- there is no GPU here)."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "sdb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sdb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from stable_diffusion_pytorch_b200 import _lib, build
    build.build()
    handle = _lib.lib()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(handle, n), f"{n} declared in sdb200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(names)
    assert handle.sdb_version() == 100


def test_argument_errors_surface_without_gpu():
    import ctypes as C
    from stable_diffusion_pytorch_b200 import _lib
    L = _lib.lib()
    # NULL pointers are rejected before any CUDA call: mirrors the reference's asserts -> SDB_EINVAL
    assert L.sdb_gemm(None, None) == -1
    assert b"null" in L.sdb_last_error()
    assert L.sdb_attention(None, 0, None, 0, None, 0, None, 0, 1, 1, 1, 1, 20, 1.0, 1, None) == -1
    assert L.sdb_groupnorm_chunks(4096) == 64 and L.sdb_groupnorm_chunks(256) == 4 and L.sdb_groupnorm_chunks(1) == 1


def test_no_cpu_fallback():
    import pytest
    import torch
    from stable_diffusion_pytorch_b200 import ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.silu(torch.zeros(4))
