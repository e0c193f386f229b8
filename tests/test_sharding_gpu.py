"""GPU (>= 2 devices): the sampling path's one collective over NCCL -- tools/check_sharded_nccl.py under torchrun:
`sharded_images` with the real LatentDiffusion on two ranks must equal the single-GPU run bit for bit.  Skipped on a
one-GPU box (the world_size-2 gloo tests in test_sharding_cpu.py cover the host logic there)."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharded_images_nccl_world2_equals_single_gpu():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tools", "check_sharded_nccl.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    line = json.loads([l for l in res.stdout.splitlines() if l.startswith("{")][-1])
    assert line["world"] == 2 and line["latents_bit_equal_all_ranks"] and line["images_bit_equal_all_ranks"]
    assert line["shards"] == [3, 2] and line["backend"] == "nccl"
