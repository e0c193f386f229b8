"""Dense attention at the AutoencoderKL bottleneck shape (config 3: 16 x 4096 tokens, one head of 128): ms and TFLOP/s.
SDB_ATTN_D128_UMMA=0 times the mma.sync kernel instead."""
import math, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_pytorch_b200 import ops
B, n, d = int(os.environ.get("B", 16)), int(os.environ.get("N", 4096)), 128
g = torch.Generator().manual_seed(0)
qkv = torch.randn(B, n, 3 * d, generator=g).to("cuda", torch.bfloat16)
q, k, v = qkv[:, :, :d], qkv[:, :, d:2 * d], qkv[:, :, 2 * d:]
for _ in range(3):
    o = ops.attention(q, k, v, 1, d, 1 / math.sqrt(d))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(10):
    o = ops.attention(q, k, v, 1, d, 1 / math.sqrt(d))
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"attention d=128 B={B} n={n}: {ms:.3f} ms  {4.0 * B * n * n * d / ms / 1e9:.1f} TFLOP/s  timeout flag {ops.debug_timeout_flag()}")
