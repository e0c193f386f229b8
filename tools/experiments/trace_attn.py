"""One traced launch of the tcgen05 attention kernel (SDB_FA_TRACE=1): prints CTA 0's phase stamps."""
import math, os, sys, torch
os.environ["SDB_FA_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_pytorch_b200 import ops
B, H, n, d = 32, 8, 4096, 20
qkv = torch.randn(B, n, 3 * H * 32, device="cuda").to(torch.bfloat16)
qkv.view(B, n, 3, H, 32)[..., d:] = 0
qkv.view(B, n, 3, H, 32)[:, :, 2, :, d] = 1
q, k, v = qkv[:, :, :256], qkv[:, :, 256:512], qkv[:, :, 512:]
for _ in range(2):
    ops.attention_p32(q, k, v, H, d, 1 / math.sqrt(d))
torch.cuda.synchronize()
