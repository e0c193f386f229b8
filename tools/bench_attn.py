"""Times the self-attention kernels at the config-2 shape (UNet batch 32, 8 heads, 4096 tokens, d=20)."""
import math, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_pytorch_b200 import ops

def timeit(fn, it=10, warm=3):
    for _ in range(warm): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

B, H, n, d = int(os.environ.get("B", 32)), 8, int(os.environ.get("N", 4096)), 20
qkv = torch.randn(B, n, 3 * H * 32, device="cuda").to(torch.bfloat16)
qkv.view(B, n, 3, H, 32)[..., d:] = 0
qkv.view(B, n, 3, H, 32)[:, :, 2, :, d] = 1
q, k, v = qkv[:, :, :256], qkv[:, :, 256:512], qkv[:, :, 512:]
ms = timeit(lambda: ops.attention_p32(q, k, v, H, d, 1 / math.sqrt(d)))
pairs = B * H * n * n
print(f"attention_p32 poly={os.environ.get('SDB_FA_POLY','default')}: {ms:.3f} ms  {pairs/ms/1e9:.2f} Gpairs/ms... "
      f"{pairs/(ms*1e-3)/148/1.965e9:.2f} pairs/clk/SM  {4*pairs*d/(ms*1e-3)/1e12:.1f} TFLOP/s(alg)")
# cross-attention over a 77-token context (same queries)
ctx = torch.randn(B, 77, 2 * H * 32, device="cuda").to(torch.bfloat16)
ctx.view(B, 77, 2, H, 32)[..., d:] = 0
ctx.view(B, 77, 2, H, 32)[:, :, 1, :, d] = 1
kc, vc = ctx[:, :, :256], ctx[:, :, 256:]
ms = timeit(lambda: ops.attention_p32(q, kc, vc, H, d, 1 / math.sqrt(d)))
by = (q.numel() + B * n * H * d) * 2
print(f"attention_p32 m=77: {ms*1e3:.1f} us  {by/ms/1e6:.0f} GB/s (Q in + O out)")
if os.environ.get("OLD"):
    qkv2 = torch.randn(B, n, 480, device="cuda").to(torch.bfloat16)
    ms = timeit(lambda: ops.attention(qkv2[:, :, :160], qkv2[:, :, 160:320], qkv2[:, :, 320:], H, d, 1 / math.sqrt(d)))
    print(f"attention (mma.sync): {ms:.3f} ms")
print("timeout flag", ops.debug_timeout_flag())
if os.environ.get("BWD"):
    # backward of the self-attention at the same shape (dQ and dK/dV on tcgen05, lse handed over by the forward)
    from stable_diffusion_pytorch_b200 import autograd_ops
    Bb = int(os.environ.get("BB", 64))
    qkvb = qkv[:1].expand(Bb, -1, -1).contiguous() if Bb != B else qkv
    qb, kb, vb = (t.detach().requires_grad_(True) for t in (qkvb[:, :, :256], qkvb[:, :, 256:512], qkvb[:, :, 512:]))
    out = autograd_ops.attention_p32(qb, kb, vb, H, d, 1 / math.sqrt(d))
    g = torch.randn_like(out)
    ms = timeit(lambda: torch.autograd.grad(out, (qb, kb, vb), g, retain_graph=True), it=5, warm=2)
    print(f"attention_p32 backward (B={Bb}): {ms:.3f} ms  (dq + dkv kernels + delta / pad glue)")
    print("timeout flag", ops.debug_timeout_flag())
