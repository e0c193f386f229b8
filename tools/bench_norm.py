"""Times the HBM-bound normalisation kernels at the UNet's config-2 shapes (UNet batch 32)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_pytorch_b200 import ops

def timeit(fn, it=20, warm=3):
    for _ in range(warm): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

B = 32
print(f"{'op':44s} {'ms':>8s} {'GB/s':>8s}  count  total_ms")
tot = 0.0
def gn(name, H, C1, C2, count, act=1):
    global tot
    x1 = torch.randn(B, H, H, C1, device="cuda").to(torch.bfloat16)
    x2 = torch.randn(B, H, H, C2, device="cuda").to(torch.bfloat16) if C2 else None
    g = torch.ones(C1 + C2, device="cuda"); b = torch.zeros(C1 + C2, device="cuda")
    sc = ops.GroupNormScratch()
    ms = timeit(lambda: ops.group_norm(x1, g, b, 32, act, x2=x2, scratch=sc))
    by = 2.0 * B * H * H * (C1 + C2) * 2      # algorithmic: one read + one write (the three-kernel route reads twice)
    old = ops._GN_FUSED
    ops._GN_FUSED = False
    ms3 = timeit(lambda: ops.group_norm(x1, g, b, 32, act, x2=x2, scratch=sc))
    ops._GN_FUSED = old
    print(f"{name:44s} {ms:8.4f} {by/ms/1e6:8.1f}  {count:5d}  {ms*count:7.3f}   (three-kernel route {ms3:.4f} ms)")
    tot += ms * count
def ln(name, rows, C, count, res=False):
    global tot
    x = torch.randn(rows, C, device="cuda").to(torch.bfloat16)
    r = torch.randn(rows, C, device="cuda").to(torch.bfloat16) if res else None
    g = torch.ones(C, device="cuda"); b = torch.zeros(C, device="cuda")
    ms = timeit(lambda: ops.add_layernorm(x, r, g, b))
    by = 2.0 * rows * C * (3 if res else 2)
    print(f"{name:44s} {ms:8.4f} {by/ms/1e6:8.1f}  {count:5d}  {ms*count:7.3f}")
    tot += ms * count
gn("GN+SiLU 160 @64 (stats+apply)", 64, 160, 0, 13)
gn("GN 160 @64 no act (transformer in)", 64, 160, 0, 5, act=0)
gn("GN+SiLU 320 @32", 32, 320, 0, 12)
gn("GN+SiLU 320+320 @32 (virtual concat)", 32, 320, 320, 2)
gn("GN+SiLU 320+160 @32", 32, 320, 160, 1)
gn("GN+SiLU 320+160 @64", 64, 320, 160, 1)
gn("GN+SiLU 160+160 @64", 64, 160, 160, 2)
ln("LN 160 rows=131072", 131072, 160, 30)
ln("LN 320 rows=32768", 32768, 320, 6)
ln("add+LN 160 rows=131072", 131072, 160, 0, res=True)
print("sum of (ms x count):", round(tot, 3), "ms")
# thin 4-channel ends (CUDA-core kernels, NCHW <-> NHWC)
x = torch.randn(16, 4, 64, 64, device="cuda").to(torch.bfloat16)
w = torch.randn(160, 4, 3, 3, device="cuda") * 0.1
bi = torch.zeros(160, device="cuda")
ms = timeit(lambda: ops.conv3x3_in(x, w, bi, torch.bfloat16, repeat=2))
print(f"{'conv3x3_in 4->160 @64 (x2 virtual repeat)':44s} {ms:8.4f} {2.0*B*64*64*160/ms/1e6:8.1f}")
y = torch.randn(B, 64, 64, 160, device="cuda").to(torch.bfloat16)
wo = torch.randn(4, 160, 3, 3, device="cuda") * 0.05
bo = torch.zeros(4, device="cuda")
ms = timeit(lambda: ops.conv3x3_out(y, wo, bo, torch.float32))
print(f"{'conv3x3_out 160->4 @64':44s} {ms:8.4f} {2.0*B*64*64*160/ms/1e6:8.1f}")
