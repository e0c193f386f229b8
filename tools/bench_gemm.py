"""Times the tcgen05 GEMM / implicit-GEMM conv kernel at the UNet's config-2 shapes (UNet batch 32)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_pytorch_b200 import ops

def timeit(fn, it=10, warm=3):
    for _ in range(warm): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

B = 32
dev = "cuda"
tot = 0.0
print(f"{'op':34s} {'ms':>8s} {'TFLOP/s':>8s} {'GB/s':>8s}  count  total_ms")
ONLY = os.environ.get("ONLY")   # substring filter, e.g. ONLY=geglu for an ncu capture of one shape
def conv(name, H, cin, cout, count, down=False, temb=False, res=False):
    global tot
    if ONLY and ONLY not in name: return
    x = torch.randn(B, H, H, cin, device=dev).to(torch.bfloat16)
    w = (torch.randn(cout, 3, 3, cin, device=dev) * 0.02).to(torch.bfloat16)
    b = torch.zeros(cout, device=dev)
    t = torch.randn(B, cout, device=dev) if temb else None
    Ho = H // 2 if down else H
    r = torch.randn(B, Ho, Ho, cout, device=dev).to(torch.bfloat16) if res else None
    ms = timeit(lambda: ops.conv3x3(x, w, b, t, r, down2=down))
    fl = 2.0 * B * H * H * 9 * cin * cout
    by = 2.0 * (x.numel() + w.numel() + B * Ho * Ho * cout * (2 if res else 1))
    print(f"{name:34s} {ms:8.3f} {fl/ms/1e9:8.1f} {by/ms/1e6:8.1f}  {count:5d}  {ms*count:7.3f}")
    tot += ms * count
def gemm(name, M, K, N, count, act=0, res=False, bias=True):
    global tot
    if ONLY and ONLY not in name: return
    a = torch.randn(M, K, device=dev).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
    b = torch.zeros(N, device=dev) if bias else None
    no = N // 2 if act == 2 else N
    r = torch.randn(M, no, device=dev).to(torch.bfloat16) if res else None
    ms = timeit(lambda: ops.gemm(a, w, b, r, act))
    fl = 2.0 * M * K * N
    by = 2.0 * (a.numel() + w.numel() + M * no * (2 if res else 1))
    print(f"{name:34s} {ms:8.3f} {fl/ms/1e9:8.1f} {by/ms/1e6:8.1f}  {count:5d}  {ms*count:7.3f}")
    tot += ms * count
M0, M1 = B * 64 * 64, B * 32 * 32
conv("conv 160->160 @64 (+temb)", 64, 160, 160, 4, temb=True)
conv("conv 160->160 @64 (+res)", 64, 160, 160, 7, res=True)
conv("conv 160->160 @64 down2", 64, 160, 160, 1, down=True)
conv("conv 160->320 @32", 32, 160, 320, 1, temb=True)
conv("conv 320->320 @32", 32, 320, 320, 11, res=True)
conv("conv 640->320 @32", 32, 640, 320, 2, temb=True)
conv("conv 480->320 @32", 32, 480, 320, 1, temb=True)
conv("conv 320->320 @64 (upsample)", 64, 320, 320, 1)
conv("conv 480->160 @64", 64, 480, 160, 1, temb=True)
conv("conv 320->160 @64", 64, 320, 160, 2, temb=True)
gemm("qkv 160->768 (M=131072)", M0, 160, 768, 10, bias=True)
gemm("q 160->256", M0, 160, 256, 10, bias=False)
gemm("attn-out 160->160 +res", M0, 160, 160, 20, res=True)
gemm("proj_in/out 160->160", M0, 160, 160, 10, res=True)
gemm("geglu 160->1280", M0, 160, 1280, 10, act=2)
gemm("ffn-out 640->160 +res", M0, 640, 160, 10, res=True)
gemm("skip1x1 480->160", M0, 480, 160, 1)
gemm("skip1x1 320->160", M0, 320, 160, 2)
gemm("mid qkv 320->768 (M=32768)", M1, 320, 768, 2)
gemm("mid geglu 320->2560", M1, 320, 2560, 2, act=2)
gemm("mid ffn-out 1280->320", M1, 1280, 320, 2, res=True)
print("sum of (ms x count):", round(tot, 3), "ms   timeout flag", ops.debug_timeout_flag())
