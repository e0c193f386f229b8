import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_pytorch_b200 import ops
def timeit(fn, it=30, warm=5):
    for _ in range(warm): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
M = 131072
for K in (160, 640):
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(160, K, device="cuda") * 0.05).to(torch.bfloat16)
    b = torch.randn(160, device="cuda")
    r = torch.randn(M, 160, device="cuda").to(torch.bfloat16)
    g = torch.ones(160, device="cuda"); be = torch.zeros(160, device="cuda")
    out = torch.empty(M, 160, device="cuda", dtype=torch.bfloat16)
    t0 = timeit(lambda: ops.gemm(a, w, b, residual=r, out=out))
    t1 = timeit(lambda: ops.gemm(a, w, b, residual=r, out=out, ln=(g, be)))
    print(f"K={K}: +res {t0*1e3:.1f} us   +res+LN {t1*1e3:.1f} us")
