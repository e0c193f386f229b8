"""Pipeline trace of the tcgen05 GEMM kernel: builds a private copy of the library with -DSDB_GEMM_TRACE
(tools/_trace/libsdb200.so), runs one GEMM shape and prints, for CTA 0's first tiles, when each role
(TMA producer, MMA issuer, epilogue DMA warp, the two epilogue warpgroups) passed its hand-off points.
  python tools/trace_gemm.py build      # here (needs nvcc only)
  M=131072 K=160 N=1280 ACT=2 python tools/trace_gemm.py   # on the GPU box"""
import ctypes, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tools", "_trace", "libsdb200.so")
if len(sys.argv) > 1 and sys.argv[1] == "build":
    from stable_diffusion_pytorch_b200 import build as b
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = [b._nvcc(), *b.NVCC_FLAGS, "-DSDB_GEMM_TRACE", "-o", OUT, os.path.join(b.CSRC, "sdb200.cu")]
    subprocess.run(cmd, check=True, cwd=b.CSRC)
    print("built", OUT)
    sys.exit(0)
import torch
from stable_diffusion_pytorch_b200 import _lib
_lib.LIB_PATH = OUT
from stable_diffusion_pytorch_b200 import ops
M, K, N, ACT, RES = (int(os.environ.get(k, d)) for k, d in (("M", 131072), ("K", 160), ("N", 1280), ("ACT", 2), ("RES", 0)))
a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
b = torch.zeros(N, device="cuda")
r = torch.randn(M, N // 2 if ACT == 2 else N, device="cuda").to(torch.bfloat16) if RES else None
for _ in range(3):
    ops.gemm(a, w, b, r, ACT)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (5 * 16 * 8))()
assert _lib.lib().sdb_debug_gemm_trace(buf) == 0
t = list(buf)
t0 = min(x for x in t if x > 0)
names = ["producer (tile start, loads issued)", "MMA (wait tempty, got it, all issued)",
         "DMA (top, prepped next, got stg_full, stores issued)", "WG0 (top, res_full, tfull, done)", "WG1 (top, res_full, tfull, done)"]
for role in range(5):
    print(names[role])
    for j in range(12):
        ev = t[(role * 16 + j) * 8:(role * 16 + j) * 8 + 4]
        print("   ", j if role < 3 else 2 * j + (role - 3), " ".join(f"{(x - t0) if x else -1:8d}" for x in ev))
cta = (ctypes.c_longlong * (160 * 4))()
assert _lib.lib().sdb_debug_gemm_cta_ns(cta) == 0
c = list(cta)
rows = [(c[4 * i], c[4 * i + 1], c[4 * i + 2], c[4 * i + 3]) for i in range(160) if c[4 * i]]
g0 = min(r[0] for r in rows)
print("CTAs", len(rows), "kernel span (ns)", max(r[1] for r in rows) - g0)
print("start offsets ns  min/max", min(r[0] - g0 for r in rows), max(r[0] - g0 for r in rows))
dur = sorted(r[1] - r[0] for r in rows)
print("CTA duration ns  min/median/max", dur[0], dur[len(dur) // 2], dur[-1])
clk = sorted((r[3] - r[2]) / max(r[1] - r[0], 1) for r in rows)
print("clock64 per ns  min/median/max", clk[0], clk[len(clk) // 2], clk[-1])
print("CTA0: ns", rows[0][1] - rows[0][0], "clk", rows[0][3] - rows[0][2])
