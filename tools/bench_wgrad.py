"""Weight-gradient kernel (csrc/wgrad_umma.cu) on the config-4 shapes: device time (CUDA events) and TFLOP/s.

    python tools/bench_wgrad.py [--batch 64] [--iters 20]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    from stable_diffusion_pytorch_b200 import ops
    dev = torch.device("cuda")
    B = a.batch
    rows = []
    for (H, Cin, Cout) in [(64, 160, 160), (32, 320, 320), (32, 640, 320), (64, 320, 160), (32, 160, 320)]:
        x = torch.randn(B, H, H, Cin, device=dev).bfloat16()
        dy = torch.randn(B, H, H, Cout, device=dev).bfloat16()
        ms = timeit(lambda: ops.wgrad_conv3x3(dy, x), a.iters)
        fl = 2.0 * 9 * Cin * Cout * B * H * H
        rows.append({"op": f"wgrad conv3x3 {Cin}->{Cout} @{H}x{H} B={B}", "ms": round(ms, 4), "tflops": round(fl / ms * 1e-9, 1)})
    for (M, N, K) in [(B * 4096, 768, 160), (B * 4096, 1280, 160), (B * 4096, 160, 640), (B * 4096, 160, 160),
                      (B * 1024, 320, 1280)]:
        x = torch.randn(M, K, device=dev).bfloat16()
        dy = torch.randn(M, N, device=dev).bfloat16()
        ms = timeit(lambda: ops.wgrad_linear(dy, x), a.iters)
        fl = 2.0 * M * N * K
        rows.append({"op": f"wgrad linear M={M} N={N} K={K}", "ms": round(ms, 4), "tflops": round(fl / ms * 1e-9, 1),
                     "gbs": round((M * (N + K) * 2) / ms * 1e-6, 1)})
    for r in rows:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
