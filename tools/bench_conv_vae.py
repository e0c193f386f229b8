"""The AutoencoderKL decoder's 3x3 convolutions (config 3: [128, 256, 512, 512], batch 16) timed one by one:
ms and executed TFLOP/s per shape.  SDB_CONV_HALO=0 / SDB_CONV_PAIR=0 switch the loader variants for A/B runs."""
import math, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_pytorch_b200 import ops, _pack

SHAPES = [  # B, H, W, Cin, Cout, up2
    (16, 512, 512, 128, 128, False), (16, 256, 256, 256, 256, False), (16, 256, 256, 256, 128, False),
    (16, 128, 128, 512, 512, False), (16, 128, 128, 512, 256, False), (16, 64, 64, 512, 512, False),
    (16, 256, 256, 256, 256, True), (16, 128, 128, 512, 512, True), (16, 64, 64, 512, 512, True)]
only = os.environ.get("ONLY")
for B, H, W, Cin, Cout, up2 in SHAPES:
    name = f"conv {Cin}->{Cout} @{H}{' up2' if up2 else ''}"
    if only and only not in name:
        continue
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, H, W, Cin, generator=g).to("cuda", torch.bfloat16)
    bias = torch.randn(Cout, generator=g).cuda()
    if up2:
        w = _pack.conv3x3_up_w((torch.randn(Cout, Cin, 3, 3, generator=g) / math.sqrt(9 * Cin)).cuda())
    else:
        w = (torch.randn(Cout, 3, 3, Cin, generator=g) / math.sqrt(9 * Cin)).to("cuda", torch.bfloat16)
    out = torch.empty(B, H * (2 if up2 else 1), W * (2 if up2 else 1), Cout, device="cuda", dtype=torch.bfloat16)
    for _ in range(2):
        ops.conv3x3(x, w, bias, up2=up2, out=out)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    it = 5
    torch.cuda.synchronize()
    e0.record()
    for _ in range(it):
        ops.conv3x3(x, w, bias, up2=up2, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / it
    taps = 16 if up2 else 9          # executed: four phases x 2x2 taps
    gf = 2.0 * B * H * W * taps * Cin * Cout / 1e9
    print(f"{name:28s} {ms:8.3f} ms  {gf / ms:8.1f} TFLOP/s executed")
print("timeout flag", ops.debug_timeout_flag())
