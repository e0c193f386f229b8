"""Config 3 of BASELINE.json: AutoencoderKL-shape ([128,256,512,512]) decode of 64x64x4 latents to 512x512
images, batch 16, bf16, random-init weights.  Prints images/s and algorithmic TFLOP/s (38 306 GF per 16 images,
SURVEY.md 8d) and, with LAYERS=1, the per-kernel launch count of one decode."""
import json, os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_pytorch_b200 import ops
from stable_diffusion_pytorch_b200.models import AutoEncoderKL, AutoencoderConfig

B, S = int(os.environ.get("B", 16)), int(os.environ.get("S", 64))
torch.manual_seed(0)
m = AutoEncoderKL(AutoencoderConfig(autoencoder_channels_list=[128, 256, 512, 512], groups=32)).eval()
with torch.no_grad():
    for p in m.parameters():
        if p.abs().sum() == 0:
            p.normal_(0, 0.02)
m = m.to("cuda")
z = torch.randn(B, 4, S, S, device="cuda", dtype=torch.bfloat16)
with torch.no_grad():
    for _ in range(2):
        y = m.decode(z)
    torch.cuda.synchronize()
    if os.environ.get("PROFILE"):      # one decode between cudaProfilerStart/Stop for `ncu --profile-from-start off`
        torch.cuda.profiler.start()
        m.decode(z)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        sys.exit(0)
    n0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    it = 5
    e0.record()
    for _ in range(it):
        y = m.decode(z)
    e1.record()
    torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / it
gf = 38306.0 * (B / 16.0) * (S * S / 4096.0)
print(json.dumps({"metric": "vae_decode_images_per_sec", "value": B / (ms * 1e-3), "ms_per_decode": ms,
                  "batch": B, "latent": [4, S, S], "image": list(y.shape[1:]), "alg_tflops": gf / ms,
                  "launches_per_decode": (ops.launch_count() - n0) // it, "finite": bool(torch.isfinite(y.float()).all()),
                  "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30, "timeout_flag": ops.debug_timeout_flag()}))
