"""tcgen05 weight-gradient kernel (csrc/wgrad_umma.cu) against fp64 torch on the same bf16 operands."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module", autouse=True)
def _gpu():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from stable_diffusion_pytorch_b200 import ops
    yield
    assert ops.debug_timeout_flag() == 0


def rel(a, b):
    return float((a.double().cpu() - b.double().cpu()).norm() / b.double().norm())


@pytest.mark.parametrize("M,N,K", [(64, 128, 64), (1000, 160, 160), (4096 + 37, 1280, 160), (300, 96, 640),
                                   (77 * 4, 512, 768), (2048, 160, 1280), (130, 776, 128)])
def test_wgrad_linear(M, N, K):
    from stable_diffusion_pytorch_b200 import ops
    g = torch.Generator().manual_seed(M + N + K)
    dy = torch.randn(M, N, generator=g).bfloat16()
    x = torch.randn(M, K, generator=g).bfloat16()
    ref = dy.double().T @ x.double()
    out = ops.wgrad_linear(dy.to(DEV), x.to(DEV))
    assert out is not None and tuple(out.shape) == (N, K)
    assert rel(out, ref) < 1e-5


def test_wgrad_linear_strided_views_and_accumulate():
    from stable_diffusion_pytorch_b200 import ops
    g = torch.Generator().manual_seed(9)
    M = 777
    big = torch.randn(M, 3 * 256, generator=g).bfloat16().to(DEV)       # fused QKV output: three strided views
    x = torch.randn(M, 160, generator=g).bfloat16().to(DEV)
    dw = torch.zeros(256, 160 + 64, device=DEV)
    ops.wgrad_linear(big[:, 256:512], x, out=dw[:, :160])
    x2 = torch.randn(M, 64, generator=g).bfloat16().to(DEV)
    ops.wgrad_linear(big[:, 256:512], x2, out=dw[:, 160:])
    ref = big[:, 256:512].double().T @ torch.cat([x, x2], 1).double()
    assert rel(dw, ref) < 1e-5
    assert ops.wgrad_linear(big[:, :256], torch.zeros(M, 20, device=DEV, dtype=torch.bfloat16)) is None   # K % 64


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(2, 64, 64, 160, 160), (3, 32, 32, 320, 160), (5, 8, 8, 640, 320),
                                            (7, 4, 4, 320, 320), (2, 16, 16, 480, 320), (1, 8, 8, 128, 16),
                                            (3, 2, 2, 64, 72)])
def test_wgrad_conv3x3(B, H, W, Cin, Cout):
    from stable_diffusion_pytorch_b200 import ops
    g = torch.Generator().manual_seed(B * H + Cin)
    x = torch.randn(B, Cin, H, W, generator=g).bfloat16()
    dy = torch.randn(B, Cout, H, W, generator=g).bfloat16()
    xd = x.double().requires_grad_(False)
    w = torch.zeros(Cout, Cin, 3, 3, dtype=torch.float64, requires_grad=True)
    y = torch.nn.functional.conv2d(xd, w, padding=1)
    (gw,) = torch.autograd.grad(y, w, dy.double())
    out = ops.wgrad_conv3x3(dy.permute(0, 2, 3, 1).contiguous().to(DEV), x.permute(0, 2, 3, 1).contiguous().to(DEV))
    assert out is not None
    assert rel(out.permute(0, 3, 1, 2), gw) < 1e-5          # [Cout,3,3,Cin] -> OIHW
