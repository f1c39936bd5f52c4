"""tcgen05 GEMM vs an fp32 PyTorch reference of the same op."""
import pytest
import torch

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    return A @ B.t()


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 128, 256), (304, 200, 136), (1000, 10 * 8, 512),
                                   (4096, 256, 1024), (64, 64, 8192), (512, 512, 256), (712, 520, 192), (1024, 2048, 64)])
def test_gemm_matches_fp32(a_mn, b_mn, M, N, K):
    from layerwise_b200.ops.gemm import gemm
    torch.manual_seed(0)
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(torch.bfloat16)
    ref = _ref(a, b, a_mn, b_mn)
    out = gemm(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32)
    torch.cuda.synchronize()
    err = (out - ref).abs().max() / ref.abs().max()
    assert err < 2e-3, float(err)
    out16 = gemm(a, b, a_mn=a_mn, b_mn=b_mn)
    assert (out16.float() - ref).abs().max() / ref.abs().max() < 1e-2


@pytest.mark.parametrize("splits", [1, 2, 7])
def test_gemm_splitk_bias_relu_accumulate(splits):
    from layerwise_b200.ops.gemm import gemm
    torch.manual_seed(1)
    M, N, K = 200, 96, 1024
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    b = torch.randn(N, K, device="cuda").to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    base = torch.randn(M, N, device="cuda")
    out = base.clone()
    gemm(a, b, out=out, bias=bias, relu=True, accumulate=True, splits=splits)
    ref = torch.relu(a.float() @ b.float().t() + bias + base)
    assert (out - ref).abs().max() / ref.abs().max() < 2e-3


def test_linear_autograd_matches_torch():
    from layerwise_b200.ops.gemm import linear
    torch.manual_seed(2)
    x = torch.randn(512, 1024, device="cuda").to(torch.bfloat16).requires_grad_(True)
    w = (torch.randn(4096, 1024, device="cuda") * 0.03).to(torch.bfloat16).requires_grad_(True)
    b = torch.randn(4096, device="cuda").to(torch.bfloat16).requires_grad_(True)
    y = linear(x, w, b)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr, wr, br = [t.detach().float().requires_grad_(True) for t in (x, w, b)]
    yr = torch.nn.functional.linear(xr, wr, br)
    yr.backward(dy.float())
    for got, exp in ((y, yr), (x.grad, xr.grad), (w.grad, wr.grad), (b.grad, br.grad)):
        assert (got.float() - exp).abs().max() / exp.abs().max() < 1.5e-2


def test_module_dispatch_uses_native_linear():
    from layerwise_b200 import _native
    from layerwise_b200.ops import nn as lnn
    lin = lnn.Linear(256, 512).cuda().to(torch.bfloat16)
    x = torch.randn(64, 256, device="cuda").to(torch.bfloat16)
    _native.counters.reset()
    y = lin(x)
    assert _native.counters.kernels >= 1
    ref = torch.nn.functional.linear(x.float(), lin.weight.float(), lin.bias.float())
    assert (y.float() - ref).abs().max() / ref.abs().max() < 1e-2
