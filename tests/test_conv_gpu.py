"""Persistent tcgen05 implicit-GEMM convolution vs an fp32 PyTorch reference."""
import pytest
import torch
import torch.nn.functional as F

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


def _mk(n, ci, co, h, w, k):
    torch.manual_seed(0)
    x = torch.randn(n, ci, h, w, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(co, ci, k, k, device="cuda") / (ci * k * k) ** 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    return x, wt


@pytest.mark.parametrize("n,ci,co,h,w", [(2, 64, 64, 8, 8), (3, 128, 256, 14, 14), (5, 256, 64, 56, 56), (1, 64, 192, 7, 9)])
def test_conv1x1_forward_stats_and_backward(n, ci, co, h, w):
    from layerwise_b200.ops.conv import conv2d
    x, wt = _mk(n, ci, co, h, w, 1)
    x1, w1 = x.clone().requires_grad_(True), wt.clone().requires_grad_(True)
    y, stats = conv2d(x1, w1, None, want_stats=True)
    ref_in, ref_w = x.float().requires_grad_(True), wt.float().requires_grad_(True)
    yr = F.conv2d(ref_in, ref_w)
    assert y.shape == yr.shape and y.is_contiguous(memory_format=torch.channels_last)
    assert (y.float() - yr).abs().max() <= 2e-2 * yr.abs().max()
    # epilogue statistics = column sums of the rounded output
    s = stats.sum(0)
    yf = y.float()
    assert torch.allclose(s[0], yf.sum((0, 2, 3)), rtol=1e-3, atol=1e-2)
    assert torch.allclose(s[1], (yf * yf).sum((0, 2, 3)), rtol=1e-3, atol=1e-2)
    dy = torch.randn_like(y)
    y.backward(dy)
    yr.backward(dy.float())
    assert (x1.grad.float() - ref_in.grad).abs().max() <= 2e-2 * ref_in.grad.abs().max()
    assert (w1.grad.float() - ref_w.grad).abs().max() <= 2e-2 * ref_w.grad.abs().max()


@pytest.mark.parametrize("n,ci,co,h,w", [(2, 64, 64, 8, 8), (3, 64, 128, 32, 32), (4, 128, 128, 28, 28), (3, 256, 256, 14, 14),
                                         (5, 512, 512, 7, 7), (2, 64, 64, 56, 56), (16, 64, 64, 4, 4)])
def test_conv3x3_forward_stats_and_dgrad(n, ci, co, h, w):
    from layerwise_b200.ops.conv import conv2d
    x, wt = _mk(n, ci, co, h, w, 3)
    x1, w1 = x.clone().requires_grad_(True), wt.clone().requires_grad_(True)
    y, stats = conv2d(x1, w1, None, want_stats=True)
    ref_in, ref_w = x.float().requires_grad_(True), wt.float().requires_grad_(True)
    yr = F.conv2d(ref_in, ref_w, padding=1)
    assert (y.float() - yr).abs().max() <= 2e-2 * yr.abs().max()
    s = stats.sum(0)
    yf = y.float()
    assert torch.allclose(s[0], yf.sum((0, 2, 3)), rtol=1e-3, atol=2e-2)
    assert torch.allclose(s[1], (yf * yf).sum((0, 2, 3)), rtol=1e-3, atol=2e-2)
    dy = torch.randn_like(y)
    y.backward(dy)
    yr.backward(dy.float())
    assert (x1.grad.float() - ref_in.grad).abs().max() <= 2e-2 * ref_in.grad.abs().max()
    assert (w1.grad.float() - ref_w.grad).abs().max() <= 3e-2 * ref_w.grad.abs().max()


def test_fork_conv1x1_fused_residual_gradient():
    from layerwise_b200.ops.conv import fork_conv1x1
    torch.manual_seed(0)
    conv = torch.nn.Conv2d(256, 64, 1, bias=False).cuda().to(torch.bfloat16).to(memory_format=torch.channels_last)
    x = torch.randn(4, 256, 14, 14, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    x1 = x.clone().requires_grad_(True)
    y, ident = fork_conv1x1(conv, x1)
    gy, gi = torch.randn_like(y), torch.randn_like(ident)
    torch.autograd.backward([y, ident], [gy, gi])
    gw1 = conv.weight.grad.clone()
    conv.weight.grad = None
    x2 = x.float().requires_grad_(True)
    w2 = conv.weight.detach().float().requires_grad_(True)
    y2 = F.conv2d(x2, w2)
    torch.autograd.backward([y2, x2 * 1.0], [gy.float(), gi.float()])
    assert (y.float() - y2).abs().max() <= 2e-2 * y2.abs().max()
    assert (x1.grad.float() - x2.grad).abs().max() <= 2e-2 * x2.grad.abs().max()
    assert (gw1.float() - w2.grad).abs().max() <= 3e-2 * w2.grad.abs().max()
