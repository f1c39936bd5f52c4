"""Persistent tcgen05 implicit-GEMM convolution vs an fp32 PyTorch reference."""
import os

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


def test_fork_premasks_gradient_for_the_producing_batchnorm():
    """Two bottlenecks in a row: the second block's fork multiplies its input gradient with the first block's ReLU mask
    (kernel epilogue) and the first block's bn3 backward hands dy on without masking or writing a residual gradient.
    Gradients must match the stock-op pipeline."""
    import copy
    from layerwise_b200.models import resnet as zoo
    from layerwise_b200.ops import nn as lnn
    torch.manual_seed(0)
    blocks = torch.nn.Sequential(zoo.Bottleneck(256, 64), zoo.Bottleneck(256, 64), zoo.Bottleneck(256, 64)).cuda()
    blocks = blocks.to(memory_format=torch.channels_last).to(torch.bfloat16)
    for m in blocks.modules():
        if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
            m.float()
            torch.nn.init.uniform_(m.weight, 0.5, 1.5)
            torch.nn.init.normal_(m.bias, 0.0, 0.5)
    ref = copy.deepcopy(blocks)
    x0 = torch.randn(8, 256, 12, 12, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    g = torch.randn(8, 256, 12, 12, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)

    def run(net, native):
        lnn.set_backend("auto" if native else "torch")
        try:
            x = x0.clone().requires_grad_(True)
            y1 = net[0](x)
            if native:
                assert y1.relu_mask_taken[0] is False
            y2 = net[1](y1)
            if native and not os.environ.get("LWC_NO_PREMASK"):
                assert y1.relu_mask_taken[0] is True and y2.relu_mask_taken[0] is False
            y3 = net[2](y2)
            (y3 * g).sum().backward()
            return x.grad.float(), [p.grad.float() for p in net.parameters()]
        finally:
            lnn.set_backend("auto")

    # (1) pre-masking is an exact re-association of the same arithmetic: bit-identical to the path that masks inside
    #     the batch-norm backward and writes a separate residual gradient
    gx_a, gp_a = run(blocks, True)
    os.environ["LWC_NO_PREMASK"] = "1"
    try:
        for p_ in blocks.parameters():
            p_.grad = None
        gx_c, gp_c = run(blocks, True)
    finally:
        del os.environ["LWC_NO_PREMASK"]
    assert torch.equal(gx_a, gx_c)
    for a, c in zip(gp_a, gp_c):
        assert torch.equal(a, c)
    # (2) against the stock-op pipeline: a ReLU whose bf16 pre-activation rounds to the other side of zero flips a whole
    #     gradient path, so compare in relative L2 and loosely (nine batch-norms deep, bf16)
    gx_b, gp_b = run(ref, False)

    def rel(a, b):
        return float((a - b).norm() / (b.norm() + 1e-12))
    assert rel(gx_a, gx_b) < 0.25, rel(gx_a, gx_b)
    for a, b in zip(gp_a, gp_b):
        assert rel(a, b) < 0.25, (a.shape, rel(a, b))
