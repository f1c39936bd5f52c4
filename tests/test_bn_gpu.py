"""Fused batch-norm(+add)(+ReLU) kernels vs an fp32 PyTorch reference."""
import pytest
import torch
import torch.nn.functional as F

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


def _reference(x, w, b, res, relu, eps=1e-5):
    xf = x.float()
    mean = xf.mean((0, 2, 3), keepdim=True)
    var = xf.var((0, 2, 3), unbiased=False, keepdim=True)
    y = (xf - mean) * torch.rsqrt(var + eps) * w.view(1, -1, 1, 1) + b.view(1, -1, 1, 1)
    if res is not None:
        y = y + res.float()
    return F.relu(y) if relu else y


@pytest.mark.parametrize("dtype,tol", [(torch.bfloat16, 3e-2), (torch.float32, 2e-4)])
@pytest.mark.parametrize("shape", [(8, 64, 28, 28), (4, 2048, 7, 7), (3, 192, 5, 9), (16, 256, 14, 14)])
@pytest.mark.parametrize("relu,use_res", [(False, False), (True, False), (True, True)])
def test_bn_forward_backward(dtype, tol, shape, relu, use_res):
    from layerwise_b200.ops.bn import batch_norm_act
    torch.manual_seed(0)
    N, Cc, H, W = shape
    x = (torch.randn(shape, device="cuda") * 2 + 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    res = torch.randn(shape, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last) if use_res else None
    w = (torch.rand(Cc, device="cuda") + 0.5)
    b = torch.randn(Cc, device="cuda") * 0.1
    rm, rv = torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")
    x1 = x.clone().requires_grad_(True)
    w1, b1 = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    r1 = res.clone().requires_grad_(True) if use_res else None
    y = batch_norm_act(x1, w1, b1, rm, rv, r1, True, 0.1, 1e-5, relu)
    dy = torch.randn_like(y)
    y.backward(dy)

    x2 = x.detach().float().requires_grad_(True)
    w2, b2 = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    r2 = res.detach().float().requires_grad_(True) if use_res else None
    yr = _reference(x2, w2, b2, r2, relu)
    yr.backward(dy.float())

    def close(a, e, scale=1.0):
        return (a.float() - e).abs().max() <= tol * scale * max(1.0, float(e.abs().max()))
    assert close(y, yr)
    assert close(x1.grad, x2.grad)
    assert close(w1.grad, w2.grad, 4.0) and close(b1.grad, b2.grad, 4.0)
    if use_res:
        assert close(r1.grad, r2.grad)
    # running statistics follow torch's convention (unbiased variance)
    xf = x.float()
    assert torch.allclose(rm, 0.1 * xf.mean((0, 2, 3)), atol=1e-3)
    assert torch.allclose(rv, 0.9 + 0.1 * xf.var((0, 2, 3), unbiased=True), rtol=1e-3, atol=1e-3)


def test_bn_eval_uses_running_stats():
    from layerwise_b200.ops import nn as lnn
    bn = lnn.BatchNorm2d(64).cuda()
    bn.running_mean.normal_()
    bn.running_var.uniform_(0.5, 2.0)
    bn.eval()
    x = torch.randn(4, 64, 9, 9, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = bn(x, relu=True)
    ref = F.relu(F.batch_norm(x.float(), bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.1, 1e-5))
    assert (y.float() - ref).abs().max() < 3e-2


def test_module_matches_torch_module_in_training():
    from layerwise_b200 import _native
    from layerwise_b200.ops import nn as lnn
    torch.manual_seed(1)
    a, b = lnn.BatchNorm2d(128).cuda(), torch.nn.BatchNorm2d(128).cuda()
    x = torch.randn(8, 128, 10, 10, device="cuda").contiguous(memory_format=torch.channels_last)
    _native.counters.reset()
    ya, yb = a(x), b(x)
    assert _native.counters.kernels == 3
    assert torch.allclose(ya, yb, atol=1e-4)
    assert torch.allclose(a.running_var, b.running_var, atol=1e-5) and int(a.num_batches_tracked) == 1


def test_num_batches_tracked_is_bumped_by_the_finalize_kernel():
    from layerwise_b200.ops import nn as lnn
    bn = lnn.BatchNorm2d(64).cuda()
    x = torch.randn(8, 64, 6, 6, device="cuda").contiguous(memory_format=torch.channels_last)
    for _ in range(3):
        bn(x, relu=True)
    bn.eval()
    bn(x)
    torch.cuda.synchronize()
    assert int(bn.num_batches_tracked) == 3


def test_relu_bitmask_replaces_saved_output():
    """With a fused residual the backward reads a 1-bit-per-element mask written by the forward, not y."""
    from layerwise_b200.ops import bn as bnops
    torch.manual_seed(0)
    x = torch.randn(6, 64, 9, 7, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    r = torch.randn_like(x).requires_grad_(True)
    w = torch.rand(64, device="cuda").add_(0.5).requires_grad_(True)
    b = torch.randn(64, device="cuda").requires_grad_(True)
    y = bnops.batch_norm_act(x, w, b, torch.zeros(64, device="cuda"), torch.ones(64, device="cuda"), r, True, 0.1, 1e-5, True)
    saved = y.grad_fn.saved_tensors
    assert saved[1] is None and saved[6] is not None and saved[6].dtype == torch.uint8      # no y, a mask
    mask = saved[6].view(6 * 9 * 7, 8)
    bits = torch.stack([(mask >> k) & 1 for k in range(8)], dim=-1).reshape(6, 9, 7, 64).permute(0, 3, 1, 2).bool()
    assert torch.equal(bits, y > 0)
    dy = torch.randn_like(y)
    y.backward(dy)
    assert torch.equal(r.grad, dy * (y > 0))
