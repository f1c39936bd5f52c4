"""Small model ops of csrc/misc.cu vs plain PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("shape", [(4, 64, 8, 8), (3, 1000), (5, 13)])
def test_relu_and_scale_forward_backward(dtype, shape):
    from layerwise_b200.ops import misc
    torch.manual_seed(0)
    x = torch.randn(*shape, device="cuda").to(dtype)
    if len(shape) == 4:
        x = x.contiguous(memory_format=torch.channels_last)
    x1 = x.clone().requires_grad_(True)
    y = misc.relu(x1)
    g = torch.randn_like(y)
    y.backward(g)
    assert torch.equal(y, torch.relu(x))
    assert torch.equal(x1.grad, g * (x > 0))
    x2 = x.clone().requires_grad_(True)
    z = misc.scale(x2, 0.125)
    z.backward(g)
    assert torch.equal(z, x * 0.125)                 # a power of two: exact in every dtype
    assert torch.equal(x2.grad, g * 0.125)


@pytest.mark.parametrize("n", [4096 * 33, 1000, 7])
def test_dropout_statistics_mask_and_backward(n):
    from layerwise_b200.ops import misc
    torch.manual_seed(3)
    x = torch.ones(n, device="cuda", dtype=torch.bfloat16).requires_grad_(True)
    y = misc.dropout(x, 0.5, True)
    kept = y != 0
    assert torch.all((y == 0) | (y == 2))
    if n > 1000:
        assert abs(float(kept.float().mean()) - 0.5) < 0.01
    g = torch.full_like(y, 3.0)
    y.backward(g)
    assert torch.equal(x.grad, kept.to(g.dtype) * 6.0)       # same mask, same 1 / (1 - p) scale
    y2 = misc.dropout(x.detach(), 0.5, True)
    if n > 1000:
        assert not torch.equal(y2 != 0, kept)                # a new call draws new bits
    assert misc.dropout(x, 0.5, False) is x                  # eval mode: identity


@pytest.mark.parametrize("n,c,ih,iw,oh,ow", [(4, 2048, 7, 7, 1, 1), (3, 512, 1, 1, 7, 7), (2, 64, 9, 5, 4, 3), (2, 128, 8, 8, 8, 8)])
def test_adaptive_avg_pool_matches_torch(n, c, ih, iw, oh, ow):
    from layerwise_b200.ops import misc
    torch.manual_seed(0)
    x = torch.randn(n, c, ih, iw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    x1 = x.clone().requires_grad_(True)
    y = misc.adaptive_avg_pool2d(x1, (oh, ow))
    xr = x.float().requires_grad_(True)
    yr = F.adaptive_avg_pool2d(xr, (oh, ow))
    assert y.shape == yr.shape
    assert torch.allclose(y.float(), yr, rtol=1e-2, atol=1e-2)
    g = torch.randn_like(y)
    y.backward(g)
    yr.backward(g.float())
    assert torch.allclose(x1.grad.float(), xr.grad, rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("n,c", [(64, 1000), (33, 10), (7, 5)])
def test_correct_and_topk_counters(n, c):
    from layerwise_b200.ops import misc
    torch.manual_seed(0)
    logits = torch.randn(n, c, device="cuda").to(torch.bfloat16)       # bf16: plenty of exact ties
    target = torch.randint(0, c, (n,), device="cuda")
    got = misc.correct(logits, target)
    assert got.dtype == torch.bool and torch.equal(got, logits.max(dim=1)[1] == target)
    cnt = misc.topk_hits(logits, target)
    cnt = misc.topk_hits(logits, target, cnt)                          # accumulates
    pred = logits.float().topk(min(5, c), 1)[1]
    hit = pred == target.view(-1, 1)
    # ties: torch.topk may order equal values differently, so compare top-1 exactly via max() and top-5 by value rank
    top1 = int((logits.max(dim=1)[1] == target).sum())
    xt = logits.float().gather(1, target.view(-1, 1))
    rank = ((logits.float() > xt) | ((logits.float() == xt) & (torch.arange(c, device="cuda").view(1, -1) < target.view(-1, 1)))).sum(1)
    top5 = int((rank < 5).sum())
    assert cnt.tolist() == [2.0 * top1, 2.0 * top5]
    assert abs(int(hit.sum()) - top5) <= n // 8 + 1                    # sanity against torch.topk up to tie order


def test_models_route_small_ops_to_native_kernels():
    from layerwise_b200 import _native
    from layerwise_b200.models import alexnet as am
    torch.manual_seed(0)
    net = am.AlexNet().cuda().to(torch.bfloat16).to(memory_format=torch.channels_last)
    batch = {"input": torch.randn(16, 3, 32, 32, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last),
             "target": torch.randint(0, 10, (16,), device="cuda")}
    _native.counters.reset()
    out = net(batch)
    out["loss"].sum().backward()
    assert out["correct"].dtype == torch.bool and torch.isfinite(out["loss"]).all()
    assert _native.counters.kernels >= 5 + 2 + 2 + 1          # 5 convs, 2 dropouts, ReLUs, correct


def test_lagged_reader_on_device_scalars():
    from layerwise_b200.utils import LaggedReader
    r = LaggedReader(depth=1)
    got = []
    for i in range(6):
        v = r.push(torch.full((), float(i), device="cuda") * 2)
        if v is not None:
            got.append(v)
    got += r.drain()
    assert got == [0.0, 2.0, 4.0, 6.0, 8.0, 10.0]
    assert all(s.is_pinned() for s in r._free)
