import pytest
import torch
import torch.nn.functional as F

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("reduction", ["none", "mean", "sum"])
@pytest.mark.parametrize("N,C", [(7, 10), (224, 1000), (33, 4097)])
def test_cross_entropy_matches_torch(dtype, tol, reduction, N, C):
    from layerwise_b200 import _native
    from layerwise_b200.ops import nn as lnn
    torch.manual_seed(0)
    x = (torch.randn(N, C, device="cuda") * 3).to(dtype)
    t = torch.randint(0, C, (N,), device="cuda")
    a = x.clone().requires_grad_(True)
    b = x.detach().float().requires_grad_(True)
    _native.counters.reset()
    la = lnn.CrossEntropyLoss(reduction=reduction)(a, t)
    assert _native.counters.kernels == 1
    lb = F.cross_entropy(b, t, reduction=reduction)
    g = torch.randn_like(lb)
    la.backward(g.to(la.dtype))
    lb.backward(g)
    assert torch.allclose(la.float(), lb, atol=tol * 10, rtol=tol)
    assert (a.grad.float() - b.grad).abs().max() <= tol * max(1.0, float(b.grad.abs().max())) * 2


def test_mean_reduction_divides_by_the_number_of_valid_targets():
    """ignore_index targets get zero loss and gradient and do not count in the mean (torch semantics)."""
    import torch
    from layerwise_b200.ops import nn as lnn
    torch.manual_seed(0)
    logits = torch.randn(37, 11, device="cuda").to(torch.bfloat16)
    target = torch.randint(0, 11, (37,), device="cuda")
    target[::5] = -100
    crit = lnn.CrossEntropyLoss().cuda()
    a = logits.clone().requires_grad_(True)
    la = crit(a, target)
    la.backward()
    b = logits.float().requires_grad_(True)
    lb = torch.nn.functional.cross_entropy(b, target)
    lb.backward()
    assert torch.allclose(la.float(), lb, rtol=2e-2, atol=2e-2)
    assert torch.allclose(a.grad.float(), b.grad, rtol=3e-2, atol=2e-3)
    assert float(a.grad[0].abs().sum()) == 0.0
