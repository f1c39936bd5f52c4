import pytest
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200.ops.sgd import FusedSGD, allow_lowp_grad

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("momentum,nesterov,wd,damp", [(0.0, False, 0.0, 0.0), (0.9, True, 5e-4, 0.0),
                                                       (0.9, False, 1e-4, 0.1)])
def test_fused_sgd_matches_torch(momentum, nesterov, wd, damp):
    torch.manual_seed(0)
    shapes = [(64,), (130, 33), (8192,), (8193,), (3, 3, 64, 128)]
    a = [torch.randn(s, device="cuda", requires_grad=True) for s in shapes]
    b = [t.detach().clone().requires_grad_(True) for t in a]
    oa = FusedSGD(a, lr=0.1, momentum=momentum, nesterov=nesterov, weight_decay=wd, dampening=damp)
    ob = torch.optim.SGD(b, lr=0.1, momentum=momentum, nesterov=nesterov, weight_decay=wd, dampening=damp)
    for step in range(3):
        for x, y in zip(a, b):
            g = torch.randn_like(x)
            x.grad = g.clone()
            y.grad = g.clone()
        oa.step()
        ob.step()
    for x, y in zip(a, b):
        assert torch.allclose(x, y, atol=1e-5, rtol=1e-5)
    assert oa.state_dict()["param_groups"][0]["momentum"] == momentum


def test_fused_sgd_bf16_model_copy_and_grads():
    torch.manual_seed(0)
    master = [allow_lowp_grad(torch.randn(5000, device="cuda", requires_grad=True))]
    model = [master[0].detach().to(torch.bfloat16)]
    opt = FusedSGD(master, lr=0.05, momentum=0.9, nesterov=True, weight_decay=1e-4, model_params=model,
                   grad_scale=1.0 / 128)
    ref = master[0].detach().clone()
    g = torch.randn(5000, device="cuda").to(torch.bfloat16)
    master[0].grad = g.clone()
    opt.step()
    d = g.float() / 128 + 1e-4 * ref
    exp = ref - 0.05 * (d + 0.9 * d)
    assert torch.allclose(master[0], exp, atol=1e-5)
    assert torch.equal(model[0], master[0].detach().to(torch.bfloat16))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_fused_sgd_low_precision_weights(dtype):
    """16-bit weights without fp32 masters (dawn.py --dtype bf16): fp32 arithmetic, one rounding per step."""
    from layerwise_b200 import _native
    torch.manual_seed(0)
    shapes = [(64,), (130, 33), (8193,), (3, 3, 64, 128)]
    a = [torch.randn(s, device="cuda").to(dtype).requires_grad_(True) for s in shapes]
    w = [t.detach().float() for t in a]                    # fp32 emulation of the same update
    m = [None] * len(a)
    opt = FusedSGD(a, lr=0.1, momentum=0.9, nesterov=True, weight_decay=5e-4)
    n0 = _native.counters.kernels
    for step in range(3):
        for i, x in enumerate(a):
            g = torch.randn_like(x)
            x.grad = g.clone()
            d = g.float() + 5e-4 * w[i]
            m[i] = d if m[i] is None else 0.9 * m[i] + d
            m[i] = m[i].to(dtype).float()                  # the momentum buffer is stored in `dtype`
            w[i] = (w[i] - 0.1 * (d + 0.9 * m[i])).to(dtype).float()
        opt.step()
    assert _native.counters.kernels - n0 == 3              # one fused launch per step, no per-tensor fallback
    tol = 2e-2 if dtype == torch.bfloat16 else 3e-3
    for x, y in zip(a, w):
        assert x.dtype == dtype
        assert torch.allclose(x.float(), y, atol=tol, rtol=tol)
    assert opt.state[a[0]]["momentum_buffer"].dtype == dtype
