"""DDP wrappers over gloo (world_size 2): dense averaging and Random-k + error feedback."""
import torch

import layerwise_b200  # noqa: F401

from _mp import run_gloo


def _net():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(12, 16), torch.nn.BatchNorm1d(16), torch.nn.ReLU(), torch.nn.Linear(16, 4))


def _worker_dense(rank, world):
    from layerwise_b200.parallel.ddp import DistributedDataParallel
    net = _net()
    with torch.no_grad():                       # perturb rank 1: the wrapper must re-broadcast rank 0's weights
        if rank == 1:
            for p in net.parameters():
                p.add_(1.0)
    ddp = DistributedDataParallel(net, check_reduction=True)
    g = torch.Generator().manual_seed(10 + rank)
    x = torch.randn(8, 12, generator=g)
    ddp(x).sum().backward()
    ddp.finish_reduction()
    return {"w": [p.detach().clone() for p in net.parameters()], "g": [p.grad.clone() for p in net.parameters()], "x": x}


def test_dense_ddp_broadcasts_and_averages():
    r0, r1 = run_gloo(_worker_dense, 2)
    for a, b in zip(r0["w"], r1["w"]):
        assert torch.equal(a, b)
    for a, b in zip(r0["g"], r1["g"]):
        assert torch.equal(a, b)
    # oracle: mean of the two local gradients
    grads = []
    for x in (r0["x"], r1["x"]):
        net = _net()
        net(x).sum().backward()
        grads.append([p.grad.clone() for p in net.parameters()])
    for a, b, got in zip(grads[0], grads[1], r0["g"]):
        assert torch.allclose((a + b) / 2, got, atol=1e-6)


def _worker_randk(rank, world):
    from layerwise_b200.ops import compress_ref as cr
    from layerwise_b200.parallel.ddp import RandomKSparsifiedDDP
    net = _net()
    ddp = RandomKSparsifiedDDP(net, randk=0.25, seed=1234)
    total = [torch.zeros(p.numel()) for p in net.parameters()]
    sent = [torch.zeros(p.numel()) for p in net.parameters()]
    nz = []
    for step in range(3):
        g = torch.Generator().manual_seed(100 * step + rank)
        x = torch.randn(8, 12, generator=g)
        net.zero_grad()
        ddp(x).sum().backward()
        ddp.finish_reduction()
        nz.append([int((p.grad != 0).sum()) for p in net.parameters()])
    eps = [e.clone() for e in ddp.epsilon]
    return {"nz": nz, "eps": eps, "keep": [cr.randomk_keep_count(p.numel(), 0.25) for p in net.parameters()],
            "g": [p.grad.clone() for p in net.parameters()]}


def test_randomk_sparsified_ddp_shared_masks_and_residuals():
    r0, r1 = run_gloo(_worker_randk, 2)
    for step_nz in r0["nz"]:
        for n, k in zip(step_nz, r0["keep"]):
            assert n <= k                      # identical masks on both ranks: at most ceil(n*randk) survivors
    for a, b in zip(r0["g"], r1["g"]):
        assert torch.equal(a, b)
    assert any(float(e.abs().sum()) > 0 for e in r0["eps"])     # error feedback holds the unsent part
