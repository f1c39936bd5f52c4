"""Oracle compressors: semantics from SURVEY 2.2 and parity with the reference's own code."""
import math
import os
import sys

import pytest
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200.ops import compress_ref as cr

REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")


@pytest.mark.parametrize("n,K,kept", [(1000, 0.01, 11), (1000, 0.1, 101), (1000, 0.5, 501), (64, 0.01, 1)])
def test_topk_keep_count(n, K, kept):
    g = torch.randn(n)
    c = cr.compress_dense(g, "Topk", K=K)
    assert int((c != 0).sum()) == kept == cr.topk_keep_count(n, K)
    # survivors are the largest magnitudes and are unchanged
    idx = g.abs().topk(kept).indices
    assert torch.equal(c[idx], g[idx])


def test_topk_ties_and_zeros():
    g = torch.zeros(100)
    g[:5] = torch.tensor([1.0, -2.0, 3.0, -4.0, 5.0])
    c = cr.compress_dense(g, "Topk", K=0.1)  # rank 90 falls among the zeros: threshold 0, all kept
    assert torch.equal(c, g)
    g = torch.tensor([1.0, 1.0, 1.0, 1.0, 0.5, 0.25])
    c = cr.compress_dense(g, "Topk", K=0.5)   # rank ceil(6*.5)=3 -> thres 1.0, four ties kept
    assert int((c != 0).sum()) == 4
    assert torch.equal(cr.compress_dense(g, "Topk", K=1.0), g)  # no exception at K=1


@pytest.mark.parametrize("n,K,kept", [(1000, 0.01, 10), (1000, 0.015, 15), (999, 0.5, 500)])
def test_randomk_exact_count(n, K, kept):
    g = torch.randn(n) + 3.0
    c = cr.compress_dense(g, "Randomk", K=K)
    assert int((c != 0).sum()) == kept == cr.randomk_keep_count(n, K)
    nz = c != 0
    assert torch.equal(c[nz], g[nz])  # no rescale


def test_randomk_injected_keys():
    g = torch.arange(1.0, 11.0)
    keys = torch.tensor([5, 9, 1, 9, 0, 3, 7, 2, 8, 4])
    c = cr.compress_dense(g, "Randomk", K=0.3, rand=keys)
    assert c.nonzero().view(-1).tolist() == [1, 3, 8]


def test_thresholds():
    g = torch.tensor([0.5, -0.125, 0.0625, -2.0, 0.25])
    assert cr.compress_dense(g, "Thresholdv", V=0.125).tolist() == [0.5, -0.125, 0.0, -2.0, 0.25]
    assert cr.compress_dense(g, "AdaptiveThreshold").tolist() == [0.0, 0.0, 0.0, -2.0, 0.0]
    g2 = torch.tensor([1.0, -2.0, 0.99])
    assert cr.compress_dense(g2, "AdaptiveThreshold").tolist() == [1.0, -2.0, 0.0]


def test_terngrad_levels_and_unbiased():
    torch.manual_seed(0)
    g = torch.randn(2000)
    s = g.abs().max()
    c = cr.compress_dense(g, "TernGrad")
    vals = set(torch.unique(c).tolist())
    assert vals <= {0.0, float(s), float(-s)}
    acc = torch.zeros_like(g)
    reps = 400
    for _ in range(reps):
        acc += cr.compress_dense(g, "TernGrad")
    assert (acc / reps - g).abs().mean() < 0.12
    assert torch.equal(cr.compress_dense(torch.zeros(8), "TernGrad"), torch.zeros(8))


def test_dithering_unbiased_levels_and_zero_guard():
    torch.manual_seed(1)
    g = torch.randn(512)
    q = 15
    c = cr.compress_dense(g, "RandomDithering", qstates=q)
    lv = (c.abs() / g.norm() * q)
    assert torch.allclose(lv, lv.round(), atol=1e-4)
    acc = torch.zeros_like(g)
    for _ in range(300):
        acc += cr.compress_dense(g, "QSGD", qstates=q)   # alias (D8)
    assert (acc / 300 - g).abs().mean() < 0.12
    z = cr.compress_dense(torch.zeros(16), "RandomDithering", qstates=255)
    assert torch.equal(z, torch.zeros(16))
    znan = cr.compress_dense(torch.zeros(16), "RandomDithering", qstates=255, zero_guard=False)
    assert torch.isnan(znan).all()   # reference behaviour (D9)


def test_identity_fallthroughs():
    g = torch.randn(10)
    for m, kw in (("Topk", dict(K=0)), ("Randomk", dict(K=None)), ("Thresholdv", dict(V=0)),
                  ("RandomDithering", dict(qstates=0)), ("bogus", {}), (None, {})):
        assert torch.equal(cr.compress_dense(g, m, **kw), g)
    assert cr.canonical_method("QSGD") == cr.DITHERING


def test_error_feedback_conserves_mass():
    torch.manual_seed(2)
    g = torch.randn(1000)
    res = torch.zeros(1000)
    sent = torch.zeros(1000)
    total = torch.zeros(1000)
    for _ in range(5):
        total += g
        sent += cr.compress_with_feedback(g, res, "Topk", K=0.05)
    assert torch.allclose(sent + res, total, atol=1e-5)


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "CIFAR10")), reason="reference tree not installed")
@pytest.mark.parametrize("method,kw", [("Topk", dict(K=0.01)), ("Topk", dict(K=0.3)),
                                       ("Thresholdv", dict(V=0.5)), ("AdaptiveThreshold", {}),
                                       ("none", {})])
def test_matches_reference_function(method, kw):
    """Deterministic methods must equal the reference's layerwise function bit for bit (W=1, gloo)."""
    import torch.distributed as dist
    sys.path.insert(0, REF)
    try:
        from CIFAR10.core import layerwise_compressed_comm as ref_fn
    finally:
        sys.path.remove(REF)
    if not dist.is_initialized():
        dist.init_process_group("gloo", init_method="tcp://127.0.0.1:29541", world_size=1, rank=0)
    torch.manual_seed(3)
    lin = torch.nn.Linear(40, 50)
    for p in lin.parameters():
        p.grad = torch.randn_like(p)
    mine = [cr.compress_dense(p.grad.clone(), method, K=kw.get("K"), V=kw.get("V")) for p in lin.parameters()]
    ref_fn(lin, 1, method, kw.get("K"), kw.get("V"), None)
    for a, p in zip(mine, lin.parameters()):
        assert torch.equal(a, p.grad)
