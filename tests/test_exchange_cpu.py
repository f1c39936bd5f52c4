"""BASELINE config #1: ResNet-9, layer-wise Top-k 1% + EF, CPU/gloo, world_size=2, synthetic 32x32."""
import pytest
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200.ops import compress_ref as cr
from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger

from _mp import run_gloo


def _small_model():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(20, 30), torch.nn.ReLU(), torch.nn.Linear(30, 5))


def _worker_exchange(rank, world, compress, method, K, V, Q, ef, steps):
    model = _small_model()
    cfg = CompressionConfig(compress=compress, method=method, ratio=K, threshold=V, qstates=Q,
                            error_feedback=ef)
    ex = GradientExchanger(model, cfg, engine="torch")
    outs = []
    for s in range(steps):
        g = torch.Generator().manual_seed(100 * s + rank)
        for p in model.parameters():
            p.grad = torch.randn(p.shape, generator=g)
        local = [p.grad.clone() for p in model.parameters()]
        ex.exchange()
        outs.append((local, [p.grad.clone() for p in model.parameters()]))
    return outs


@pytest.mark.parametrize("compress", ["layerwise", "entiremodel"])
@pytest.mark.parametrize("method,K,V", [("Topk", 0.1, None), ("Thresholdv", None, 1.0),
                                        ("AdaptiveThreshold", None, None), ("none", None, None)])
def test_deterministic_methods_two_ranks(compress, method, K, V):
    res = run_gloo(_worker_exchange, 2, (compress, method, K, V, None, False, 1))
    (loc0, out0), (loc1, out1) = res[0][0], res[1][0]
    for a, b in zip(out0, out1):
        assert torch.equal(a, b)            # replicas agree
    if compress == "layerwise":
        exp = [(cr.compress_dense(a, method, K, V) + cr.compress_dense(b, method, K, V)) / 2
               for a, b in zip(loc0, loc1)]
    else:
        fa = torch.cat([t.reshape(-1) for t in loc0])
        fb = torch.cat([t.reshape(-1) for t in loc1])
        flat = (cr.compress_dense(fa, method, K, V) + cr.compress_dense(fb, method, K, V)) / 2
        exp, off = [], 0
        for t in loc0:
            exp.append(flat[off:off + t.numel()].view_as(t))
            off += t.numel()
    for e, o in zip(exp, out0):
        assert torch.allclose(e, o, atol=1e-6)


def test_error_feedback_accumulates_across_steps():
    steps = 4
    res = run_gloo(_worker_exchange, 2, ("layerwise", "Topk", 0.05, None, None, True, steps))
    # replay rank 0 and rank 1 residuals with the oracle
    nparam = len(res[0][0][0])
    resid = [[torch.zeros(res[r][0][0][i].numel()) for i in range(nparam)] for r in range(2)]
    for s in range(steps):
        for i in range(nparam):
            cs = [cr.compress_with_feedback(res[r][s][0][i], resid[r][i], "Topk", K=0.05) for r in range(2)]
            assert torch.allclose((cs[0] + cs[1]) / 2, res[0][s][1][i], atol=1e-6)
            assert torch.equal(res[0][s][1][i], res[1][s][1][i])


def test_shared_seed_randomk_masks_coincide():
    res = run_gloo(_worker_exchange, 2, ("layerwise", "Randomk", 0.1, None, None, False, 1))
    out = res[0][0][1]
    for t in out:
        # identical masks on both ranks: exactly ceil(n*K) non-zeros survive the sum
        assert int((t != 0).sum()) == cr.randomk_keep_count(t.numel(), 0.1)


def _worker_resnet9(rank, world):
    from layerwise_b200.models import build_cifar_model
    from layerwise_b200.ops.sgd import SGD
    from layerwise_b200.train.cifar import run_batches
    torch.manual_seed(0)
    model = build_cifar_model("resnet9")
    cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.01, error_feedback=True)
    ex = GradientExchanger(model, cfg, engine="torch")
    opt = SGD(model.parameters(), lr=1e-4, momentum=0.9, weight_decay=5e-4, nesterov=True)
    g = torch.Generator().manual_seed(rank)
    batches = [{"input": torch.randn(8, 3, 32, 32, generator=g), "target": torch.randint(0, 10, (8,), generator=g)}
               for _ in range(2)]
    stats = run_batches(model, batches, True, ex, opt.step)
    return {"w": [p.detach().clone() for p in model.parameters()], "loss": stats.mean("loss"), "step": ex.step}


def test_resnet9_topk_ef_two_ranks_plumbing():
    r0, r1 = run_gloo(_worker_resnet9, 2)
    assert r0["step"] == 2
    for a, b in zip(r0["w"], r1["w"]):
        assert torch.equal(a, b)  # replicas stay in lock-step although inputs differ
    assert r0["loss"] == r0["loss"]  # finite


def test_check_health_reports_new_timeouts():
    import pytest
    import torch
    from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger
    p = torch.nn.Parameter(torch.zeros(8))
    ex = GradientExchanger([p], CompressionConfig(error_feedback=True))
    assert ex.check_health() == {} or "timeouts" in ex.check_health()       # the torch engine keeps no counters

    class _Eng:                                                            # an engine with counters
        def __init__(self):
            self.n = 0

        def stats(self):
            return {"sent": 5, "dropped": 1, "timeouts": self.n, "bytes": 40}
    ex.engine = _Eng()
    assert ex.check_health()["dropped"] == 1
    ex.engine.n = 2
    with pytest.raises(RuntimeError, match="2 peer flag wait"):
        ex.check_health()
    assert ex.check_health()["timeouts"] == 2                              # already reported: no second raise
    ex.engine.n = 3
    assert ex.check_health(raise_on_timeout=False)["timeouts"] == 3
    # without error feedback a clamped entry would be lost for good: that is an error, not a statistic
    ex2 = GradientExchanger([p], CompressionConfig(error_feedback=False))
    ex2.engine = _Eng()
    with pytest.raises(RuntimeError, match="did not fit"):
        ex2.check_health()
