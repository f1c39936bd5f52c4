"""End-to-end training sanity on the GPU: the fused pipeline must actually learn, and track the stock-op pipeline."""
import pytest
import torch

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


def _run(native: bool, steps: int, method: str, ratio: float, lr: float = 0.02, momentum: float = 0.9):
    from layerwise_b200.ops import nn as lnn
    from layerwise_b200.parallel.exchange import CompressionConfig
    from layerwise_b200.train.imagenet import Trainer
    lnn.set_backend("auto" if native else "torch")
    try:
        torch.manual_seed(0)
        cfg = CompressionConfig(compress="layerwise", method=method, ratio=ratio, error_feedback=True)
        tr = Trainer(arch="resnet50", num_classes=16, dtype="bf16", cfg=cfg, lr=lr, momentum=momentum, weight_decay=1e-4,
                     device="cuda:0", overlap=native, engine="auto" if native else "torch", init_bn0=False)
        g = torch.Generator().manual_seed(1)
        x = torch.randint(0, 256, (32, 3, 64, 64), dtype=torch.uint8, generator=g).cuda()
        y = torch.randint(0, 16, (32,), generator=g).cuda()
        losses = []
        for _ in range(steps):
            _, loss = tr.train_step(x, y)
            losses.append(float(loss))
        return losses
    finally:
        lnn.set_backend("auto")


def test_fused_pipeline_memorises_a_batch():
    # gentle learning rate: at 0.02 + momentum this tiny problem is chaotic (the loss spikes before it recovers) and
    # whether step 60 sits below a threshold depends on which algorithms cuDNN's autotuner picked in this process
    losses = _run(True, 60, "Topk", 0.1, lr=0.005)
    assert all(l == l for l in losses)                       # finite
    assert min(losses[-10:]) < 0.8 * losses[0], losses[::8]


def test_fused_pipeline_tracks_stock_pipeline():
    # a gentle learning rate keeps the problem well conditioned (at lr 0.02 + momentum even the stock pipeline's
    # loss spikes to ~36 before recovering, and any rounding difference is amplified chaotically)
    a = _run(True, 12, "Topk", 0.25, lr=0.003, momentum=0.0)
    b = _run(False, 12, "Topk", 0.25, lr=0.003, momentum=0.0)
    # same data, same init, same compressor: the two loss curves must stay close (bf16 + different kernels)
    assert abs(a[0] - b[0]) < 0.05 * abs(b[0]) + 0.05
    assert b[-1] < b[0] and a[-1] < a[0]
    assert abs(a[-1] - b[-1]) < 0.15 * abs(b[-1]) + 0.1, (a, b)
