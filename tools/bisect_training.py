#!/usr/bin/env python
"""Loss trajectories of the same tiny training problem with native components switched on one at a time."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import conv as oconv  # noqa: E402
from layerwise_b200.ops import nn as lnn  # noqa: E402
from layerwise_b200.parallel.exchange import CompressionConfig  # noqa: E402
from layerwise_b200.train.imagenet import Trainer  # noqa: E402

_orig_fork = oconv.fork_conv1x1


def run(name, backend, engine, s2d=True, fork=True, overlap=True, steps=12, method="Topk", native_ops=None):
    lnn.set_backend(backend)
    if not fork:
        oconv.fork_conv1x1 = lambda m, x: (m(x), x)
    else:
        oconv.fork_conv1x1 = _orig_fork
    saved = {}
    if native_ops is not None:      # enable only a subset of native layer ops
        import layerwise_b200.ops.bn as obn, layerwise_b200.ops.pool as opool, layerwise_b200.ops.loss as oloss, layerwise_b200.ops.gemm as ogemm
        for mod, fn, key in ((obn, "maybe_batch_norm", "bn"), (opool, "maybe_max_pool", "pool"), (oloss, "maybe_cross_entropy", "ce"), (ogemm, "maybe_linear", "linear")):
            saved[(mod, fn)] = getattr(mod, fn)
            if key not in native_ops:
                setattr(mod, fn, lambda *a, **k: None)
    try:
        torch.manual_seed(0)
        cfg = CompressionConfig(compress="layerwise", method=method, ratio=0.25, error_feedback=True)
        tr = Trainer(arch="resnet50", num_classes=16, dtype="bf16", cfg=cfg, lr=0.02, momentum=0.9, weight_decay=1e-4,
                     device="cuda:0", overlap=overlap and engine != "torch", engine=engine, init_bn0=False)
        tr.normalize.s2d = tr.normalize.s2d and s2d
        g = torch.Generator().manual_seed(1)
        x = torch.randint(0, 256, (32, 3, 64, 64), dtype=torch.uint8, generator=g).cuda()
        y = torch.randint(0, 16, (32,), generator=g).cuda()
        losses = [float(tr.train_step(x, y)[1].detach()) for _ in range(steps)]
        print(f"{name:34s} " + " ".join(f"{l:7.3f}" for l in losses), flush=True)
    finally:
        for (mod, fn), f in saved.items():
            setattr(mod, fn, f)
        lnn.set_backend("auto")
        oconv.fork_conv1x1 = _orig_fork


run("stock ops + torch engine", "torch", "torch", s2d=False, fork=False)
run("stock ops + torch engine (again)", "torch", "torch", s2d=False, fork=False)
run("stock ops + fused engine/SGD", "torch", "auto", s2d=False, fork=False)
run("stock ops + fused, no overlap", "torch", "auto", s2d=False, fork=False, overlap=False)
run("stock ops + fused, method none", "torch", "auto", s2d=False, fork=False, method="none")
run("stock ops + torch, method none", "torch", "torch", s2d=False, fork=False, method="none")
run("native bn only + torch engine", "auto", "torch", s2d=False, fork=False, native_ops={"bn"})
run("native pool only + torch engine", "auto", "torch", s2d=False, fork=False, native_ops={"pool"})
run("native ce only + torch engine", "auto", "torch", s2d=False, fork=False, native_ops={"ce"})
run("native linear only + torch engine", "auto", "torch", s2d=False, fork=False, native_ops={"linear"})
run("s2d stem only + torch engine", "torch", "torch", s2d=True, fork=False)
run("fork conv only + torch engine", "torch", "torch", s2d=False, fork=True)
run("all native", "auto", "auto")
