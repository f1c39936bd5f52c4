"""CIFAR-10 training loop (the ``dawn.py`` recipe).

Call structure follows the reference (CIFAR10/core.py:303-341): ``run_batches`` ->
``train_epoch`` -> ``train``; per step: forward, ``loss.sum().backward()``, gradient
exchange (layer-wise / entire-model / plain averaging), optimiser step, zero grads.
Differences: the exchange is a ``GradientExchanger`` (fused CUDA engine on B200, reference
math on CPU/gloo); no per-step host sync (stats stay on device until the epoch summary);
process-group creation accepts NCCL as well as gloo.
"""
from __future__ import annotations

from typing import Optional

import torch

from ..parallel import comm
from ..parallel.exchange import CompressionConfig, GradientExchanger
from ..utils.logging import StatsLogger
from ..utils.misc import union
from ..utils.timing import Timer


def run_batches(model, batches, training, exchanger: Optional[GradientExchanger] = None,
                optimizer_step=None, stats=None, max_batches: Optional[int] = None):
    stats = stats or StatsLogger(("loss", "correct"))
    model.train(training)
    ctx = torch.enable_grad() if training else torch.no_grad()
    with ctx:
        for i, batch in enumerate(batches):
            if max_batches is not None and i >= max_batches:
                break
            output = model(batch)
            stats.append(output)
            if training:
                output["loss"].sum().backward()
                if exchanger is not None:
                    exchanger.exchange()
                optimizer_step()
                model.zero_grad(set_to_none=False)
    return stats


def train_epoch(model, train_batches, test_batches, optimizer_step, timer, exchanger=None,
                test_time_in_total=True, max_batches=None):
    train_stats = run_batches(model, train_batches, True, exchanger, optimizer_step,
                              max_batches=max_batches)
    train_time = timer()
    test_stats = run_batches(model, test_batches, False, max_batches=max_batches)
    test_time = timer(test_time_in_total)
    return {
        "train time": train_time, "train loss": train_stats.mean("loss"),
        "train acc": train_stats.mean("correct"),
        "test time": test_time, "test loss": test_stats.mean("loss"),
        "test acc": test_stats.mean("correct"),
        "total time": timer.total_time,
    }


def train(model, optimizer, train_batches, test_batches, epochs, master_address=None,
          world_size=1, rank=0, loggers=(), test_time_in_total=True, timer=None, compress=None,
          method=None, K=None, V=None, qstates=None, error_feedback=False, backend=None,
          engine="auto", max_batches=None, cfg: Optional[CompressionConfig] = None):
    """Train for ``epochs``; returns the last epoch's summary dict."""
    if world_size > 1 or master_address:
        comm.init_process_group(backend=backend, init_method=master_address,
                                world_size=world_size, rank=rank)
    cfg = cfg or CompressionConfig(compress=compress or "none", method=method or "none",
                                   ratio=K or 0, threshold=V or 0, qstates=qstates or 0,
                                   error_feedback=error_feedback)
    exchanger = GradientExchanger([p for p in model.parameters() if p.requires_grad], cfg,
                                  engine=engine)
    timer = timer or Timer(torch.cuda.synchronize if torch.cuda.is_available() else None)
    summary = {}
    for epoch in range(epochs):
        stats = train_epoch(model, train_batches, test_batches, optimizer.step, timer, exchanger,
                            test_time_in_total=test_time_in_total, max_batches=max_batches)
        lr = optimizer.param_values()["lr"] * train_batches.batch_size \
            if hasattr(optimizer, "param_values") else optimizer.param_groups[0]["lr"]
        summary = union({"epoch": epoch + 1, "lr": lr}, stats)
        for logger in loggers:
            logger.append(summary)
    return summary


def warmup(model, batch_size: int, device=None, dtype=torch.float32, channels_last: bool = False):
    """One forward/backward on random inputs before timing starts (kernel selection, allocator
    warm-up, native-library load) -- the role of the reference's ``warmup_cudnn``
    (CIFAR10/torch_backend.py:18-29)."""
    device = device or next(model.parameters()).device
    x = torch.rand(batch_size, 3, 32, 32, device=device).to(dtype)
    if channels_last:
        x = x.contiguous(memory_format=torch.channels_last)
    batch = {"input": x, "target": torch.randint(0, 10, (batch_size,), device=device)}
    model.train(True)
    out = model(batch)
    out["loss"].sum().backward()
    model.zero_grad(set_to_none=False)
    if torch.cuda.is_available():
        torch.cuda.synchronize()
