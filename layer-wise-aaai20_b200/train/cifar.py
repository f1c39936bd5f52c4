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


def _zero_grads(model, exchanger):
    eng = getattr(exchanger, "engine", None)
    if eng is not None and hasattr(eng, "zero_grad"):
        eng.zero_grad()          # fused engine: grads dropped (autograd's next tensors are read in place) or one memset
    else:
        model.zero_grad(set_to_none=False)


class GraphedTrainStep:
    """Whole CIFAR training step (forward, ``loss.sum().backward()``, compressed exchange, SGD, zero grads) captured
    once as a CUDA graph and replayed per batch -- the step is launch-bound (~80 % GPU-busy when stepped eagerly).

    Needs the fused exchange engine (its step counter moves to device memory) and a ``ScheduledOptimizer`` over
    ``FusedSGD`` whose only scheduled hyper-parameter is the learning rate (mirrored to device memory before every
    replay).  The first ``warmup`` batches of a shape step eagerly (cuDNN autotuning, momentum buffers, plans), the
    next one is captured; batches of another shape keep stepping eagerly.
    """

    def __init__(self, model, exchanger, optimizer, warmup: int = 3):
        self.model, self.exchanger, self.optimizer = model, exchanger, optimizer
        self.warmup = warmup
        self.graph = None
        self.shape = None
        self._seen_shape, self._seen = None, 0
        self._side = None
        eng = getattr(exchanger, "engine", None)
        inner = getattr(optimizer, "_opt", None)
        scheduled = [k for k, v in getattr(optimizer, "opt_params", {}).items() if callable(v) and k != "lr"]
        self.ok = (eng is not None and hasattr(eng, "enable_graph_mode") and inner is not None
                   and hasattr(inner, "enable_device_lr") and not scheduled)

    def _eager(self, batch):
        output = self.model(batch)
        output["loss"].sum().backward()
        self.exchanger.exchange()
        self.optimizer.step()
        _zero_grads(self.model, self.exchanger)
        return output

    def _body(self, batch):
        """The captured part: like ``_eager`` but the optimiser's host-side schedule update stays outside."""
        output = self.model(batch)
        output["loss"].sum().backward()
        self.exchanger.exchange()
        self.optimizer._opt.step()
        _zero_grads(self.model, self.exchanger)
        return {"loss": output["loss"].detach(), "correct": output["correct"].detach()}

    def _advance_schedule(self):
        opt = self.optimizer
        opt.step_number += 1
        opt._opt.param_groups[0].update(**opt.param_values())
        opt._opt.sync_device_lr()

    def _capture(self, batch):
        from .. import _native
        eng = self.exchanger.engine
        dev = batch["input"].device
        self.static = {k: v.clone() for k, v in batch.items() if torch.is_tensor(v)}
        eng.enable_graph_mode()
        self.optimizer._opt.enable_device_lr(dev)
        torch.cuda.synchronize(dev)
        graph = torch.cuda.CUDAGraph()
        n0, step0 = _native.counters.kernels, eng.step
        with torch.cuda.graph(graph, capture_error_mode="thread_local"):
            self.out = self._body(self.static)
        self.launches = _native.counters.kernels - n0
        _native.counters.add(-self.launches)
        eng.step = step0
        self.graph = graph
        self.shape = tuple(batch["input"].shape)

    def __call__(self, batch):
        if not self.ok or not batch["input"].is_cuda:
            return self._eager(batch)
        shape = tuple(batch["input"].shape)
        if self.graph is None:
            self._seen = self._seen + 1 if shape == self._seen_shape else 1
            self._seen_shape = shape
            if self._seen <= self.warmup:
                # warm-up steps run on a side stream: autograd state created on the legacy default stream would make
                # the later capture fail ("legacy stream depend on a capturing blocking stream")
                if self._side is None:
                    self._side = torch.cuda.Stream(device=batch["input"].device)
                cur = torch.cuda.current_stream()
                self._side.wait_stream(cur)
                with torch.cuda.stream(self._side):
                    out = self._eager(batch)
                cur.wait_stream(self._side)
                return out
            self._capture(batch)                  # executes nothing; this batch is the first replay
        if shape != self.shape:
            return self._eager(batch)
        from .. import _native
        for k, v in self.static.items():
            v.copy_(batch[k], non_blocking=True)
        self._advance_schedule()
        self.graph.replay()
        _native.counters.add(self.launches)
        self.exchanger.engine.note_replay()
        # the static outputs are overwritten by the next replay: hand out copies
        return {k: v.clone() for k, v in self.out.items()}


def run_batches(model, batches, training, exchanger: Optional[GradientExchanger] = None,
                optimizer_step=None, stats=None, max_batches: Optional[int] = None, graphed=None, phase_timer=None):
    """``phase_timer`` (utils.timing.PhaseTimer): CUDA-event time + NVTX range per phase of every training step."""
    from ..utils.timing import maybe_phase
    stats = stats or StatsLogger(("loss", "correct"))
    model.train(training)
    ctx = torch.enable_grad() if training else torch.no_grad()
    with ctx:
        for i, batch in enumerate(batches):
            if max_batches is not None and i >= max_batches:
                break
            if training and graphed is not None:
                stats.append(graphed(batch))
                continue
            with maybe_phase(phase_timer, "forward" if training else "eval forward"):
                output = model(batch)
            stats.append(output)
            if training:
                with maybe_phase(phase_timer, "backward (+hooked compress/send)"):
                    output["loss"].sum().backward()
                if exchanger is not None:
                    with maybe_phase(phase_timer, "exchange (exposed: join + receive)"):
                        exchanger.exchange()
                with maybe_phase(phase_timer, "optimizer"):
                    optimizer_step()
                    _zero_grads(model, exchanger)
    return stats


def train_epoch(model, train_batches, test_batches, optimizer_step, timer, exchanger=None,
                test_time_in_total=True, max_batches=None, graphed=None, phase_timer=None):
    train_stats = run_batches(model, train_batches, True, exchanger, optimizer_step,
                              max_batches=max_batches, graphed=graphed, phase_timer=phase_timer)
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
          engine="auto", max_batches=None, cfg: Optional[CompressionConfig] = None, cuda_graph: bool = False,
          phase_times: bool = False):
    """Train for ``epochs``; returns the last epoch's summary dict.  ``phase_times``: print a per-phase device-time
    table (forward / backward / exposed exchange / optimizer, max over ranks) after every epoch."""
    if world_size > 1 or master_address:
        comm.init_process_group(backend=backend, init_method=master_address,
                                world_size=world_size, rank=rank)
    cfg = cfg or CompressionConfig(compress=compress or "none", method=method or "none",
                                   ratio=K or 0, threshold=V or 0, qstates=qstates or 0,
                                   error_feedback=error_feedback)
    exchanger = GradientExchanger([p for p in model.parameters() if p.requires_grad], cfg,
                                  engine=engine)
    timer = timer or Timer(torch.cuda.synchronize if torch.cuda.is_available() else None)
    graphed = GraphedTrainStep(model, exchanger, optimizer) if (cuda_graph and not phase_times) else None
    from ..utils.timing import PhaseTimer
    ptimer = PhaseTimer() if phase_times else None
    summary = {}
    for epoch in range(epochs):
        stats = train_epoch(model, train_batches, test_batches, optimizer.step, timer, exchanger,
                            test_time_in_total=test_time_in_total, max_batches=max_batches, graphed=graphed,
                            phase_timer=ptimer)
        if ptimer is not None:
            nsteps = min(len(train_batches), max_batches) if max_batches else len(train_batches)
            table = ptimer.table(nsteps)
            ptimer.reset()
            if rank == 0:
                print(table, flush=True)
        exchanger.check_health()             # a peer whose flags stopped arriving raises here instead of hanging the job
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
