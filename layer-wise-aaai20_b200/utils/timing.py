"""Wall-clock and device timers.

``Timer`` mirrors the behaviour of the reference's lap timer
(CIFAR10/core.py:14-27): each call returns the time since the previous call and
optionally adds it to ``total_time``.  ``DeviceTimer``/``PhaseTimer`` are the
B200 additions: CUDA-event timing on the launching stream, reported
as a max over ranks.
"""
from __future__ import annotations

import contextlib
import time
from collections import defaultdict
from typing import Callable, Dict, Optional

import torch


def localtime() -> str:
    return time.strftime("%Y-%m-%d %H:%M:%S", time.localtime())


class Timer:
    """Lap timer; ``synch`` (e.g. ``torch.cuda.synchronize``) runs before each reading."""

    def __init__(self, synch: Optional[Callable[[], None]] = None):
        self.synch = synch or (lambda: None)
        self.synch()
        self.times = [time.time()]
        self.total_time = 0.0

    def __call__(self, include_in_total: bool = True) -> float:
        self.synch()
        self.times.append(time.time())
        lap = self.times[-1] - self.times[-2]
        if include_in_total:
            self.total_time += lap
        return lap


class DeviceTimer:
    """CUDA-event stopwatch on the current stream (falls back to wall clock on CPU)."""

    def __init__(self, device: Optional[torch.device] = None):
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self._t0 = None
        self._e0 = self._e1 = None

    def start(self, stream=None):
        if self.cuda:
            self._e0 = torch.cuda.Event(enable_timing=True)
            self._e1 = torch.cuda.Event(enable_timing=True)
            self._e0.record(stream)
        else:
            self._t0 = time.perf_counter()
        return self

    def stop(self, stream=None):
        if self.cuda:
            self._e1.record(stream)
        else:
            self._t1 = time.perf_counter()
        return self

    def elapsed_ms(self) -> float:
        if self.cuda:
            self._e1.synchronize()
            return float(self._e0.elapsed_time(self._e1))
        return (self._t1 - self._t0) * 1e3


class PhaseTimer:
    """Accumulates device time per named phase (compress / exchange / fwd / bwd / opt).

    Events are recorded without synchronising; ``summary()`` synchronises once and
    (if a process group is up) reduces each phase with MAX over ranks.
    """

    def __init__(self, enabled: bool = True):
        self.enabled = enabled and torch.cuda.is_available()
        self._pending = defaultdict(list)
        self._wall: Dict[str, float] = defaultdict(float)
        self._open = {}

    def start(self, name: str, stream=None):
        if self.enabled:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            self._open[name] = e
        else:
            self._open[name] = time.perf_counter()

    def stop(self, name: str, stream=None):
        if name not in self._open:
            return
        if self.enabled:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            self._pending[name].append((self._open.pop(name), e))
        else:
            self._wall[name] += (time.perf_counter() - self._open.pop(name)) * 1e3

    def summary(self, reduce_max: bool = True) -> Dict[str, float]:
        out = dict(self._wall)
        if self.enabled:
            torch.cuda.synchronize()
            for k, evs in self._pending.items():
                out[k] = out.get(k, 0.0) + sum(a.elapsed_time(b) for a, b in evs)
        import torch.distributed as dist
        if reduce_max and dist.is_available() and dist.is_initialized() and out:
            keys = sorted(out)
            dev = "cuda" if (self.enabled and dist.get_backend() == "nccl") else "cpu"
            t = torch.tensor([out[k] for k in keys], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out = {k: float(v) for k, v in zip(keys, t.tolist())}
        return out

    def reset(self):
        self._pending.clear()
        self._wall.clear()
        self._open.clear()

    @contextlib.contextmanager
    def phase(self, name: str, stream=None):
        """``with timer.phase('forward'):`` -- a device-timed phase that is also an NVTX range."""
        from .nvtx import nvtx_range
        with nvtx_range(name):
            self.start(name, stream)
            try:
                yield
            finally:
                self.stop(name, stream)

    def table(self, steps: int = 1, reduce_max: bool = True) -> str:
        """Per-phase milliseconds per step (max over ranks) as a fixed-width table."""
        s = self.summary(reduce_max)
        tot = sum(s.values()) or 1.0
        lines = [f"{'phase':<28s}{'ms/step':>10s}{'share':>8s}"]
        for k, v in sorted(s.items(), key=lambda kv: -kv[1]):
            lines.append(f"{k:<28s}{v / max(steps, 1):10.3f}{100.0 * v / tot:7.1f}%")
        return "\n".join(lines)


@contextlib.contextmanager
def maybe_phase(timer: Optional["PhaseTimer"], name: str):
    """No-op without a timer: the hot loops call this unconditionally."""
    if timer is None:
        yield
    else:
        with timer.phase(name):
            yield
