"""Device scalars read on the host a fixed number of steps late.

``float(loss)`` right after a step makes the host wait for the step it has just launched, and the device then idles
until the next launch arrives.  ``LaggedReader.push(loss)`` queues an asynchronous copy of the scalar into pinned host
memory behind the step and hands back the value pushed ``depth`` calls earlier -- which has long landed -- so every
step's result still reaches the host, one step late, without a bubble.  The reference reads ``reduced_loss.item()``
every iteration (IMAGENET/training/train_imagenet_nv.py:432-441); its trainers here read their device-side meters once
per ``print_freq`` steps instead, this is the per-step variant for callers that want every value.
"""
from __future__ import annotations

from collections import deque
from typing import Deque, List, Optional, Tuple

import torch


class LaggedReader:
    def __init__(self, depth: int = 1):
        self.depth = max(0, int(depth))
        self._pending: Deque[Tuple[torch.Tensor, Optional[torch.cuda.Event]]] = deque()
        self._free: List[torch.Tensor] = []

    def _slot(self, like: torch.Tensor) -> torch.Tensor:
        for i, s in enumerate(self._free):
            if s.dtype == like.dtype:
                return self._free.pop(i)
        pin = like.is_cuda and torch.cuda.is_available()
        return torch.empty((), dtype=like.dtype, pin_memory=pin)

    def _take(self) -> float:
        host, ev = self._pending.popleft()
        if ev is not None:
            ev.synchronize()
        v = float(host)
        self._free.append(host)
        return v

    def push(self, scalar: torch.Tensor) -> Optional[float]:
        """Queue ``scalar`` (0-d tensor) for read-back; returns the value queued ``depth`` pushes ago, if any."""
        t = scalar.detach().reshape(())
        host = self._slot(t)
        host.copy_(t, non_blocking=True)
        ev = None
        if t.is_cuda:
            ev = torch.cuda.Event()
            ev.record()
        self._pending.append((host, ev))
        return self._take() if len(self._pending) > self.depth else None

    def drain(self) -> List[float]:
        """Wait for and return everything still queued (call once after the last step)."""
        out = []
        while self._pending:
            out.append(self._take())
        return out
