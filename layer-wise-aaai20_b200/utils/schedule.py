"""Learning-rate schedules.

``PiecewiseLinear`` is the CIFAR schedule primitive (reference CIFAR10/core.py:157-159).
``PhaseScheduler`` implements the ImageNet *phases* mini-language
(train_imagenet_nv.py:602-651; phases built in IMAGENET/train.py:60-149): each LR
phase is ``{'ep': e | (e0, e1), 'lr': v | (v0, v1)}``; a pair means a linear ramp
evaluated per iteration.
"""
from __future__ import annotations

from collections import namedtuple
from collections.abc import Iterable
from typing import Callable, List, Optional

import numpy as np


class PiecewiseLinear(namedtuple("PiecewiseLinear", ("knots", "vals"))):
    def __call__(self, t):
        return float(np.interp([t], self.knots, self.vals)[0])


def listify(p=None, q=None):
    if p is None:
        p = []
    elif isinstance(p, str) or not isinstance(p, Iterable):
        p = [p]
    else:
        p = list(p)
    n = q if isinstance(q, int) else 1 if q is None else len(q)
    if len(p) == 1:
        p = p * n
    return p


class PhaseScheduler:
    def __init__(self, optimizer, phases: List[dict], on_change: Optional[Callable] = None):
        self.optimizer = optimizer
        self.current_lr = None
        self.on_change = on_change
        self.phases = [self._format(dict(p)) for p in phases]
        self.tot_epochs = max(max(p["ep"]) for p in self.phases)

    @staticmethod
    def _format(phase):
        phase["ep"] = listify(phase["ep"])
        phase["lr"] = listify(phase["lr"])
        if len(phase["lr"]) == 2 and len(phase["ep"]) != 2:
            raise ValueError("a linear LR phase needs an (start, end) epoch pair")
        return phase

    def get_current_phase(self, epoch):
        for phase in reversed(self.phases):
            if epoch >= phase["ep"][0]:
                return phase
        raise ValueError("epoch precedes the first LR phase")

    def get_lr(self, epoch, batch_curr, batch_tot):
        phase = self.get_current_phase(epoch)
        if len(phase["lr"]) == 1:
            return phase["lr"][0]
        lr0, lr1 = phase["lr"]
        ep0, ep1 = phase["ep"]
        if "epoch_step" in phase:
            batch_curr = 0
        steps = (ep1 - ep0) * batch_tot
        cur = (epoch - ep0) * batch_tot + batch_curr
        return lr0 + cur * (lr1 - lr0) / steps

    def update_lr(self, epoch, batch_num, batch_tot):
        lr = self.get_lr(epoch, batch_num, batch_tot)
        if lr == self.current_lr:
            return lr
        if self.on_change is not None and (batch_num == 1 or batch_num == batch_tot):
            self.on_change(self.current_lr, lr)
        self.current_lr = lr
        if hasattr(self.optimizer, "set_lr"):
            self.optimizer.set_lr(lr)
        else:
            for group in self.optimizer.param_groups:
                group["lr"] = lr
        return lr
