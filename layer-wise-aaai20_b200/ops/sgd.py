"""SGD with momentum / Nesterov / weight decay: fused multi-tensor CUDA step + schedules.

``FusedSGD`` subclasses ``torch.optim.SGD`` so its ``state_dict`` is byte-compatible with
the reference's checkpoints (the ``optimizer`` key of train_imagenet_nv.py:663-669) --
only ``step`` is replaced: on CUDA one multi-tensor kernel launch per dtype group updates
every parameter (``d = g*gscale + wd*w; v = mu*v + (1-damp)*d; w -= lr*(d + mu*v | v)``),
optionally also writing the bf16/fp16 model copy of fp32 master weights (the reference's
``master_params_to_model_params`` pass, fp16util.py:125-139, folded into the update).
The reference's torch-1.1 optimiser issues ~4 launches per tensor from a Python loop.

``ScheduledOptimizer`` is the CIFAR wrapper whose hyper-parameters may be callables of
the step number (reference TorchOptimiser, CIFAR10/torch_backend.py:122-143).
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional

import torch


def allow_lowp_grad(param: torch.Tensor) -> torch.Tensor:
    """Let an fp32 master weight carry a bf16/fp16 ``.grad`` (torch >= 2.9 checks ``grad_dtype``)."""
    if hasattr(param, "grad_dtype"):
        try:
            param.grad_dtype = None
        except Exception:
            pass
    return param


class FusedSGD(torch.optim.SGD):
    def __init__(self, params, lr=0.0, momentum=0.0, dampening=0.0, weight_decay=0.0,
                 nesterov=False, grad_scale: float = 1.0, model_params: Optional[List] = None):
        super().__init__(params, lr=lr, momentum=momentum, dampening=dampening,
                         weight_decay=weight_decay, nesterov=nesterov)
        self.grad_scale = grad_scale
        # optional low-precision copies, aligned with the flattened param order
        self._model_copy: Dict[int, torch.Tensor] = {}
        if model_params is not None:
            flat = [p for g in self.param_groups for p in g["params"]]
            for master, model in zip(flat, model_params):
                if model is not master:
                    self._model_copy[id(master)] = model
        self._plans = {}
        self._lr_dev = None          # graph-safe mode: one device scalar per param group
        self._lr_dev_val = {}

    def set_lr(self, lr: float):
        for g in self.param_groups:
            g["lr"] = lr

    def enable_device_lr(self, device):
        """Keep each group's learning rate in device memory (read by the kernel) so that a captured step can be
        replayed with a changing schedule; ``sync_device_lr`` uploads the values that changed."""
        if self._lr_dev is None:
            self._lr_dev = {gi: torch.zeros(1, dtype=torch.float32, device=device)
                            for gi in range(len(self.param_groups))}
        self.sync_device_lr()

    def sync_device_lr(self):
        if self._lr_dev is None:
            return
        for gi, g in enumerate(self.param_groups):
            lr = float(g["lr"])
            if self._lr_dev_val.get(gi) != lr:
                self._lr_dev[gi].fill_(lr)
                self._lr_dev_val[gi] = lr

    def _native(self, p) -> bool:
        if not p.is_cuda:
            return False
        from .. import _native
        return _native.available()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            params = [p for p in group["params"] if p.grad is not None]
            if not params:
                continue
            if self._native(params[0]):
                from . import sgd_native
                sgd_native.step_group(self, gi, group, params)
            else:
                self._step_reference(group, params)
        return loss

    def _step_reference(self, group, params):
        lr, mu, damp = group["lr"], group["momentum"], group["dampening"]
        wd, nesterov = group["weight_decay"], group["nesterov"]
        for p in params:
            d = p.grad.detach().to(p.dtype)
            if self.grad_scale != 1.0:
                d = d * self.grad_scale
            if wd != 0:
                d = d.add(p, alpha=wd)
            if mu != 0:
                st = self.state[p]
                buf = st.get("momentum_buffer")
                if buf is None:
                    buf = st["momentum_buffer"] = d.clone()
                else:
                    buf.mul_(mu).add_(d, alpha=1 - damp)
                d = d.add(buf, alpha=mu) if nesterov else buf
            p.add_(d, alpha=-lr)
            model = self._model_copy.get(id(p))
            if model is not None:
                model.data.copy_(p)


class ScheduledOptimizer:
    """Optimiser whose hyper-parameters can be functions of the step count."""

    def __init__(self, weights, optimizer_cls, step_number: int = 0, **opt_params):
        self.weights = list(weights)
        self.step_number = step_number
        self.opt_params = opt_params
        self._opt = optimizer_cls(self.weights, **self.param_values())

    def param_values(self):
        return {k: v(self.step_number) if callable(v) else v for k, v in self.opt_params.items()}

    def step(self):
        self.step_number += 1
        self._opt.param_groups[0].update(**self.param_values())
        self._opt.step()

    def zero_grad(self, set_to_none: bool = False):
        self._opt.zero_grad(set_to_none=set_to_none)

    def state_dict(self):
        return {"step_number": self.step_number, "opt": self._opt.state_dict()}

    def load_state_dict(self, sd):
        self.step_number = sd["step_number"]
        self._opt.load_state_dict(sd["opt"])

    @property
    def param_groups(self):
        return self._opt.param_groups

    def __repr__(self):
        return repr(self._opt)


def SGD(weights, lr=0, momentum=0, weight_decay=0, dampening=0, nesterov=False) -> ScheduledOptimizer:
    """CIFAR recipe helper (reference torch_backend.py:140-143) backed by ``FusedSGD``."""
    return ScheduledOptimizer(weights, FusedSGD, lr=lr, momentum=momentum,
                              weight_decay=weight_decay, dampening=dampening, nesterov=nesterov)
