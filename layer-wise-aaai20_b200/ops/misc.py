"""Small model ops on hand-written kernels (``csrc/misc.cu``): stand-alone ReLU, logit scale, dropout, adaptive /
global average pooling, argmax-correct and top-1 / top-5 counters.

Reference ops: ``nn.ReLU`` after biased layers (CIFAR10/alexnet.py:31-36, CIFAR10/vgg16.py:24-30), ``Mul(0.125)``
(CIFAR10/torch_backend.py:72-77), ``nn.Dropout`` (alexnet.py:30,33; vgg16.py:26,29), ``nn.AdaptiveAvgPool2d``
(IMAGENET/training/resnet.py:109, vgg16.py:22), ``Correct`` (torch_backend.py:88-90), ``accuracy`` / ``correct``
(IMAGENET/training/train_imagenet_nv.py:679-689).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _native
from .._native import abi

_DT = (torch.float32, torch.bfloat16, torch.float16)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dense(x: torch.Tensor) -> bool:
    """Any dense layout works for elementwise ops as long as input and output share it."""
    return x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last))


def usable(x: torch.Tensor) -> bool:
    return x.is_cuda and x.dtype in _DT and _native.available() and x.numel() > 0 and _dense(x) and x.data_ptr() % 16 == 0


def _eltwise(x, aux, op: int, alpha: float = 1.0):
    y = torch.empty_like(x)
    abi.check(_native.lib().lwc_eltwise(x.data_ptr(), 0 if aux is None else aux.data_ptr(), y.data_ptr(), x.numel(), op,
                                        float(alpha), abi.dtype_code(x.dtype), _stream()), "lwc_eltwise")
    _native.counters.add(1)
    return y


class _ReLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = _eltwise(x, None, 0)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        if dy.stride() != y.stride():
            dy = dy.contiguous(memory_format=torch.channels_last) if (y.dim() == 4 and not y.is_contiguous()) else dy.contiguous()
        return _eltwise(dy, y, 2)


def relu(x: torch.Tensor) -> torch.Tensor:
    return _ReLU.apply(x) if usable(x) else torch.relu(x)


class _Scale(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, alpha):
        ctx.alpha = alpha
        return _eltwise(x, None, 1, alpha)

    @staticmethod
    def backward(ctx, dy):
        if not _dense(dy):
            dy = dy.contiguous()
        return _eltwise(dy, None, 1, ctx.alpha), None


def scale(x: torch.Tensor, alpha: float) -> torch.Tensor:
    return _Scale.apply(x, float(alpha)) if usable(x) else x * alpha


_DROPOUT_CALLS = {"n": 0}


class _Dropout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed, offset):
        n = x.numel()
        y = torch.empty_like(x)
        mask = torch.empty((n + 7) // 8, dtype=torch.uint8, device=x.device)
        abi.check(_native.lib().lwc_dropout_forward(x.data_ptr(), y.data_ptr(), mask.data_ptr(), n, float(p), int(seed), int(offset),
                                                    abi.dtype_code(x.dtype), _stream()), "lwc_dropout_forward")
        _native.counters.add(1)
        ctx.save_for_backward(mask)
        ctx.p = p
        return y

    @staticmethod
    def backward(ctx, dy):
        (mask,) = ctx.saved_tensors
        if not dy.is_contiguous():
            dy = dy.contiguous()
        dx = torch.empty_like(dy)
        abi.check(_native.lib().lwc_dropout_backward(dy.data_ptr(), mask.data_ptr(), dx.data_ptr(), dy.numel(), float(ctx.p),
                                                     abi.dtype_code(dy.dtype), _stream()), "lwc_dropout_backward")
        _native.counters.add(1)
        return dx, None, None, None


def dropout(x: torch.Tensor, p: float, training: bool = True, seed: Optional[int] = None) -> torch.Tensor:
    """Inverted dropout.  The Philox stream is keyed by (seed, call counter): every call draws fresh bits, and two
    processes seeded alike draw the same masks (the seed comes from ``torch.initial_seed()`` unless given)."""
    if not training or p <= 0.0:
        return x
    if not (usable(x) and x.is_contiguous() and p < 1.0):
        return torch.nn.functional.dropout(x, p, True)
    _DROPOUT_CALLS["n"] += 1
    s = torch.initial_seed() if seed is None else seed
    return _Dropout.apply(x, float(p), s & 0xFFFFFFFFFFFFFFFF, _DROPOUT_CALLS["n"] << 40)


class _AdaptiveAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, oh, ow):
        n, c, ih, iw = x.shape
        y = torch.empty((n, c, oh, ow), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        abi.check(_native.lib().lwc_adaptive_avgpool(x.data_ptr(), y.data_ptr(), n, ih, iw, oh, ow, c, 0, abi.dtype_code(x.dtype),
                                                     _stream()), "lwc_adaptive_avgpool")
        _native.counters.add(1)
        ctx.shape = (n, c, ih, iw, oh, ow)
        return y

    @staticmethod
    def backward(ctx, dy):
        n, c, ih, iw, oh, ow = ctx.shape
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((n, c, ih, iw), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        abi.check(_native.lib().lwc_adaptive_avgpool(dy.data_ptr(), dx.data_ptr(), n, ih, iw, oh, ow, c, 1, abi.dtype_code(dy.dtype),
                                                     _stream()), "lwc_adaptive_avgpool")
        _native.counters.add(1)
        return dx, None, None


def adaptive_avg_pool2d(x: torch.Tensor, output_size) -> torch.Tensor:
    oh, ow = (output_size, output_size) if isinstance(output_size, int) else tuple(output_size)
    ok = (x.is_cuda and x.dim() == 4 and x.dtype in _DT and _native.available() and x.shape[1] % 8 == 0
          and x.is_contiguous(memory_format=torch.channels_last) and oh is not None and ow is not None)
    if not ok:
        return torch.nn.functional.adaptive_avg_pool2d(x, output_size)
    return _AdaptiveAvgPool.apply(x, int(oh), int(ow))


def correct(logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """``logits.max(1)[1] == target`` as a bool tensor, one kernel."""
    if not (logits.is_cuda and logits.dim() == 2 and logits.dtype in _DT and _native.available() and logits.is_contiguous()
            and target.dtype == torch.int64 and target.is_contiguous()):
        return logits.max(dim=1)[1] == target
    n, c = logits.shape
    out = torch.empty(n, dtype=torch.uint8, device=logits.device)
    abi.check(_native.lib().lwc_accuracy(logits.data_ptr(), target.data_ptr(), out.data_ptr(), 0, n, c, abi.dtype_code(logits.dtype),
                                         _stream()), "lwc_accuracy")
    _native.counters.add(1)
    return out.view(torch.bool)


def topk_hits(logits: torch.Tensor, target: torch.Tensor, counters: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Adds (top-1 hits, top-5 hits) of this batch to ``counters`` (float32 [2] on the device; created when None) and
    returns it.  Nothing is read back: callers accumulate over a whole epoch and synchronise once."""
    if counters is None:
        counters = torch.zeros(2, dtype=torch.float32, device=logits.device)
    ok = (logits.is_cuda and logits.dim() == 2 and logits.dtype in _DT and _native.available() and logits.is_contiguous()
          and target.dtype == torch.int64 and target.is_contiguous())
    if not ok:
        k = min(5, logits.shape[1])
        pred = logits.float().topk(k, 1)[1]
        hit = pred == target.view(-1, 1)
        counters += torch.stack([hit[:, :1].sum(), hit.sum()]).to(counters.dtype)
        return counters
    n, c = logits.shape
    abi.check(_native.lib().lwc_accuracy(logits.data_ptr(), target.data_ptr(), 0, counters.data_ptr(), n, c,
                                         abi.dtype_code(logits.dtype), _stream()), "lwc_accuracy")
    _native.counters.add(1)
    return counters
