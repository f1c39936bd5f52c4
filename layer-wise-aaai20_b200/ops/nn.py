"""Layer modules used by the model zoo.

Every class subclasses the matching ``torch.nn`` module (so ``isinstance`` checks,
``state_dict`` keys and initialisers behave as usual) and routes ``forward`` to the
hand-written sm_100a kernels when (a) the tensors live on a CUDA device, (b) the
native library is loaded and (c) the native backend is enabled for that op (see
``set_backend``).  On CPU, or for shapes a kernel does not cover, the stock
PyTorch implementation runs -- that path is also the numerical oracle in tests.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

_BACKEND = {"name": "auto"}


def set_backend(name: str):
    """``'auto'`` (native kernels where available), ``'torch'`` (stock ops only)."""
    if name not in ("auto", "torch", "native"):
        raise ValueError(name)
    _BACKEND["name"] = name


def get_backend() -> str:
    return _BACKEND["name"]


def _native_ok(x: torch.Tensor) -> bool:
    if _BACKEND["name"] == "torch" or not x.is_cuda:
        return False
    from .. import _native
    return _native.available()


class Conv2d(nn.Conv2d):
    def forward(self, x):
        if _native_ok(x):
            from . import conv as _conv
            y = _conv.maybe_conv2d(self, x, relu=getattr(self, "fuse_relu", False))
            if y is not None:
                return y
        y = super().forward(x)
        return F.relu(y) if getattr(self, "fuse_relu", False) else y


class Linear(nn.Linear):
    def forward(self, x):
        if _native_ok(x):
            from . import gemm as _gemm
            y = _gemm.maybe_linear(self, x)
            if y is not None:
                return y
        y = super().forward(x)
        return F.relu(y) if getattr(self, "fuse_relu", False) else y


class BatchNorm2d(nn.BatchNorm2d):
    """Batch norm with an optional fused epilogue: ``relu(bn(x) + residual)``.

    ``forward(x, residual=None, relu=False)`` -- callers that used to write
    ``relu(bn(x))`` or ``relu(bn(x) + identity)`` pass the flags instead, and the fused
    sm_100a kernels (``ops/bn.py``) do it in one pass structure; on CPU / unsupported
    layouts the same math runs with stock ops.
    """

    def forward(self, x, residual=None, relu=False, fuse_bwd=False):
        """``fuse_bwd``: the output has exactly ONE consumer, a convolution -- its data gradient may then produce this
        layer's backward reduction sums in its epilogue (``ops/conv.py::conv_dgrad_bnstats``)."""
        relu = relu or getattr(self, "fuse_relu", False)      # graph nets mark conv-bn-relu triples
        if _native_ok(x):
            from . import bn as _bn
            y = _bn.maybe_batch_norm(self, x, residual, relu, fuse_bwd=fuse_bwd)
            if y is not None:
                return y
        y = super().forward(x)
        if residual is not None:
            y = y + residual
        return F.relu(y) if relu else y


class ReLU(nn.ReLU):
    """``nn.ReLU``; ``already_applied=True`` marks a node whose producer (a fused batch norm) has applied the
    activation already, so the node is kept for graph/name compatibility but does no work."""

    def forward(self, x):
        if getattr(self, "already_applied", False):
            return x
        if _native_ok(x) and not (self.inplace and x.requires_grad and x.is_leaf):
            from . import misc as _misc
            return _misc.relu(x)
        return super().forward(x)


class Dropout(nn.Dropout):
    """Inverted dropout on the Philox kernel of ``csrc/misc.cu`` (1-bit mask saved for the backward pass)."""

    def forward(self, x):
        if self.training and _native_ok(x):
            from . import misc as _misc
            return _misc.dropout(x, self.p, True)
        return super().forward(x)


class AdaptiveAvgPool2d(nn.AdaptiveAvgPool2d):
    def forward(self, x):
        if _native_ok(x):
            from . import misc as _misc
            return _misc.adaptive_avg_pool2d(x, self.output_size)
        return super().forward(x)


class MaxPool2d(nn.MaxPool2d):
    def forward(self, x):
        if _native_ok(x):
            from . import pool as _pool
            y = _pool.maybe_max_pool(self, x)
            if y is not None:
                return y
        return super().forward(x)


class CrossEntropyLoss(nn.CrossEntropyLoss):
    def forward(self, logits, target):
        if _native_ok(logits):
            from . import loss as _loss
            y = _loss.maybe_cross_entropy(self, logits, target)
            if y is not None:
                return y
        return super().forward(logits.float() if logits.dtype != torch.float32 else logits, target)


def add_relu(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """``relu(x + y)`` -- the residual join (reference resnet.py:89-90)."""
    return F.relu(x + y)
