"""Cross-entropy through ``csrc/loss.cu``: one forward kernel (per-sample loss + log-sum-exp) and one backward
kernel (softmax - onehot, scaled), replacing ATen's log_softmax + nll_loss pairs.

Reference ops: ``nn.CrossEntropyLoss(reduction='none')`` (CIFAR10/dawn.py:85) and the mean-reduced criterion of
IMAGENET/training/train_imagenet_nv.py:187.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native
from .._native import abi


class _CrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, reduction):
        lib = _native.lib()
        N, Cc = logits.shape
        loss = torch.empty(N, dtype=torch.float32, device=logits.device)
        lse = torch.empty(N, dtype=torch.float32, device=logits.device)
        st = torch.cuda.current_stream().cuda_stream
        abi.check(lib.lwc_cross_entropy_forward(logits.data_ptr(), target.data_ptr(), loss.data_ptr(), lse.data_ptr(),
                                                N, Cc, abi.dtype_code(logits.dtype), C.c_void_p(st)), "lwc_cross_entropy_forward")
        _native.counters.add(1)
        ctx.save_for_backward(logits, target, lse)
        ctx.reduction = reduction
        if reduction == "none":
            return loss
        return loss.sum() if reduction == "sum" else loss.mean()

    @staticmethod
    def backward(ctx, g):
        lib = _native.lib()
        logits, target, lse = ctx.saved_tensors
        N, Cc = logits.shape
        d = torch.empty_like(logits)
        st = torch.cuda.current_stream().cuda_stream
        if ctx.reduction == "none":
            gp, scale = g.float().contiguous(), 1.0
            gptr = gp.data_ptr()
        else:
            # scalar upstream gradient: fold it in without a host sync by broadcasting it per sample
            gp = (g.float() * (1.0 / N if ctx.reduction == "mean" else 1.0)).expand(N).contiguous()
            gptr, scale = gp.data_ptr(), 1.0
        abi.check(lib.lwc_cross_entropy_backward(logits.data_ptr(), target.data_ptr(), lse.data_ptr(), gptr, d.data_ptr(),
                                                 N, Cc, abi.dtype_code(logits.dtype), C.c_float(scale), C.c_void_p(st)),
                  "lwc_cross_entropy_backward")
        _native.counters.add(1)
        return d, None, None


def maybe_cross_entropy(module, logits, target):
    if (logits.dim() != 2 or not logits.is_contiguous() or target.dtype != torch.int64 or target.dim() != 1
            or module.weight is not None or module.label_smoothing != 0.0 or module.ignore_index >= 0
            or logits.dtype not in (torch.float32, torch.bfloat16, torch.float16)):
        return None
    return _CrossEntropy.apply(logits, target.contiguous(), module.reduction)
