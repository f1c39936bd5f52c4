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
    def forward(ctx, logits, target, reduction, ignore_index=-100, trust_targets=False):
        lib = _native.lib()
        N, Cc = logits.shape
        loss = torch.empty(N, dtype=torch.float32, device=logits.device)
        lse = torch.empty(N, dtype=torch.float32, device=logits.device)
        st = torch.cuda.current_stream().cuda_stream
        abi.check(lib.lwc_cross_entropy_forward(logits.data_ptr(), target.data_ptr(), loss.data_ptr(), lse.data_ptr(),
                                                N, Cc, abi.dtype_code(logits.dtype), C.c_void_p(st)), "lwc_cross_entropy_forward")
        _native.counters.add(1)
        ctx.reduction = reduction
        # The kernel gives targets outside [0, C) -- which includes the default ignore_index -100 -- zero loss and zero
        # gradient.  torch's 'mean' divides by the number of NON-ignored targets: count them on the device (no host
        # sync) unless the caller vouches that every target is a class index (`trust_targets`, the trainers do).
        cnt = None
        if reduction == "mean" and not trust_targets:
            cnt = (target != ignore_index).sum().clamp_min(1).to(torch.float32)
        ctx.save_for_backward(logits, target, lse, cnt)
        if reduction == "none":
            return loss
        if reduction == "sum":
            return loss.sum()
        return loss.mean() if cnt is None else loss.sum() / cnt

    @staticmethod
    def backward(ctx, g):
        lib = _native.lib()
        logits, target, lse, cnt = ctx.saved_tensors
        N, Cc = logits.shape
        d = torch.empty_like(logits)
        st = torch.cuda.current_stream().cuda_stream
        if ctx.reduction == "none":
            gp, scale = g.float().contiguous(), 1.0
            gptr = gp.data_ptr()
        else:
            # scalar upstream gradient: fold it in without a host sync by broadcasting it per sample
            if ctx.reduction == "mean":
                gp = (g.float() / (cnt if cnt is not None else float(N))).expand(N).contiguous()
            else:
                gp = g.float().expand(N).contiguous()
            gptr, scale = gp.data_ptr(), 1.0
        abi.check(lib.lwc_cross_entropy_backward(logits.data_ptr(), target.data_ptr(), lse.data_ptr(), gptr, d.data_ptr(),
                                                 N, Cc, abi.dtype_code(logits.dtype), C.c_float(scale), C.c_void_p(st)),
                  "lwc_cross_entropy_backward")
        _native.counters.add(1)
        return d, None, None, None, None


def maybe_cross_entropy(module, logits, target):
    if (logits.dim() != 2 or not logits.is_contiguous() or target.dtype != torch.int64 or target.dim() != 1
            or module.weight is not None or module.label_smoothing != 0.0 or module.ignore_index >= 0
            or logits.dtype not in (torch.float32, torch.bfloat16, torch.float16)):
        return None
    return _CrossEntropy.apply(logits, target.contiguous(), module.reduction, int(module.ignore_index),
                               bool(getattr(module, "trust_targets", False)))
