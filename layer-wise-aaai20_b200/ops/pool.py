"""Channels-last max pooling through ``csrc/pool.cu`` (forward + byte argmax, gather backward).

Reference ops replaced: ``nn.MaxPool2d`` 2x2/2, 4x4/4 (CIFAR10/dawn.py:51,:70-71) and 3x3/2 pad 1
(IMAGENET/training/resnet.py:104).
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native
from .._native import abi


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


def supported(module, x) -> bool:
    if not x.is_cuda or x.dim() != 4 or x.dtype not in (torch.bfloat16, torch.float16, torch.float32):
        return False
    if not x.is_contiguous(memory_format=torch.channels_last) or x.shape[1] % 8:
        return False
    k, s, p, d = _pair(module.kernel_size), _pair(module.stride or module.kernel_size), _pair(module.padding), _pair(module.dilation)
    if k[0] != k[1] or s[0] != s[1] or p[0] != p[1] or d != (1, 1) or module.ceil_mode or module.return_indices:
        return False
    return k[0] <= 15 and 2 * p[0] <= k[0]


class _MaxPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, pad):
        lib = _native.lib()
        N, Cc, H, W = x.shape
        OH, OW = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
        y = torch.empty((N, Cc, OH, OW), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        need_grad = x.requires_grad and torch.is_grad_enabled()
        idx = torch.empty(N * OH * OW * Cc, dtype=torch.uint8, device=x.device) if ctx.needs_input_grad[0] else None
        st = torch.cuda.current_stream().cuda_stream
        abi.check(lib.lwc_maxpool_forward(x.data_ptr(), y.data_ptr(), 0 if idx is None else idx.data_ptr(), N, H, W, Cc,
                                          k, s, pad, abi.dtype_code(x.dtype), C.c_void_p(st)), "lwc_maxpool_forward")
        _native.counters.add(1)
        ctx.geom = (N, H, W, Cc, k, s, pad)
        ctx.save_for_backward(idx)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _native.lib()
        (idx,) = ctx.saved_tensors
        N, H, W, Cc, k, s, pad = ctx.geom
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        st = torch.cuda.current_stream().cuda_stream
        abi.check(lib.lwc_maxpool_backward(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, Cc, k, s, pad,
                                           abi.dtype_code(dy.dtype), C.c_void_p(st)), "lwc_maxpool_backward")
        _native.counters.add(1)
        return dx, None, None, None


def maybe_max_pool(module, x):
    if not supported(module, x):
        return None
    k = _pair(module.kernel_size)[0]
    s = _pair(module.stride or module.kernel_size)[0]
    return _MaxPool.apply(x, k, s, _pair(module.padding)[0])
