"""Convolution front end for the persistent tcgen05 implicit-GEMM kernel (``csrc/conv.cu``).

Covers stride-1 "same" convolutions with 1x1 and 3x3 kernels on channels-last bf16 activations whose channel
counts are multiples of 64: forward (optionally emitting batch-norm statistics partials from the epilogue) and the
data gradient (a convolution with the spatially flipped, channel-transposed kernel).  Weight gradients of 1x1
layers go through the split-K MN-major GEMM (``ops/gemm.py``); everything else falls back to the stock op.

Reference ops: ``nn.Conv2d`` forward / backward through cuDNN (IMAGENET/training/resnet.py:21-24,:55-69,
CIFAR10/dawn.py:23-35).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import torch

from .. import _native
from .._native import abi


def _nhwc(x: torch.Tensor) -> bool:
    return x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last)


def supported(x: torch.Tensor, weight: torch.Tensor, stride, padding, dilation, groups) -> bool:
    if not (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and _nhwc(x)):
        return False
    co, ci, kh, kw = weight.shape
    if kh != kw or kh not in (1, 3) or tuple(stride) != (1, 1) or tuple(dilation) != (1, 1) or groups != 1:
        return False
    if tuple(padding) != (kh // 2, kh // 2) or ci % 64 or co % 64:
        return False
    return weight.is_contiguous(memory_format=torch.channels_last) or kh == 1


def tiles_m(n: int, h: int, w: int, kh: int) -> Tuple[int, int]:
    rows = C.c_int(0)
    t = int(_native.lib().lwc_conv_tiles_m(n, h, w, kh, C.byref(rows)))
    return t, int(rows.value)


def conv_forward(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
                 want_stats: bool = False, b_mn: bool = False, out_channels: Optional[int] = None):
    """``y = conv(x, weight)``; returns ``(y, stats)`` with ``stats`` = ``[tiles_m, 2, Cout]`` fp32 or ``None``.

    ``weight`` is the channels-last forward weight ``[Cout, Cin, k, k]`` (memory order Cout, k, k, Cin).  With
    ``b_mn`` (1x1 only) ``weight`` is used as stored ``[K][N]``, which turns the call into the data-gradient GEMM.
    """
    lib = _native.lib()
    n, ci, h, w = x.shape
    kh = weight.shape[2]
    co = out_channels if out_channels is not None else weight.shape[0]
    y = torch.empty((n, co, h, w), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    stats = None
    if want_stats:
        tm, _ = tiles_m(n, h, w, kh)
        stats = torch.empty((tm, 2, co), dtype=torch.float32, device=x.device)
    if bias is not None:
        bias = bias.float().contiguous()
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.lwc_conv_forward(x.data_ptr(), weight.data_ptr(), y.data_ptr(), n, h, w, ci, co, kh, int(b_mn),
                              0 if bias is None else bias.data_ptr(), int(relu), 0 if stats is None else stats.data_ptr(),
                              C.c_void_p(st))
    abi.check(rc, "lwc_conv_forward")
    _native.counters.add(1)
    return y, stats


def flipped_transposed(weight: torch.Tensor) -> torch.Tensor:
    """Kernel of the data-gradient convolution: ``Wd[ci, co, a, b] = W[co, ci, k-1-a, k-1-b]`` (channels-last)."""
    return weight.flip(2, 3).transpose(0, 1).contiguous(memory_format=torch.channels_last)


class _ConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, want_stats):
        y, stats = conv_forward(x, weight, bias, want_stats=want_stats)
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        ctx.mark_non_differentiable(*([stats] if stats is not None else []))
        return (y, stats) if stats is not None else (y, None)

    @staticmethod
    def backward(ctx, dy, _dstats=None):
        x, weight = ctx.saved_tensors
        if not _nhwc(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        kh = weight.shape[2]
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            if kh == 1:        # dX = dY * W with W used as stored [Cout][Cin] (MN-major B): no transpose pass
                dx, _ = conv_forward(dy, weight, b_mn=True, out_channels=weight.shape[1])
            else:
                dx, _ = conv_forward(dy, flipped_transposed(weight))
        if ctx.needs_input_grad[1]:
            if kh == 1:
                from .gemm import gemm
                n, co, h, w = dy.shape
                dy2 = dy.permute(0, 2, 3, 1).reshape(-1, co)
                x2 = x.permute(0, 2, 3, 1).reshape(-1, x.shape[1])
                dw = gemm(dy2, x2, a_mn=True, b_mn=True).view(co, x.shape[1], 1, 1).contiguous(memory_format=torch.channels_last)
            else:
                dw = torch.ops.aten.convolution_backward(dy, x, weight, None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1,
                                                         [False, True, False])[1]
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy.float().sum((0, 2, 3)).to(dy.dtype)
        return dx, dw, db, None


def conv2d(x, weight, bias=None, want_stats: bool = False):
    """Autograd-aware native convolution; returns ``(y, stats_or_None)``."""
    return _ConvFn.apply(x, weight, bias, want_stats)


_ENABLED = {"on": False}


def enable(flag: bool = True):
    """Route eligible ``ops.nn.Conv2d`` layers through the native kernel (off by default: see DESIGN.md)."""
    _ENABLED["on"] = bool(flag)


def maybe_conv2d(module, x):
    """Module dispatch; returns ``None`` when the stock implementation must run."""
    if not _ENABLED["on"]:
        return None
    if not supported(x, module.weight, module.stride, module.padding, module.dilation, module.groups):
        return None
    y, stats = conv2d(x, module.weight, module.bias, want_stats=getattr(module, "emit_bn_stats", False))
    if stats is not None:
        y.bn_stats = stats
    return y


# ---- residual fork: conv1x1(x) and the identity branch share one fused backward --------------------------------
def conv1x1_dgrad_add(dy: torch.Tensor, weight: torch.Tensor, addend: torch.Tensor,
                      relu_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``dX = dY * W + addend`` in one pass of the native kernel (W = forward weight ``[Cout, Cin, 1, 1]``); with
    ``relu_mask`` (the bit mask a fused batch-norm wrote for the tensor ``X``) also ``dX *= mask``."""
    lib = _native.lib()
    n, co, h, w = dy.shape
    ci = weight.shape[1]
    dx = torch.empty((n, ci, h, w), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
    st = torch.cuda.current_stream().cuda_stream
    if relu_mask is not None:
        rc = lib.lwc_conv1x1_add_mask(dy.data_ptr(), weight.data_ptr(), addend.data_ptr(), relu_mask.data_ptr(), dx.data_ptr(),
                                      n, h, w, co, ci, 1, C.c_void_p(st))
        abi.check(rc, "lwc_conv1x1_add_mask")
    else:
        rc = lib.lwc_conv1x1_add(dy.data_ptr(), weight.data_ptr(), addend.data_ptr(), dx.data_ptr(), n, h, w, co, ci, 1,
                                 C.c_void_p(st))
        abi.check(rc, "lwc_conv1x1_add")
    _native.counters.add(1)
    return dx


def apply_bitmask(t: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """``t * mask`` for a channels-last ``t`` [N,C,H,W] and a bit mask [N*H*W*C/8] bytes (bit j of byte b <-> channel
    8 b + j of the same pixel): the plain-op counterpart of the kernel epilogue."""
    n, c, h, w = t.shape
    bits = torch.stack([(mask >> k) & 1 for k in range(8)], dim=-1).reshape(n, h, w, c).permute(0, 3, 1, 2)
    return t * bits.to(t.dtype)


class _ForkConv1x1(torch.autograd.Function):
    """``(conv1x1(x, w), x)``: the block input feeds both the first 1x1 convolution and the skip connection.  In the
    backward pass the two gradients w.r.t. ``x`` would be summed by a separate elementwise kernel; here the data
    gradient runs on the tcgen05 kernel whose epilogue adds the skip-connection gradient on the way out."""

    @staticmethod
    def forward(ctx, x, weight, relu_mask=None):
        ctx.save_for_backward(x, weight, relu_mask)
        y = torch.nn.functional.conv2d(x, weight)
        return y, x.view_as(x)

    @staticmethod
    def backward(ctx, dy, d_identity):
        x, weight, relu_mask = ctx.saved_tensors
        dx = dw = None
        if not _nhwc(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        if ctx.needs_input_grad[0]:
            if d_identity is not None and _nhwc(d_identity) and d_identity.dtype == dy.dtype:
                dx = conv1x1_dgrad_add(dy, weight, d_identity, relu_mask)
            else:
                dx, _ = conv_forward(dy, weight, b_mn=True, out_channels=weight.shape[1])
                if d_identity is not None:
                    dx = dx + d_identity
                if relu_mask is not None:           # this function promised the producer a masked gradient
                    dx = apply_bitmask(dx, relu_mask)
        if ctx.needs_input_grad[1]:
            dw = torch.ops.aten.convolution_backward(dy, x, weight, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1,
                                                     [False, True, False])[1]
        return dx, dw, None


def fork_conv1x1(module, x):
    """``(module(x), x)`` with the fused backward when the layer qualifies; plain ops otherwise.

    If ``x`` is the output of a fused batch-norm + residual + ReLU (it then carries ``relu_mask`` / ``relu_mask_taken``,
    see ``ops/bn.py``) and ``x`` has no consumer other than this fork, the fork also multiplies its input gradient with
    that ReLU mask in the kernel epilogue and tells the batch-norm so: its backward then neither applies the mask nor
    writes a separate gradient for the residual branch (it hands the incoming gradient on as is)."""
    w = module.weight
    ok = (x.is_cuda and _native.available() and torch.is_grad_enabled() and x.requires_grad and module.bias is None
          and tuple(module.kernel_size) == (1, 1) and tuple(module.stride) == (1, 1) and tuple(module.padding) == (0, 0)
          and module.groups == 1 and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and _nhwc(x)
          and w.shape[0] % 64 == 0 and w.shape[1] % 64 == 0)
    if not ok:
        return module(x), x
    mask = getattr(x, "relu_mask", None)
    taken = getattr(x, "relu_mask_taken", None)
    if mask is not None and taken is not None and w.shape[1] % 128 == 0 and not os.environ.get("LWC_NO_PREMASK"):
        taken[0] = True
        return _ForkConv1x1.apply(x, w, mask)
    return _ForkConv1x1.apply(x, w)
