"""tcgen05 GEMM front end: raw ``gemm`` call, ``linear`` autograd function, module dispatch.

The kernel (``csrc/gemm.cu``) multiplies bf16 operands with fp32 accumulation in tensor
memory.  Operand layouts are expressed by *major-ness* so that forward, dgrad and wgrad of a
linear / 1x1-convolution layer all run without materialising a transpose:

    forward  y  = x  w^T + b : gemm(x,  w)                      (both K-major)
    dgrad    dx = dy w       : gemm(dy, w,  b_mn=True)
    wgrad    dw = dy^T x     : gemm(dy, x,  a_mn=True, b_mn=True)   (split-K over the batch)

Reference ops replaced: ``nn.Linear`` / 1x1 ``nn.Conv2d`` through cuBLAS / cuDNN
(CIFAR10/alexnet.py:31-36, CIFAR10/vgg16.py:24-30, IMAGENET/training/resnet.py:55-69,:110).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _native
from .._native import abi

_WS = {}


def _workspace(nbytes: int, device) -> torch.Tensor:
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() * 4 < nbytes:
        ws = torch.empty(max(nbytes // 4, 1 << 20), dtype=torch.float32, device=device)
        _WS[key] = ws
    return ws


def supported(*tensors) -> bool:
    for t in tensors:
        if t.dtype != torch.bfloat16 or not t.is_cuda or t.dim() != 2 or t.stride(1) != 1:
            return False
        if t.stride(0) % 8 or t.data_ptr() % 16:
            return False
    return True


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, out_dtype=torch.bfloat16, bias: Optional[torch.Tensor] = None,
         relu: bool = False, accumulate: bool = False, splits: Optional[int] = None) -> torch.Tensor:
    """``C[M,N] (+)= A B^T`` on the tensor cores.

    ``a`` is ``[M,K]`` (or stored ``[K,M]`` when ``a_mn``); ``b`` is ``[N,K]`` (or stored ``[K,N]``
    when ``b_mn``).  Rows may be strided (``stride(0) % 8 == 0``), the inner dim must be dense.
    """
    lib = _native.lib()
    if not supported(a, b):
        raise ValueError("gemm: operands must be 2-D bf16 CUDA tensors with dense rows, 16-byte aligned")
    K, M = (a.shape if a_mn else a.shape[::-1])
    Kb, N = (b.shape if b_mn else b.shape[::-1])
    if K != Kb:
        raise ValueError(f"gemm: reduction dims differ ({K} vs {Kb})")
    if out is None:
        out = torch.empty(M, N, dtype=out_dtype, device=a.device)
    if splits is None:
        splits = int(lib.lwc_gemm_pick_splits(M, N, K))
    ws_ptr = 0
    if splits > 1:
        ws = _workspace(int(lib.lwc_gemm_workspace_bytes(M, N, splits)), a.device)
        ws_ptr = ws.data_ptr()
    if bias is not None:
        bias = bias.float().contiguous()
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.lwc_gemm_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, a.stride(0), b.stride(0),
                           out.stride(0), int(a_mn), int(b_mn), abi.dtype_code(out.dtype),
                           0 if bias is None else bias.data_ptr(), int(relu), int(accumulate), splits, ws_ptr,
                           C.c_void_p(st))
    abi.check(rc, "lwc_gemm_bf16")
    _native.counters.add(2 if splits > 1 else 1)
    return out


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias):
        ctx.save_for_backward(x, w)
        ctx.has_bias = bias is not None
        return gemm(x, w, bias=bias)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy, w, b_mn=True)
        if ctx.needs_input_grad[1]:
            dw = gemm(dy, x, a_mn=True, b_mn=True)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy.float().sum(0).to(dy.dtype)
        return dx, dw, db


def linear(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``x @ w.T + bias`` for 2-D bf16 ``x`` with autograd through the tcgen05 kernel."""
    return _LinearFn.apply(x, w, bias)


def maybe_linear(module, x):
    """Module dispatch: native path for bf16 inputs whose shapes satisfy the TMA constraints."""
    w = module.weight
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or x.dim() != 2:
        return None
    if not x.is_contiguous():
        x = x.contiguous()
    if not supported(x, w) or w.shape[0] % 8:
        return None
    return linear(x, w, module.bias)
