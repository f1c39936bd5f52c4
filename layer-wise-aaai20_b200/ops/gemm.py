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


# ---- Linear layers: per shape, the faster of the GEMM kernel and the 1x1 path of the convolution kernel -----------------
# A linear layer is a 1x1 convolution over M "pixels".  For the skinny GEMMs of the classifiers (M = 512 rows against
# 4096 x 25088 weights) the persistent tap-table kernel (csrc/convx.cu: 256-wide tiles, 8 epilogue warps) is ahead of
# csrc/gemm.cu for forward and data gradient, the split-K MN-major GEMM is ahead for the weight gradient
# (profiles/linear_bench_r2.json); the choice is timed once per shape like the convolution tile widths.
_LIN_TUNE = {}


def _as_pixels(t: torch.Tensor) -> torch.Tensor:
    return t.view(t.shape[0], t.shape[1], 1, 1)           # [M, C, 1, 1]: channels-last and contiguous at once


def _pick(key, cands):
    """cands: {name: thunk}; returns the cached / fastest name (first one while a CUDA graph is being captured)."""
    name = _LIN_TUNE.get(key)
    if name is not None:
        return name
    names = list(cands)
    if len(names) == 1 or torch.cuda.is_current_stream_capturing():
        return names[0]
    best, best_t = names[0], float("inf")
    for n in names:
        cands[n]()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            cands[n]()
        e1.record()
        e1.synchronize()
        t = e0.elapsed_time(e1)
        if t < best_t:
            best, best_t = n, t
    _LIN_TUNE[key] = best
    return best


def linear_forward(x, w, bias=None, relu: bool = False):
    from . import conv as _conv
    M, K = x.shape
    N = w.shape[0]
    cands = {"gemm": lambda: gemm(x, w, bias=bias, relu=relu)}
    if N % 64 == 0 and K % 8 == 0 and _conv.enabled():
        cands["convx"] = lambda: _conv.conv_forward(_as_pixels(x), _as_pixels(w), bias, relu=relu, stride=1, padding=0)[0].view(M, N)
    return cands[_pick(("f", M, K, N, bias is not None, relu), cands)]()


def linear_dgrad(dy, w):
    from . import conv as _conv
    M, N = dy.shape
    K = w.shape[1]
    cands = {"gemm": lambda: gemm(dy, w, b_mn=True)}
    if N % 64 == 0 and K % 64 == 0 and _conv.enabled():
        cands["convx"] = lambda: _conv.conv_dgrad(_as_pixels(dy), _as_pixels(w), (1, 1), 1, 0).view(M, K)
    return cands[_pick(("d", M, K, N), cands)]()


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, relu=False):
        y = linear_forward(x, w, bias, relu)
        ctx.save_for_backward(x, w, y if relu else None)
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        dy = dy.contiguous()
        if y is not None:
            dy = dy * (y > 0)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = linear_dgrad(dy, w)
        if ctx.needs_input_grad[1]:
            dw = gemm(dy, x, a_mn=True, b_mn=True)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy.float().sum(0).to(dy.dtype)
        return dx, dw, db, None


def linear(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False) -> torch.Tensor:
    """``relu?(x @ w.T + bias)`` for 2-D bf16 ``x`` with autograd through the tcgen05 kernels."""
    return _LinearFn.apply(x, w, bias, relu)


def maybe_linear(module, x):
    """Module dispatch: native path for bf16 inputs whose shapes satisfy the TMA constraints."""
    w = module.weight
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or x.dim() != 2:
        return None
    if not x.is_contiguous():
        x = x.contiguous()
    if not supported(x, w):
        return None
    n = w.shape[0]
    relu = bool(getattr(module, "fuse_relu", False))
    if n % 8:
        # class heads (10 outputs): zero rows pad the weight to one 64-wide tile, the extra outputs are sliced away;
        # the padding and the slice are differentiable, so the gradients come back in the parameter's own shape
        pad = 64 - n % 64
        wp = torch.nn.functional.pad(w, (0, 0, 0, pad))
        bp = None if module.bias is None else torch.nn.functional.pad(module.bias, (0, pad))
        return linear(x, wp, bp, relu)[:, :n]
    return linear(x, w, module.bias, relu)
