"""Input normalisation: uint8 NCHW batch -> normalised channels-last activations in one kernel, optionally
in 2x2 space-to-depth form for the 7x7/2 ResNet stem (``csrc/input.cu``).

Reference op: ``BatchTransformDataLoader.process_tensors`` (IMAGENET/training/dataloader.py:88-93).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn.functional as F

from .. import _native
from .._native import abi

IMAGENET_MEAN = (0.485 * 255, 0.456 * 255, 0.406 * 255)
IMAGENET_STD = (0.229 * 255, 0.224 * 255, 0.225 * 255)


def space_to_depth_reference(xn: torch.Tensor) -> torch.Tensor:
    """Normalised NCHW [N,3,H,W] -> [N,16,H/2+3,W/2+3] (12 used channels, 2/1 zero border); pure torch."""
    n, c, h, w = xn.shape
    x = xn.reshape(n, c, h // 2, 2, w // 2, 2).permute(0, 3, 5, 1, 2, 4).reshape(n, 4 * c, h // 2, w // 2)
    x = F.pad(x, (2, 1, 2, 1, 0, 16 - 4 * c))
    return x


def s2d_stem_weight(w7: torch.Tensor) -> torch.Tensor:
    """[Co,3,7,7] stride-2 pad-3 kernel -> the equivalent [Co,16,4,4] stride-1 kernel on the s2d input."""
    co = w7.shape[0]
    w8 = F.pad(w7, (1, 0, 1, 0))                                   # zero row / column in front
    w = w8.reshape(co, 3, 4, 2, 4, 2).permute(0, 3, 5, 1, 2, 4).reshape(co, 12, 4, 4)
    return F.pad(w, (0, 0, 0, 0, 0, 4))


_WS = {}


def stem_wgrad_native_ok(oh: int, ow: int, dtype, cout: int = 64) -> bool:
    """Can ``csrc/stem.cu`` compute the stem's weight gradient for this output size?"""
    return (_native.available() and dtype == torch.bfloat16 and cout == 64 and oh % 2 == 0 and ow % 16 == 0
            and 16 <= ow <= 144)


def stem_wgrad(dy: torch.Tensor, xs: torch.Tensor, out_dtype=torch.bfloat16, scale: float = 1.0) -> torch.Tensor:
    """Weight gradient [64,3,7,7] (channels-last) of the 7x7/2 stem from ``dy`` [N,64,OH,OW] (channels-last bf16) and
    the space-to-depth input ``xs`` [N,16,OH+3,OW+3] (channels-last bf16): one tcgen05 kernel + a tiny reduction."""
    n, co, oh, ow = dy.shape
    assert co == 64 and tuple(xs.shape) == (n, 16, oh + 3, ow + 3)
    assert dy.dtype == torch.bfloat16 and xs.dtype == torch.bfloat16
    if not dy.is_contiguous(memory_format=torch.channels_last):
        dy = dy.contiguous(memory_format=torch.channels_last)
    assert xs.is_contiguous(memory_format=torch.channels_last)
    lib = _native.lib()
    nbytes = int(lib.lwc_stem_wgrad_workspace_bytes(n, oh))
    key = (dy.device.index, torch.cuda.current_stream(dy.device).cuda_stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() * 4 < nbytes:
        ws = _WS[key] = torch.empty(nbytes // 4, dtype=torch.float32, device=dy.device)
    gw = torch.empty((64, 3, 7, 7), dtype=out_dtype, device=dy.device, memory_format=torch.channels_last)
    st = torch.cuda.current_stream().cuda_stream
    abi.check(lib.lwc_stem_wgrad(dy.data_ptr(), xs.data_ptr(), gw.data_ptr(), ws.data_ptr(), n, oh, ow,
                                 abi.dtype_code(out_dtype), float(scale), C.c_void_p(st)), "lwc_stem_wgrad")
    _native.counters.add(2)
    return gw


class _StemS2D(torch.autograd.Function):
    """Stem convolution: forward on the space-to-depth input (dense 4x4/1, K = 256).  Weight gradient: the tcgen05
    kernel of ``csrc/stem.cu`` straight from the space-to-depth input, or (``x3`` given: shapes the kernel does not
    take) cuDNN's wgrad in the original 7x7/2 formulation on a plain 3-channel copy -- its wgrad for the 16-channel
    4x4 form is ~2x slower than its 7x7 one on B200.  No input gradient: the stem consumes data."""

    @staticmethod
    def forward(ctx, xs, x3, w7, want_stats=False):
        ctx.native = x3 is None
        ctx.save_for_backward(xs if x3 is None else x3, w7)
        return stem_forward(xs, w7, want_stats)

    @staticmethod
    def backward(ctx, dy):
        x, w7 = ctx.saved_tensors
        if ctx.native:
            return None, None, stem_wgrad(dy, x, w7.dtype), None
        gw = torch.ops.aten.convolution_backward(dy, x, w7, None, [2, 2], [3, 3], [1, 1], False, [0, 0], 1,
                                                 [False, True, False])[1]
        return None, None, gw, None


_LAST_STATS = {"v": None}


def stem_forward_native_ok(xs: torch.Tensor, w7: torch.Tensor) -> bool:
    from . import conv as _conv
    return (_conv.enabled() and xs.is_cuda and _native.available() and xs.dtype == torch.bfloat16
            and w7.dtype == torch.bfloat16 and tuple(w7.shape[1:]) == (3, 7, 7) and w7.shape[0] % 64 == 0
            and xs.shape[1] == 16 and xs.is_contiguous(memory_format=torch.channels_last))


def stem_forward(xs: torch.Tensor, w7: torch.Tensor, want_stats: bool = False, bn_hint=None) -> torch.Tensor:
    """The stem convolution on the space-to-depth input: the tap-table tcgen05 kernel (four kernel-row taps over
    overlapping 128-byte windows, K = 256) or, off the GPU, the equivalent dense 4x4 convolution."""
    w = s2d_stem_weight(w7.detach())
    _LAST_STATS["v"] = None
    if not stem_forward_native_ok(xs, w7):
        if xs.is_contiguous(memory_format=torch.channels_last):
            w = w.contiguous(memory_format=torch.channels_last)
        return F.conv2d(xs, w, None, stride=1, padding=0)
    from .conv import Stats
    lib = _native.lib()
    n, _, hs, ws_ = xs.shape
    oh, ow, co = hs - 3, ws_ - 3, w7.shape[0]
    w4 = w.permute(0, 2, 3, 1).contiguous()                       # [Cout][4 rows][4 cols][16 ch]
    y = torch.empty((n, co, oh, ow), dtype=xs.dtype, device=xs.device, memory_format=torch.channels_last)
    part = None
    if want_stats:
        part = torch.empty((int(lib.lwc_convx_max_partials()), 2, co), dtype=torch.float32, device=xs.device)
    nparts = C.c_int(0)
    st = torch.cuda.current_stream().cuda_stream

    def run(hint):
        abi.check(lib.lwc_convx_stem_forward(xs.data_ptr(), w4.data_ptr(), y.data_ptr(), n, oh, ow, co,
                                             0 if part is None else part.data_ptr(), C.byref(nparts), hint, C.c_void_p(st)),
                  "lwc_convx_stem_forward")

    # per-tap boxes, or the halo variant: the four kernel-row taps read one 16 x 11-window box, weights resident
    from . import conv as _conv
    hint = bn_hint
    if hint is None:
        hint = _conv._tuned(("stem", n, oh, ow, co, want_stats), co, run, [64, 128, 1064, 1128, 3064] if co % 128 == 0 else [64, 1064, 3064])
    run(hint)
    _native.counters.add(1)
    if want_stats:
        _LAST_STATS["v"] = Stats(part, int(nparts.value), n * oh * ow)
    return y


def stem_s2d(xs: torch.Tensor, w7: torch.Tensor, want_stats: bool = False) -> torch.Tensor:
    x3 = getattr(xs, "plain3", None)
    y = None
    if w7.requires_grad and torch.is_grad_enabled():
        if (xs.is_cuda and xs.dtype == torch.bfloat16 and w7.dtype == torch.bfloat16 and tuple(w7.shape) == (64, 3, 7, 7)
                and xs.shape[1] == 16 and xs.is_contiguous(memory_format=torch.channels_last)
                and stem_wgrad_native_ok(xs.shape[2] - 3, xs.shape[3] - 3, xs.dtype, w7.shape[0])):
            y = _StemS2D.apply(xs, None, w7, want_stats)
        elif x3 is not None:
            y = _StemS2D.apply(xs, x3, w7, want_stats)
    if y is None:
        y = stem_forward(xs, w7, want_stats)
    stats, _LAST_STATS["v"] = _LAST_STATS["v"], None
    if stats is not None:
        y.bn_stats = stats
    return y


def normalize_u8(x: torch.Tensor, dtype=torch.bfloat16, s2d: bool = False, mean=IMAGENET_MEAN, std=IMAGENET_STD,
                 keep_plain: bool = False):
    """uint8 NCHW -> normalised tensor.  CUDA: one fused kernel; returns a channels-last tensor
    ([N,8,H,W] with 5 zero channels dropped by a view -> [N,3,H,W] is not possible, so the plain mode returns
    a dense channels-last [N,3,H,W] via torch ops when ``s2d`` is False)."""
    if s2d and x.is_cuda and _native.available() and x.is_contiguous() and x.dtype == torch.uint8:
        n, c, h, w = x.shape
        assert c == 3 and h % 2 == 0 and w % 2 == 0
        out = torch.empty((n, h // 2 + 3, w // 2 + 3, 16), dtype=dtype, device=x.device)
        m = (C.c_float * 3)(*mean)
        s = (C.c_float * 3)(*std)
        st = torch.cuda.current_stream().cuda_stream
        abi.check(_native.lib().lwc_normalize_u8(x.data_ptr(), out.data_ptr(), n, h, w, 16, 1, m, s,
                                                 abi.dtype_code(dtype), C.c_void_p(st)), "lwc_normalize_u8")
        _native.counters.add(1)
        res = out.permute(0, 3, 1, 2)           # logical NCHW view of NHWC storage (= channels_last)
        if keep_plain and not stem_wgrad_native_ok(h // 2, w // 2, dtype):
            # shapes csrc/stem.cu does not take: cuDNN's weight gradient is cheaper in the original 7x7 formulation,
            # which needs a plain 3-channel copy
            plain = torch.empty((n, h, w, 3), dtype=dtype, device=x.device)
            abi.check(_native.lib().lwc_normalize_u8(x.data_ptr(), plain.data_ptr(), n, h, w, 3, 0, m, s,
                                                     abi.dtype_code(dtype), C.c_void_p(st)), "lwc_normalize_u8")
            _native.counters.add(1)
            res.plain3 = plain.permute(0, 3, 1, 2)
        return res
    mean_t = torch.tensor(mean, device=x.device).view(1, 3, 1, 1)
    std_t = torch.tensor(std, device=x.device).view(1, 3, 1, 1)
    xn = (x.float() - mean_t) / std_t
    if s2d:
        return space_to_depth_reference(xn).to(dtype).contiguous(memory_format=torch.channels_last)
    return xn.to(dtype)
