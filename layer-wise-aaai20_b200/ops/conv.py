"""Convolution front end for the tcgen05 convolution kernels (``csrc/convx.cu``, ``csrc/conv.cu``).

Every convolution of the model zoo runs on hand-written sm_100a kernels: forward (optionally with bias + ReLU in
the epilogue, optionally emitting batch-norm statistics so that the following batch norm skips its statistics
pass), data gradient and weight gradient, for square kernels with ``k * k <= 16``, stride 1 or 2, channels-last
bf16 activations, ``Cout % 64 == 0`` and any ``Cin`` (3-channel inputs are stored with 8 channels; the padding to
the 64-wide K chunk happens in the TMA descriptor).  The ResNet stem runs on the same kernel through an
overlapping-window tensor map over the space-to-depth input (``ops/input.py``).

The output-tile width (64 / 128 / 256 channels) is chosen per layer shape by timing the candidates once, the way
the reference relies on ``cudnn.benchmark = True`` (CIFAR10/torch_backend.py:7, train_imagenet_nv.py:95).

Reference ops: ``nn.Conv2d`` forward / backward through cuDNN (IMAGENET/training/resnet.py:18-21,59-66,100-101,
122-126; CIFAR10/dawn.py:23-35; CIFAR10/alexnet.py:15-27; CIFAR10/vgg16.py:69-82).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

from .. import _native
from .._native import abi


def _nhwc(x: torch.Tensor) -> bool:
    return x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# ---- per-shape choice of the output-tile width -----------------------------------------------------------------
_TUNE: Dict[tuple, int] = {}
_TUNE_ON = {"on": os.environ.get("LWC_CONV_TUNE", "1") != "0"}


def autotune(flag: bool = True):
    """Time the tile-width candidates on the first call of every new layer shape (default on)."""
    _TUNE_ON["on"] = bool(flag)


def _candidates(cout: int):
    return [b for b in (64, 128, 256) if cout % b == 0] or [64]


_HALO = {"on": os.environ.get("LWC_CONV_HALO", "1") != "0"}


def _halo_candidates(kdim: int, nout: int, k: int, stride: int):
    """Extra tile hints (1000 + width) that select the halo variant of the kernel: K dimension of one 64-channel chunk,
    stride 1, several taps -- resident weights and one input box per tile instead of one per tap."""
    if not _HALO["on"] or kdim > 64 or k < 2 or k * k > 16 or stride != 1:
        return []
    return [1000 + b for b in (64, 128) if nout % b == 0]


_TWO = {"on": os.environ.get("LWC_CONV_TWO", "1") != "0"}


def _two_candidates(nout: int, aux: bool):
    """Tile hints 2000 + width: the variant compiled for two resident CTAs per SM (64- and 128-wide tiles; 64 only for the
    launches that stage a second tile per sub-tile: batch-norm backward sums, residual fork)."""
    if not _TWO["on"]:
        return []
    return [2000 + b for b in ((64,) if aux else (64, 128)) if nout % b == 0]


def _tuned(key: tuple, cout: int, run, cands=None) -> int:
    """``run(bn_hint)`` launches the kernel; returns the fastest hint for ``key`` (cached)."""
    hint = _TUNE.get(key)
    if hint is not None:
        return hint
    cands = cands or _candidates(cout)
    if not _TUNE_ON["on"] or len(cands) == 1 or torch.cuda.is_current_stream_capturing():
        return 0
    best, best_t = 0, float("inf")
    for b in cands:
        try:
            run(b)                                  # warm-up (also sets the kernel's shared-memory attribute)
        except RuntimeError:                        # a variant that does not fit this shape
            continue
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            run(b)
        e1.record()
        e1.synchronize()
        t = e0.elapsed_time(e1)
        if t < best_t:
            best, best_t = b, t
    _TUNE[key] = best
    return best


def tuned_table() -> Dict[tuple, int]:
    return dict(_TUNE)


def save_tuned(path: str) -> int:
    """Write the tile choices made so far (layer shape -> hint) as JSON; returns the number of entries.  The choices are
    the counterpart of cuDNN's benchmark cache, which the reference rebuilds on every start (``cudnn.benchmark = True``,
    CIFAR10/dawn.py:120, IMAGENET/training/train_imagenet_nv.py:141): a job that restarts, or that wants every rank to
    run the same variants, loads them instead of timing again."""
    import json
    rows = [[list(k), int(v)] for k, v in _TUNE.items()]
    tmp = f"{path}.tmp{os.getpid()}"
    with open(tmp, "w") as f:
        json.dump({"version": 1, "entries": rows}, f)
    os.replace(tmp, path)
    return len(rows)


def load_tuned(path: str) -> int:
    """Merge tile choices saved by ``save_tuned`` (entries already present win); returns the number of entries added."""
    import json
    with open(path) as f:
        doc = json.load(f)
    if doc.get("version") != 1:
        raise ValueError(f"{path}: unknown tuning-table version {doc.get('version')!r}")
    n = 0
    for k, v in doc["entries"]:
        key = tuple(k)
        if key not in _TUNE:
            _TUNE[key] = int(v)
            n += 1
    return n


def _tune_cache_from_env():
    """``LWC_CONV_TUNE_CACHE=<file>``: load it now if it exists, write it back when the process exits."""
    path = os.environ.get("LWC_CONV_TUNE_CACHE")
    if not path:
        return
    if os.path.exists(path):
        try:
            load_tuned(path)
        except (OSError, ValueError, KeyError, TypeError):
            pass                                   # a stale or truncated file only costs the timing runs
    import atexit
    atexit.register(lambda: _TUNE and save_tuned(path))


_tune_cache_from_env()


# ---- operand preparation --------------------------------------------------------------------------------------
def _prep_input(x: torch.Tensor) -> torch.Tensor:
    """Channels-last with a channel count the TMA unit can address (multiple of 8: 16-byte pixel pitch)."""
    if x.shape[1] % 8:
        x = F.pad(x, (0, 0, 0, 0, 0, 8 - x.shape[1] % 8))
    if not _nhwc(x):
        x = x.contiguous(memory_format=torch.channels_last)
    return x


def _prep_weight(weight: torch.Tensor) -> torch.Tensor:
    """[Cout, Cin, k, k] -> memory [Cout][k*k][Kc] with Kc = round_up(Cin, 64) (zero-padded only when needed)."""
    ci = weight.shape[1]
    w = weight.detach().permute(0, 2, 3, 1)        # a view when the weight is channels-last
    if ci % 64:
        w = F.pad(w, (0, 64 - ci % 64))
    return w.contiguous()


def supported(x: torch.Tensor, weight: torch.Tensor, stride, padding, dilation, groups) -> bool:
    if not (x.is_cuda and x.dim() == 4 and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16):
        return False
    if isinstance(padding, str):
        return False
    co, ci, kh, kw = weight.shape
    s, p = tuple(stride), tuple(padding)
    if kh != kw or kh * kw > 16 or s[0] != s[1] or s[0] not in (1, 2) or tuple(dilation) != (1, 1) or groups != 1:
        return False
    if p[0] != p[1] or co % 64:
        return False
    return x.shape[1] == ci


def out_hw(h: int, w: int, k: int, stride: int, pad: int) -> Tuple[int, int]:
    return (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1


# ---- kernels ----------------------------------------------------------------------------------------------------
class Stats:
    """Per-CTA batch-norm partials written by a forward kernel: ``partials[:nparts]`` = [sum, sum of squares] per
    channel over ``count`` output pixels."""
    __slots__ = ("partials", "nparts", "count")

    def __init__(self, partials, nparts, count):
        self.partials, self.nparts, self.count = partials, nparts, count

    def sums(self) -> torch.Tensor:
        return self.partials[: self.nparts].sum(0)


def conv_forward(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
                 want_stats: bool = False, stride: int = 1, padding: Optional[int] = None, bn_hint: Optional[int] = None):
    """``y = relu?(conv(x, weight) + bias?)``; returns ``(y, Stats | None)``.  ``x`` [N,Cin,H,W] channels-last bf16,
    ``weight`` [Cout,Cin,k,k] bf16."""
    lib = _native.lib()
    co, ci, k, _ = weight.shape
    pad = k // 2 if padding is None else int(padding)
    xin = _prep_input(x.detach())
    n, cx, h, w = xin.shape
    oh, ow = out_hw(h, w, k, stride, pad)
    wt = _prep_weight(weight)
    y = torch.empty((n, co, oh, ow), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    part = None
    if want_stats:
        part = torch.empty((int(lib.lwc_convx_max_partials()), 2, co), dtype=torch.float32, device=x.device)
    bptr = 0
    if bias is not None:
        bias = bias.detach().float().contiguous()
        bptr = bias.data_ptr()
    nparts = C.c_int(0)

    def run(hint):
        rc = lib.lwc_convx_forward(xin.data_ptr(), wt.data_ptr(), y.data_ptr(), n, h, w, cx, co, k, stride, pad, bptr,
                                   int(relu), 0 if part is None else part.data_ptr(), C.byref(nparts), hint, _stream())
        abi.check(rc, "lwc_convx_forward")

    if bn_hint is None:
        bn_hint = _tuned(("f", n, h, w, cx, co, k, stride, pad, want_stats), co, run,
                         _candidates(co) + _halo_candidates(cx, co, k, stride) + _two_candidates(co, False))
    run(bn_hint)
    _native.counters.add(1)
    return y, (Stats(part, int(nparts.value), n * oh * ow) if want_stats else None)


def conv_dgrad(dy: torch.Tensor, weight: torch.Tensor, in_hw: Tuple[int, int], stride: int = 1,
               padding: Optional[int] = None, bn_hint: Optional[int] = None) -> torch.Tensor:
    """Data gradient ``dX`` [N,Cin,H,W] of ``conv(x, weight)`` given ``dY`` (Cin, Cout multiples of 64)."""
    lib = _native.lib()
    co, ci, k, _ = weight.shape
    pad = k // 2 if padding is None else int(padding)
    if not _nhwc(dy):
        dy = dy.contiguous(memory_format=torch.channels_last)
    n = dy.shape[0]
    h, w = in_hw
    dx = torch.empty((n, ci, h, w), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
    if k == 1 and stride == 1 and pad == 0:
        wv = weight.detach().reshape(co, ci).contiguous()

        def run(hint):
            abi.check(lib.lwc_convx_dgrad1x1(dy.data_ptr(), wv.data_ptr(), dx.data_ptr(), n * h * w, co, ci, hint, _stream()),
                      "lwc_convx_dgrad1x1")
        nl = 1
    else:
        wt = weight.detach().permute(0, 2, 3, 1).contiguous()        # [Cout][k][k][Cin]: a view of a channels-last weight

        def run(hint):
            abi.check(lib.lwc_convx_dgrad(dy.data_ptr(), wt.data_ptr(), dx.data_ptr(), n, h, w, ci, co, k, stride, pad, hint,
                                          _stream()), "lwc_convx_dgrad")
        nl = stride * stride
    if bn_hint is None:
        bn_hint = _tuned(("d", n, h, w, ci, co, k, stride, pad), ci, run,
                         _candidates(ci) + _halo_candidates(co, ci, k, stride) + (_two_candidates(ci, False) if stride == 1 else []))
    run(bn_hint)
    _native.counters.add(nl)
    return dx


def conv_dgrad_bnstats(dy: torch.Tensor, weight: torch.Tensor, in_hw: Tuple[int, int], padding: int, link: dict,
                       bn_hint: Optional[int] = None) -> torch.Tensor:
    """Stride-1 data gradient whose output is the gradient of a batch-norm (+ ReLU) output: the same launch sums the
    batch norm's backward reductions (sum d, sum d * x) into per-CTA partials, left in ``link['partials']`` for the batch
    norm's backward (``ops/bn.py``), which then skips its reduction pass over dz and x."""
    lib = _native.lib()
    co, ci, k, _ = weight.shape
    if not _nhwc(dy):
        dy = dy.contiguous(memory_format=torch.channels_last)
    n = dy.shape[0]
    h, w = in_hw
    dx = torch.empty((n, ci, h, w), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
    wt = weight.detach().permute(0, 2, 3, 1).contiguous()
    part = torch.empty((int(lib.lwc_convx_max_partials()), 2, ci), dtype=torch.float32, device=dy.device)
    nparts = C.c_int(0)
    xb, gamma, beta, mean, invstd = link["x"], link["gamma"], link["beta"], link["mean"], link["invstd"]

    def run(hint):
        abi.check(lib.lwc_convx_dgrad_bnstats(dy.data_ptr(), wt.data_ptr(), dx.data_ptr(), n, h, w, ci, co, k, int(padding),
                                              xb.data_ptr(), 0 if gamma is None else gamma.data_ptr(),
                                              0 if beta is None else beta.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                              int(link["relu"]), part.data_ptr(), C.byref(nparts), hint, _stream()),
                  "lwc_convx_dgrad_bnstats")

    if bn_hint is None:
        bn_hint = _tuned(("db", n, h, w, ci, co, k, padding), ci, run,
                         [b for b in (64, 128) if ci % b == 0] + _halo_candidates(co, ci, k, 1) + _two_candidates(ci, True))
    run(bn_hint)
    _native.counters.add(1)
    link["partials"] = (part, int(nparts.value))
    return dx


def bn_link_usable(link, x: torch.Tensor, weight: torch.Tensor, stride: int) -> bool:
    """Can this convolution's data gradient produce the backward sums of the batch norm described by ``link``?"""
    return (link is not None and _ENABLED["on"] and _FUSE_BNB["on"] and stride == 1 and weight.shape[1] % 64 == 0
            and weight.shape[0] % 64 == 0 and link["x"].shape == x.shape and link["x"].dtype == torch.bfloat16 and _nhwc(link["x"]))


_WS: Dict[tuple, torch.Tensor] = {}


def _workspace(nbytes: int, device) -> torch.Tensor:
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() * 4 < nbytes:
        ws = _WS[key] = torch.empty(max(nbytes // 4 + 1, 1 << 20), dtype=torch.float32, device=device)
    return ws


def conv_wgrad(x: torch.Tensor, dy: torch.Tensor, weight_shape, stride: int = 1, padding: Optional[int] = None,
               out_dtype=torch.bfloat16, scale: float = 1.0, bn_hint: Optional[int] = None) -> torch.Tensor:
    """Weight gradient [Cout,Cin,k,k] (channels-last) from the layer input ``x`` and ``dY``."""
    lib = _native.lib()
    co, ci, k, _ = weight_shape
    pad = k // 2 if padding is None else int(padding)
    xin = _prep_input(x.detach())
    if not _nhwc(dy):
        dy = dy.contiguous(memory_format=torch.channels_last)
    n, cx, h, w = xin.shape
    dw = torch.empty((co, k, k, ci), dtype=out_dtype, device=x.device)

    def run(hint):
        nbytes = int(lib.lwc_convx_wgrad_workspace_bytes(n, h, w, cx, co, k, stride, pad, hint))
        if nbytes == 0:
            raise RuntimeError(f"conv_wgrad: unsupported shape {tuple(x.shape)} -> {tuple(weight_shape)}")
        ws = _workspace(nbytes, x.device)
        abi.check(lib.lwc_convx_wgrad(xin.data_ptr(), dy.data_ptr(), dw.data_ptr(), ws.data_ptr(), n, h, w, cx, co, ci, k,
                                      stride, pad, abi.dtype_code(out_dtype), float(scale), hint, _stream()), "lwc_convx_wgrad")

    if bn_hint is None:
        # weight gradient: tile width x accumulator slices per CTA (hint = 1000 * slices + width, see csrc/convx.cu)
        cands = [cap * 1000 + b for b in _candidates(co) for cap in (0, 1, 2)]
        bn_hint = _tuned(("w", n, h, w, cx, co, k, stride, pad), co, run, cands)
    run(bn_hint)
    _native.counters.add(2)
    return dw.permute(0, 3, 1, 2)                     # logical [Cout,Cin,k,k], channels-last memory


class _ConvFn(torch.autograd.Function):
    last_stats = None

    @staticmethod
    def forward(ctx, x, weight, bias, stride, padding, relu, want_stats, bn_link=None):
        y, stats = conv_forward(x, weight, bias, relu=relu, want_stats=want_stats, stride=stride, padding=padding)
        ctx.save_for_backward(x, weight, y if relu else None)
        ctx.conf = (stride, padding, relu, bias is not None)
        ctx.bn_link = bn_link if bn_link_usable(bn_link, x, weight, stride) else None
        _ConvFn.last_stats = stats
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        stride, padding, relu, has_bias = ctx.conf
        if relu:
            dy = dy * (y > 0)
        if not _nhwc(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            if ctx.bn_link is not None:
                dx = conv_dgrad_bnstats(dy, weight, (x.shape[2], x.shape[3]), padding, ctx.bn_link)
            elif weight.shape[1] % 64 == 0:
                dx = conv_dgrad(dy, weight, (x.shape[2], x.shape[3]), stride, padding)
            else:       # not reached by the model zoo (only first layers have such inputs, and they need no dX)
                dx = torch.ops.aten.convolution_backward(dy, x, weight, None, [stride] * 2, [padding] * 2, [1, 1], False,
                                                         [0, 0], 1, [True, False, False])[0]
        if ctx.needs_input_grad[1]:
            dw = conv_wgrad(x, dy, weight.shape, stride, padding, out_dtype=weight.dtype)
        if has_bias and ctx.needs_input_grad[2]:
            db = dy.float().sum((0, 2, 3)).to(dy.dtype)
        return dx, dw, db, None, None, None, None, None


def conv2d(x, weight, bias=None, want_stats: bool = False, stride: int = 1, padding: Optional[int] = None,
           relu: bool = False, bn_link=None):
    """Autograd-aware native convolution; returns ``(y, Stats | None)``.  ``bn_link``: see ``ops.bn`` (the input is the
    output of a batch norm whose backward sums this layer's data gradient can produce)."""
    pad = weight.shape[2] // 2 if padding is None else int(padding)
    _ConvFn.last_stats = None
    y = _ConvFn.apply(x, weight, bias, int(stride), pad, bool(relu), bool(want_stats), bn_link)
    stats, _ConvFn.last_stats = _ConvFn.last_stats, None
    return y, stats


_ENABLED = {"on": os.environ.get("LWC_NATIVE_CONV", "1") != "0"}


def enable(flag: bool = True):
    """Route ``ops.nn.Conv2d`` layers through the native kernels (default on)."""
    _ENABLED["on"] = bool(flag)


def enabled() -> bool:
    return _ENABLED["on"]


_FUSE_STATS = {"on": os.environ.get("LWC_CONV_BN_STATS", "1") != "0"}


def fuse_bn_stats(flag: bool = True):
    """Let convolutions marked ``emit_bn_stats`` sum the following batch norm's statistics in their epilogue."""
    _FUSE_STATS["on"] = bool(flag)


_FUSE_BNB = {"on": os.environ.get("LWC_CONV_BN_BWD", "1") != "0"}


def fuse_bn_backward(flag: bool = True):
    """Let stride-1 data gradients sum the backward reductions of the batch norm that produced their layer's input."""
    _FUSE_BNB["on"] = bool(flag)


def maybe_conv2d(module, x, relu: bool = False):
    """Module dispatch; returns ``None`` when the stock implementation must run."""
    if not _ENABLED["on"]:
        return None
    if not supported(x, module.weight, module.stride, module.padding, module.dilation, module.groups):
        return None
    want = (_FUSE_STATS["on"] and bool(getattr(module, "emit_bn_stats", False)) and module.training
            and torch.is_grad_enabled() and module.bias is None)
    y, stats = conv2d(x, module.weight, module.bias, want_stats=want, stride=module.stride[0], padding=module.padding[0],
                      relu=relu, bn_link=getattr(x, "bn_link", None))
    if stats is not None:
        y.bn_stats = stats
    return y


# ---- residual fork: conv1x1(x) and the identity branch share one fused backward --------------------------------
# Residual fork: the k_fx variant (two CTAs per SM at 56^2 / 28^2 / 14^2) and the round-1 kernel are timed against each
# other per shape.  Also summing bn3's backward reductions in that launch (a third staged tile per sub-tile) measures
# slower than the separate reduction pass it replaces (profiles/halo_probe_r2.json): LWC_FORK_LINK=1 to opt in.
_FORK_NEW = {"on": os.environ.get("LWC_FORK_NEW", "1") != "0"}
_FORK_LINK = {"on": os.environ.get("LWC_FORK_LINK", "0") == "1"}


def conv1x1_dgrad_add(dy: torch.Tensor, weight: torch.Tensor, addend: torch.Tensor,
                      relu_mask: Optional[torch.Tensor] = None, link: Optional[dict] = None,
                      bn_hint: Optional[int] = None) -> torch.Tensor:
    """``dX = dY * W + addend`` in one pass of the native kernel (W = forward weight ``[Cout, Cin, 1, 1]``); with
    ``relu_mask`` (the bit mask a fused batch-norm wrote for the tensor ``X``) also ``dX *= mask``.  With ``link`` (left by
    that batch norm, see ``ops/bn.py``) the same launch sums the batch norm's backward reductions over ``dX`` and its
    input into ``link['partials']``."""
    lib = _native.lib()
    n, co, h, w = dy.shape
    ci = weight.shape[1]
    dx = torch.empty((n, ci, h, w), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
    def run_old():
        if relu_mask is not None:
            abi.check(lib.lwc_conv1x1_add_mask(dy.data_ptr(), weight.data_ptr(), addend.data_ptr(), relu_mask.data_ptr(),
                                               dx.data_ptr(), n, h, w, co, ci, 1, _stream()), "lwc_conv1x1_add_mask")
        else:
            abi.check(lib.lwc_conv1x1_add(dy.data_ptr(), weight.data_ptr(), addend.data_ptr(), dx.data_ptr(), n, h, w, co, ci, 1,
                                          _stream()), "lwc_conv1x1_add")

    if _FORK_NEW["on"] and _ENABLED["on"]:
        wv = weight.detach().reshape(co, ci).contiguous()
        use_link = (link is not None and relu_mask is not None and _FUSE_BNB["on"] and _FORK_LINK["on"]
                    and link["x"].shape == dx.shape)
        part, nparts = None, C.c_int(0)
        if use_link:
            part = torch.empty((int(lib.lwc_convx_max_partials()), 2, ci), dtype=torch.float32, device=dy.device)

        def run(hint):
            if hint < 0:                                # the round-1 kernel, as one more candidate (no sums)
                return run_old()
            abi.check(lib.lwc_convx_dgrad1x1_fork(dy.data_ptr(), wv.data_ptr(), addend.data_ptr(),
                                                  0 if relu_mask is None else relu_mask.data_ptr(), dx.data_ptr(), n * h * w, co, ci,
                                                  link["x"].data_ptr() if use_link else 0,
                                                  link["mean"].data_ptr() if use_link else 0,
                                                  link["invstd"].data_ptr() if use_link else 0,
                                                  0 if part is None else part.data_ptr(), C.byref(nparts), hint, _stream()),
                      "lwc_convx_dgrad1x1_fork")

        if bn_hint is None:
            bn_hint = _tuned(("k", n, h, w, ci, co, relu_mask is not None, use_link), ci, run,
                             [b for b in (64, 128) if ci % b == 0]
                             + ([] if use_link else _two_candidates(ci, True) + ([-1] if ci % 128 == 0 or relu_mask is None else [])))
            if bn_hint == 0 and not _TUNE_ON["on"]:
                bn_hint = 2064 if (n * h * w >= 100000 and not use_link) else 128 if ci % 128 == 0 else 64
        run(bn_hint)
        _native.counters.add(1)
        if use_link:
            link["partials"] = (part, int(nparts.value))
        return dx
    run_old()
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
    def forward(ctx, x, weight, relu_mask=None, want_stats=False, link=None):
        ctx.save_for_backward(x, weight, relu_mask)
        ctx.link = link
        if _ENABLED["on"]:
            y, stats = conv_forward(x, weight, want_stats=want_stats, stride=1, padding=0)
            _ConvFn.last_stats = stats
        else:
            y = F.conv2d(x, weight)
        return y, x.view_as(x)

    @staticmethod
    def backward(ctx, dy, d_identity):
        x, weight, relu_mask = ctx.saved_tensors
        dx = dw = None
        if not _nhwc(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        if ctx.needs_input_grad[0]:
            if d_identity is not None and _nhwc(d_identity) and d_identity.dtype == dy.dtype:
                dx = conv1x1_dgrad_add(dy, weight, d_identity, relu_mask, ctx.link)
            else:
                dx = conv_dgrad(dy, weight, (x.shape[2], x.shape[3]), 1, 0)
                if d_identity is not None:
                    dx = dx + d_identity
                if relu_mask is not None:           # this function promised the producer a masked gradient
                    dx = apply_bitmask(dx, relu_mask)
        if ctx.needs_input_grad[1]:
            if _ENABLED["on"]:
                dw = conv_wgrad(x, dy, weight.shape, 1, 0, out_dtype=weight.dtype)
            else:
                dw = torch.ops.aten.convolution_backward(dy, x, weight, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1,
                                                         [False, True, False])[1]
        return dx, dw, None, None, None


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
    want = _FUSE_STATS["on"] and bool(getattr(module, "emit_bn_stats", False)) and module.training and _ENABLED["on"]
    mask = getattr(x, "relu_mask", None)
    taken = getattr(x, "relu_mask_taken", None)
    _ConvFn.last_stats = None
    if mask is not None and taken is not None and w.shape[1] % 128 == 0 and not os.environ.get("LWC_NO_PREMASK"):
        taken[0] = True
        y, ident = _ForkConv1x1.apply(x, w, mask, want, getattr(x, "relu_mask_link", None))
    else:
        y, ident = _ForkConv1x1.apply(x, w, None, want, None)
    stats, _ConvFn.last_stats = _ConvFn.last_stats, None
    if stats is not None:
        y.bn_stats = stats
    return y, ident
