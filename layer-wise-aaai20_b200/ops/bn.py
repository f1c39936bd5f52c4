"""Fused batch-norm (+ residual add) (+ ReLU) for channels-last activations.

``batch_norm_act`` is an autograd function over ``csrc/bn.cu``: training forward = statistics
+ finalize + apply (3 launches), backward = reduce + finalize + apply (3 launches), with the
ReLU and the residual join folded in.  The stock modules it replaces run separate statistics /
transform / ReLU / add kernels (reference: nn.BatchNorm2d + nn.ReLU(True) + ``out += residual``,
CIFAR10/dawn.py:23-28, IMAGENET/training/resnet.py:40-56,:75-95).
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
        ws = torch.empty(max(nbytes // 4 + 1, 1 << 18), dtype=torch.float32, device=device)
        _WS[key] = ws
    return ws


def _rows_channels(x: torch.Tensor):
    """(M, C) when ``x`` is a dense channels-last 4-D tensor or a contiguous 2-D one, else None."""
    if x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last):
        n, c, h, w = x.shape
        return n * h * w, c
    if x.dim() == 2 and x.is_contiguous():
        return x.shape[0], x.shape[1]
    return None


def supported(x: torch.Tensor, weight, bias, residual=None) -> bool:
    if not x.is_cuda or x.dtype not in (torch.bfloat16, torch.float16, torch.float32):
        return False
    mc = _rows_channels(x)
    if mc is None or mc[1] % 8 or mc[0] < 1:
        return False
    for t in (weight, bias):
        if t is not None and (t.dtype != torch.float32 or not t.is_contiguous()):
            return False
    if residual is not None and (residual.dtype != x.dtype or residual.shape != x.shape
                                 or _rows_channels(residual) is None):
        return False
    return True


def _ptr(t):
    return 0 if t is None else t.data_ptr()


class _BatchNormAct(torch.autograd.Function):
    last_mask = (None, None, None)      # (bit mask, taken-flag, link) of the most recent forward, picked up by batch_norm_act

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, residual, training, momentum, eps, relu,
                num_batches_tracked=None, stat_partials=None, stat_nparts=0, stat_count=0, link=None):
        lib = _native.lib()
        M, Cc = _rows_channels(x)
        y = torch.empty_like(x)       # preserves the channels-last strides
        st = torch.cuda.current_stream().cuda_stream
        mean = invstd = None
        ws_ptr = 0
        if training:
            mean = torch.empty(Cc, dtype=torch.float32, device=x.device)
            invstd = torch.empty(Cc, dtype=torch.float32, device=x.device)
            ws_ptr = _workspace(int(lib.lwc_bn_workspace_bytes(M, Cc)), x.device).data_ptr()
        # a fused residual makes the ReLU mask depend on the output: keep it as one bit per element instead of re-reading y
        mask = None
        if training and relu and residual is not None:
            mask = torch.empty(M * Cc // 8, dtype=torch.uint8, device=x.device)
        if training and stat_partials is not None:
            # the producing convolution's epilogue already summed the batch statistics: finalize + apply only
            rc = lib.lwc_convx_stats_finalize(stat_partials.data_ptr(), int(stat_nparts), Cc, int(stat_count), float(eps),
                                              float(momentum), mean.data_ptr(), invstd.data_ptr(), _ptr(running_mean),
                                              _ptr(running_var), _ptr(num_batches_tracked), C.c_void_p(st))
            abi.check(rc, "lwc_convx_stats_finalize")
            rc = lib.lwc_bn_apply_stats(x.data_ptr(), y.data_ptr(), _ptr(residual), _ptr(weight), _ptr(bias), mean.data_ptr(),
                                        invstd.data_ptr(), M, Cc, abi.dtype_code(x.dtype), int(relu), _ptr(mask), C.c_void_p(st))
            abi.check(rc, "lwc_bn_apply_stats")
            _native.counters.add(2)
        else:
            rc = lib.lwc_bn_forward_ex(x.data_ptr(), y.data_ptr(), _ptr(residual), _ptr(weight), _ptr(bias), _ptr(mean),
                                       _ptr(invstd), _ptr(running_mean), _ptr(running_var), ws_ptr, M, Cc,
                                       abi.dtype_code(x.dtype), int(training), int(relu), float(eps), float(momentum),
                                       _ptr(num_batches_tracked), _ptr(mask), C.c_void_p(st))
            abi.check(rc, "lwc_bn_forward_ex")
            _native.counters.add(3 if training else 1)
        ctx.training, ctx.relu, ctx.has_res = training, relu, residual is not None
        ctx.eps = eps
        # the single consumer of y (a stride-1 convolution) may compute this layer's backward reduction sums in its
        # data-gradient epilogue: it finds what it needs in `link` and leaves the partials there (ops/conv.py)
        ctx.link = None
        if link is not None and training and residual is None and x.dim() == 4 and x.dtype == torch.bfloat16:
            link.update(x=x, gamma=weight, beta=bias, mean=mean, invstd=invstd, relu=bool(relu))
            ctx.link = link
        # the ReLU mask is recomputed from x unless a residual was fused (then it depends on y)
        need_y = relu and (residual is not None or not training) and mask is None
        # a consumer (ops.conv.fork_conv1x1) may take over the masking of the incoming gradient: it flips taken[0]
        ctx.mask_taken = [False] if mask is not None else None
        # ... and may then also sum this layer's backward reductions over the gradient it produces (fork_link['partials'])
        ctx.fork_link = None
        if mask is not None and x.dim() == 4 and x.dtype == torch.bfloat16:
            ctx.fork_link = {"x": x, "mean": mean, "invstd": invstd}
        _BatchNormAct.last_mask = (mask, ctx.mask_taken, ctx.fork_link)
        if training:
            ctx.save_for_backward(x, y if need_y else None, weight, mean, invstd, bias, mask)
        else:
            ctx.save_for_backward(x, y if need_y else None, weight, running_mean, running_var, bias, None)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _native.lib()
        x, y, weight, mean, invstd, bias, mask = ctx.saved_tensors
        M, Cc = _rows_channels(x)
        if _rows_channels(dy) is None or dy.stride() != x.stride():
            dy = dy.contiguous(memory_format=torch.channels_last) if x.dim() == 4 else dy.contiguous()
        if not ctx.training:
            # eval-mode backward (rare): plain affine map, do it with tensor ops
            a = (weight if weight is not None else 1.0) * torch.rsqrt(invstd + ctx.eps)   # invstd slot holds running_var
            shape = (1, -1, 1, 1) if x.dim() == 4 else (1, -1)
            g = dy if not ctx.relu else dy * (y > 0)
            dx = (g.float() * a.view(shape)).to(x.dtype)
            xhat = (x.float() - mean.view(shape)) * torch.rsqrt(invstd + ctx.eps).view(shape)
            dims = (0, 2, 3) if x.dim() == 4 else (0,)
            return (dx, (g.float() * xhat).sum(dims), g.float().sum(dims), None, None,
                    g if ctx.has_res else None, None, None, None, None, None, None, None, None, None)
        link = getattr(ctx, "link", None)
        if link is not None and "partials" in link:
            part, nparts = link.pop("partials")
            dx = torch.empty_like(x)
            dgamma = torch.empty(Cc, dtype=torch.float32, device=x.device)
            dbeta = torch.empty(Cc, dtype=torch.float32, device=x.device)
            st = torch.cuda.current_stream().cuda_stream
            rc = lib.lwc_bn_backward_from_partials(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), _ptr(weight), _ptr(bias),
                                                   mean.data_ptr(), invstd.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(),
                                                   part.data_ptr(), int(nparts), M, Cc, abi.dtype_code(x.dtype), int(ctx.relu),
                                                   C.c_void_p(st))
            abi.check(rc, "lwc_bn_backward_from_partials")
            _native.counters.add(2)
            return (dx, dgamma if weight is not None else None, dbeta if weight is not None else None, None, None,
                    None, None, None, None, None, None, None, None, None, None)
        dx = torch.empty_like(x)
        premasked = ctx.mask_taken is not None and ctx.mask_taken[0]
        flink = getattr(ctx, "fork_link", None)
        if premasked and flink is not None and "partials" in flink:
            # the fork's data-gradient kernel produced dy (mask applied) and summed it: finalize + apply only
            part, nparts = flink.pop("partials")
            dgamma = torch.empty(Cc, dtype=torch.float32, device=x.device)
            dbeta = torch.empty(Cc, dtype=torch.float32, device=x.device)
            st = torch.cuda.current_stream().cuda_stream
            rc = lib.lwc_bn_backward_from_partials(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), _ptr(weight), _ptr(bias),
                                                   mean.data_ptr(), invstd.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(),
                                                   part.data_ptr(), int(nparts), M, Cc, abi.dtype_code(x.dtype), 0, C.c_void_p(st))
            abi.check(rc, "lwc_bn_backward_from_partials")
            _native.counters.add(2)
            return (dx, dgamma if weight is not None else None, dbeta if weight is not None else None, None, None,
                    dy if ctx.has_res else None, None, None, None, None, None, None, None, None, None)
        if premasked:
            # the consumer's backward already multiplied dy with this layer's ReLU mask: no mask to read, and the
            # gradient of the residual branch IS dy -- nothing to write
            mask, y, relu_flag, dres = None, None, 0, None
        else:
            relu_flag = int(ctx.relu)
            dres = torch.empty_like(x) if ctx.has_res else None
        dgamma = torch.empty(Cc, dtype=torch.float32, device=x.device)
        dbeta = torch.empty(Cc, dtype=torch.float32, device=x.device)
        ws_ptr = _workspace(int(lib.lwc_bn_workspace_bytes(M, Cc)), x.device).data_ptr()
        st = torch.cuda.current_stream().cuda_stream
        rc = lib.lwc_bn_backward_ex(dy.data_ptr(), x.data_ptr(), _ptr(y), dx.data_ptr(), _ptr(dres), _ptr(weight),
                                    _ptr(bias), mean.data_ptr(), invstd.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), ws_ptr,
                                    M, Cc, abi.dtype_code(x.dtype), relu_flag, _ptr(mask), C.c_void_p(st))
        abi.check(rc, "lwc_bn_backward_ex")
        if premasked and ctx.has_res:
            dres = dy
        _native.counters.add(3)
        return (dx, dgamma if weight is not None else None, dbeta if weight is not None else None, None, None,
                dres, None, None, None, None, None, None, None, None, None)


def batch_norm_act(x, weight, bias, running_mean, running_var, residual=None, training=True,
                   momentum=0.1, eps=1e-5, relu=False, num_batches_tracked=None, stats=None, fuse_bwd=False):
    """``relu?(batch_norm(x) + residual?)`` in one fused pass structure (see module docstring).
    ``num_batches_tracked`` (int64 CUDA scalar) is incremented inside the statistics-finalize kernel."""
    _BatchNormAct.last_mask = (None, None, None)
    link = {} if (fuse_bwd and training and residual is None and torch.is_grad_enabled()) else None
    if stats is not None and training:
        y = _BatchNormAct.apply(x, weight, bias, running_mean, running_var, residual, training, momentum, eps, relu,
                                num_batches_tracked, stats.partials, stats.nparts, stats.count, link)
    else:
        y = _BatchNormAct.apply(x, weight, bias, running_mean, running_var, residual, training, momentum, eps, relu,
                                num_batches_tracked, None, 0, 0, link)
    if link:
        y.bn_link = link               # picked up by the consuming ops.nn.Conv2d
    mask, taken, flink = _BatchNormAct.last_mask
    if mask is not None:
        y.relu_mask, y.relu_mask_taken, y.relu_mask_link = mask, taken, flink       # see ops.conv.fork_conv1x1
        _BatchNormAct.last_mask = (None, None, None)
    return y


def maybe_batch_norm(module, x, residual=None, relu=False, fuse_bwd=False):
    """Module dispatch; returns ``None`` when the stock implementation must run."""
    if module.momentum is None:          # cumulative moving average: stock implementation
        return None
    if not supported(x, module.weight, module.bias, residual):
        return None
    training = module.training or module.running_mean is None
    nbt = None
    if training and module.track_running_stats and module.num_batches_tracked is not None:
        nbt = module.num_batches_tracked
        if not (nbt.is_cuda and nbt.dtype == torch.int64):
            nbt.add_(1)
            nbt = None
    return batch_norm_act(x, module.weight, module.bias, module.running_mean, module.running_var,
                          residual, training, module.momentum, module.eps, relu, nbt,
                          stats=getattr(x, "bn_stats", None), fuse_bwd=fuse_bwd)
