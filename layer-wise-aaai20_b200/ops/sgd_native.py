"""Launch glue for the fused multi-tensor SGD kernel (``csrc/sgd.cu``)."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native
from .._native import abi


class _Plan:
    def __init__(self, key, handle, grad_dtype, model_dtype):
        self.key, self.handle, self.grad_dtype, self.model_dtype = key, handle, grad_dtype, model_dtype


def _build(opt, params, bufs):
    lib = _native.lib()
    n = len(params)
    arr = (abi.LwcSgdTensor * n)()
    model_dtype = None
    for i, p in enumerate(params):
        model = opt._model_copy.get(id(p))
        arr[i].param = p.data_ptr()
        arr[i].grad = p.grad.data_ptr()
        arr[i].mom = 0 if bufs is None else bufs[i].data_ptr()
        arr[i].model = 0 if model is None else model.data_ptr()
        arr[i].n = p.numel()
        if model is not None:
            model_dtype = model.dtype
    handle = C.c_void_p()
    abi.check(lib.lwc_sgd_plan_create(C.byref(handle), arr, n), "lwc_sgd_plan_create")
    return handle, model_dtype


def _same_dense_layout(p, g) -> bool:
    if p.stride() != g.stride() or p.size() != g.size():
        return False
    return p.is_contiguous() or (p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last))


_POISON = {}


def _poison_ptr() -> int:
    """Device address of the exchange's fail-closed word on the current device (see ``csrc/exchange.cu``)."""
    dev = torch.cuda.current_device()
    p = _POISON.get(dev)
    if p is None:
        out = C.c_void_p()
        abi.check(_native.lib().lwc_poison_ptr(C.byref(out)), "lwc_poison_ptr")
        p = _POISON[dev] = int(out.value or 0)
    return p


def step_group(opt, gi, group, params):
    """One fused launch per (dtype) subset of a param group."""
    lib = _native.lib()
    mu = group["momentum"]
    # split by (param dtype must be fp32), grad dtype
    subsets = {}
    for p in params:
        lowp = p.dtype in (torch.bfloat16, torch.float16) and p.grad.dtype == p.dtype \
            and opt._model_copy.get(id(p)) is None
        if not (p.dtype == torch.float32 or lowp) or not _same_dense_layout(p, p.grad):
            opt._step_reference(group, [p])
            continue
        # key: (grad dtype, param dtype) -- 16-bit weights without fp32 masters take the low-precision kernel
        subsets.setdefault((p.grad.dtype, p.dtype), []).append(p)
    for (gdt, pdt), ps in subsets.items():
        first = False
        bufs = None
        if mu != 0:
            # `first_step` is one flag per launch: tensors that get their first gradient later than the others (or a
            # partly restored optimizer state) take the plain per-tensor step once, so existing momentum is never reset
            fresh = [p for p in ps if opt.state[p].get("momentum_buffer") is None]
            if fresh and len(fresh) < len(ps):
                opt._step_reference(group, fresh)
                ps = [p for p in ps if not any(p is f for f in fresh)]
            if fresh and len(fresh) == len(ps) and torch.cuda.is_current_stream_capturing():
                raise RuntimeError("FusedSGD: the first optimizer step (momentum buffers are created in it) cannot be "
                                   "captured in a CUDA graph -- run at least one eager step first")
            bufs = []
            for p in ps:
                st = opt.state[p]
                if "momentum_buffer" not in st or st["momentum_buffer"] is None:
                    st["momentum_buffer"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    first = True
                if st["momentum_buffer"].dtype != p.dtype:        # e.g. a checkpoint written by another precision mode
                    st["momentum_buffer"] = st["momentum_buffer"].to(p.dtype)
                bufs.append(st["momentum_buffer"])
        key = (gi, gdt, pdt) + tuple(p.data_ptr() for p in ps) + tuple(p.grad.data_ptr() for p in ps) + \
            (tuple(b.data_ptr() for b in bufs) if bufs else ())
        plan = opt._plans.get((gi, gdt, pdt))
        if plan is None or plan.key != key:
            if plan is not None:
                lib.lwc_sgd_plan_destroy(plan.handle)
            handle, mdt = _build(opt, ps, bufs)
            plan = _Plan(key, handle, gdt, mdt)
            opt._plans[(gi, gdt, pdt)] = plan
        d = abi.LwcSgdDesc()
        d.num_tensors = len(ps)
        d.grad_dtype = abi.dtype_code(gdt)
        d.param_dtype = abi.dtype_code(pdt)
        d.model_dtype = abi.dtype_code(plan.model_dtype) if plan.model_dtype is not None else abi.LWC_F32
        d.lr, d.momentum, d.dampening = float(group["lr"]), float(mu), float(group["dampening"])
        d.weight_decay, d.grad_scale = float(group["weight_decay"]), float(opt.grad_scale)
        d.nesterov, d.first_step = int(bool(group["nesterov"])), int(first)
        lr_dev = opt._lr_dev.get(gi) if opt._lr_dev is not None else None
        if lr_dev is not None:               # graph-safe: the kernel reads the learning rate from device memory
            if not torch.cuda.is_current_stream_capturing():
                opt.sync_device_lr()
            d.lr_dev = lr_dev.data_ptr()
        d.skip_flag = _poison_ptr()
        st = torch.cuda.current_stream().cuda_stream
        abi.check(lib.lwc_sgd_step(plan.handle, C.byref(d), C.c_void_p(st)), "lwc_sgd_step")
        _native.counters.add(1)
