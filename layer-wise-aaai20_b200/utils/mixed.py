"""Mixed-precision helpers: half/bf16 model conversion and fp32 master weights.

Capabilities of the reference's fp16util (IMAGENET/training/fp16util.py:21-139):
``network_to_half`` keeps batch-norm in fp32 and casts inputs at the front;
``prep_param_lists`` builds fp32 master copies (optionally one flat tensor);
``model_grads_to_master_grads`` / ``master_params_to_model_params`` move data between
the two sets.  Generalised to bf16 (the B200 default) via ``dtype``.
"""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.nn as nn
from torch._utils import _flatten_dense_tensors, _unflatten_dense_tensors


class CastInput(nn.Module):
    def __init__(self, dtype):
        super().__init__()
        self.dtype = dtype

    def forward(self, x):
        return x.to(self.dtype)


def bn_to_float(module: nn.Module) -> nn.Module:
    """Recursively return batch-norm layers (params and buffers) to fp32."""
    if isinstance(module, nn.modules.batchnorm._BatchNorm):
        module.float()
    for child in module.children():
        bn_to_float(child)
    return module


def network_to_half(network: nn.Module, dtype=torch.float16) -> nn.Module:
    """``Sequential(cast, network.to(dtype))`` with batch-norm kept in fp32."""
    return nn.Sequential(CastInput(dtype), bn_to_float(network.to(dtype)))


def network_to_bf16(network: nn.Module) -> nn.Module:
    return network_to_half(network, torch.bfloat16)


def prep_param_lists(model: nn.Module, flat_master: bool = False) -> Tuple[List, List]:
    """Return ``(model_params, master_params)`` with fp32 masters carrying ``.grad`` slots."""
    model_params = [p for p in model.parameters() if p.requires_grad]
    if flat_master:
        try:
            flat = _flatten_dense_tensors([p.data for p in model_params]).float()
        except Exception as e:  # mixed dtypes cannot be flattened together
            raise ValueError("flat_master=True needs all parameters in one dtype") from e
        master = torch.nn.Parameter(flat)
        master.requires_grad = True
        if master.grad is None:
            master.grad = master.new_zeros(master.size())
        return model_params, [master]
    masters = [p.detach().clone().float() for p in model_params]
    for m in masters:
        m.requires_grad = True
    return model_params, masters


def model_grads_to_master_grads(model_params, master_params, flat_master: bool = False):
    if flat_master:
        master_params[0].grad.data.copy_(_flatten_dense_tensors([p.grad.data for p in model_params]))
        return
    for model, master in zip(model_params, master_params):
        if model.grad is None:
            master.grad = None
            continue
        if master.grad is None:
            master.grad = torch.empty_like(master.data)
        master.grad.data.copy_(model.grad.data)


def master_params_to_model_params(model_params, master_params, flat_master: bool = False):
    if flat_master:
        for model, master in zip(model_params,
                                 _unflatten_dense_tensors(master_params[0].data, model_params)):
            model.data.copy_(master)
        return
    for model, master in zip(model_params, master_params):
        model.data.copy_(master.data)


def bnwd_optim_params(model: nn.Module, model_params, master_params, weight_decay=None):
    """Two optimiser groups: batch-norm params with weight_decay 0, the rest default.

    Reference: IMAGENET/training/experimental_utils.py:5-22.
    """
    bn_ids = set()
    for m in model.modules():
        if isinstance(m, nn.modules.batchnorm._BatchNorm):
            bn_ids.update(id(p) for p in m.parameters(recurse=False))
    model_params, master_params = list(model_params), list(master_params)
    bn, rest = [], []
    for mp, pp in zip(model_params, master_params):
        (bn if id(mp) in bn_ids else rest).append(pp)
    groups = [{"params": bn, "weight_decay": 0.0}, {"params": rest}]
    if weight_decay is not None:
        groups[1]["weight_decay"] = weight_decay
    return [g for g in groups if g["params"]]
