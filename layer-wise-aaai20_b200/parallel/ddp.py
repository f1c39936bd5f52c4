"""Module wrappers: dense DDP and Random-k sparsifying DDP with error feedback.

Behavioural model: the reference's two hand-maintained copies of PyTorch-1.1's Python DDP
(IMAGENET/training/ddp.py:19-489 dense; IMAGENET/training/sparsified_ddp.py:20-495
Random-k + error feedback).  What they do, and how it is done here:

* constructor broadcasts parameters and buffers from rank 0 (ddp.py:190-194,
  sparsified_ddp.py:199-203)                        -> ``comm.broadcast_coalesced``
* gradients are bucketed (25 MiB, reverse order) and reduced from autograd hooks while
  backward is still running (ddp.py:394-456)        -> ``GradientExchanger(overlap=True)``;
  on B200 each bucket is one fused kernel sequence over peer memory, on CPU/gloo the
  exchange runs after backward
* Random-k: ``g += eps; mask = randperm(n) < ceil(n*randk); send g[mask]; eps = g*(1-mask)``
  with identical seeds on all ranks so only values travel (sparsified_ddp.py:403-414,:164)
  -> ``CompressionConfig(method='Randomk', error_feedback=True, shared_seed=True)``; the mask
  is a counter-based hash of (seed, step, tensor id, index), independent of hook order
* training ``forward`` re-broadcasts buffers (BN running stats) from rank 0 every
  iteration (ddp.py:373-377)                        -> ``broadcast_buffers`` flag
* ``check_reduction`` raises if a previous backward never reduced (ddp.py:312-327)

Out of scope: single-process multi-GPU replicate/scatter/gather -- one process
per GPU is the only mode, as in the reference's actual use (``device_ids=[local_rank]``).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import comm
from .exchange import CompressionConfig, GradientExchanger


class _DDPBase(nn.Module):
    def __init__(self, module: nn.Module, cfg: CompressionConfig, device_ids=None, output_device=None,
                 dim: int = 0, broadcast_buffers: bool = True, process_group=None,
                 check_reduction: bool = False, engine: str = "auto"):
        super().__init__()
        if device_ids is not None and len(device_ids) > 1:
            raise RuntimeError("only multi-process single-GPU mode is supported")
        self.module = module
        self.dim = dim
        self.process_group = process_group
        self.broadcast_buffers = broadcast_buffers
        self.check_reduction = check_reduction
        self.broadcast_bucket_size = 250 << 20
        self._buffer_bcast = None
        self._buffer_bcast_tried = False
        params = [p for p in module.parameters() if p.requires_grad]
        on_gpu = bool(params) and params[0].is_cuda
        self.exchanger = GradientExchanger(params, cfg, group=process_group, engine=engine, overlap=on_gpu)
        # every replica starts from rank 0's parameters and buffers (reference sparsified_ddp.py:199-203): through peer
        # memory when the ranks share it, else one coalesced collective
        states = [t for t in module.state_dict().values() if torch.is_tensor(t)]
        if comm.world_size(process_group) > 1 and states:
            init_bcast = self._peer_broadcaster(states)
            if init_bcast is not None:
                init_bcast.broadcast()
                torch.cuda.synchronize()
                comm.barrier(process_group)
                init_bcast.close()
            else:
                comm.broadcast_coalesced(states, 0, process_group, self.broadcast_bucket_size)
        self._hook_driven = getattr(self.exchanger.engine, "overlap", False)
        self._awaiting_reduction = False
        if not self._hook_driven and params:
            # CPU / torch engine: reduce once the whole backward has finished
            self._pending = 0
            self._nparams = len(params)
            for p in params:
                p.register_post_accumulate_grad_hook(self._count_hook)

    # torch engine: fire the exchange when the last gradient of this backward has arrived
    def _count_hook(self, param):
        self._pending += 1
        if self._pending == self._nparams:
            self._pending = 0
            self.exchanger.exchange()
            self._awaiting_reduction = False

    def _peer_broadcaster(self, tensors):
        """A peer-memory broadcaster for ``tensors`` when the ranks share NVLink peer memory (CUDA tensors, fused / dense
        engine, i.e. one node), else None (NCCL / gloo broadcast)."""
        if not tensors or not all(t.is_cuda for t in tensors):
            return None
        if getattr(self.exchanger, "engine_name", "torch") not in ("fused", "dense"):
            return None
        bc = None
        try:
            from .peer import PeerBroadcaster
            bc = PeerBroadcaster(tensors, self.process_group, root=0)
        except Exception:           # no peer access between the ranks: keep the collective
            bc = None
        # all ranks must take the same path: one that could not map the heaps sends everybody to the collective
        ok = comm.sum_tensor(torch.tensor([1.0 if bc is not None else 0.0], device=tensors[0].device), self.process_group)
        if int(ok.item()) != comm.world_size(self.process_group):
            if bc is not None:
                bc.close()
            return None
        return bc

    def _sync_buffers(self):
        """Batch-norm running statistics follow rank 0 before every training forward (reference ddp.py:373-377): one
        ``lwc_peer_broadcast`` launch on the step path, no NCCL call."""
        if self.broadcast_buffers and comm.world_size(self.process_group) > 1:
            bufs = list(self.module.buffers())
            if not bufs:
                return
            if self._buffer_bcast is None and not self._buffer_bcast_tried:
                self._buffer_bcast_tried = True
                self._buffer_bcast = self._peer_broadcaster(bufs)
            if self._buffer_bcast is not None:
                self._buffer_bcast.broadcast()
            else:
                comm.broadcast_coalesced(bufs, 0, self.process_group, self.broadcast_bucket_size)

    def forward(self, *inputs, **kwargs):
        if self.training and torch.is_grad_enabled():
            if self.check_reduction and self._awaiting_reduction:
                raise RuntimeError("the previous backward did not reduce every gradient bucket "
                                   "(call finish_reduction() / step through the wrapper)")
            self._sync_buffers()
            self._awaiting_reduction = True
        return self.module(*inputs, **kwargs)

    def finish_reduction(self):
        """Join the hook-launched exchange (fused engine); call before ``optimizer.step()``."""
        if self._hook_driven:
            self.exchanger.exchange()
        self._awaiting_reduction = False

    def train(self, mode: bool = True):
        super().train(mode)
        self.module.train(mode)
        return self


class DistributedDataParallel(_DDPBase):
    """Dense gradient averaging driven from autograd hooks (reference ddp.py)."""

    def __init__(self, module, device_ids=None, output_device=None, dim=0, broadcast_buffers=True,
                 process_group=None, bucket_cap_mb: float = 25, check_reduction=False, engine="auto"):
        cfg = CompressionConfig(compress="none", bucket_cap_mb=bucket_cap_mb)
        super().__init__(module, cfg, device_ids, output_device, dim, broadcast_buffers, process_group,
                         check_reduction, engine)


class RandomKSparsifiedDDP(_DDPBase):
    """Random-k sparsified gradient averaging with error feedback (reference sparsified_ddp.py)."""

    def __init__(self, module, device_ids=None, output_device=None, dim=0, broadcast_buffers=True,
                 process_group=None, bucket_cap_mb: float = 25, check_reduction=False, randk: float = 1.0,
                 seed: int = 2147483647, engine="auto"):
        self.randk = randk
        cfg = CompressionConfig(compress="layerwise", method="Randomk" if randk < 1.0 else "none",
                                ratio=randk, error_feedback=randk < 1.0, seed=seed, shared_seed=True,
                                bucket_cap_mb=bucket_cap_mb)
        super().__init__(module, cfg, device_ids, output_device, dim, broadcast_buffers, process_group,
                         check_reduction, engine)

    @property
    def epsilon(self):
        """Error-feedback residuals, one flat fp32 tensor per parameter."""
        eng = self.exchanger.engine
        return getattr(eng, "resid_views", None) or getattr(eng, "residual", None)


class DDP(DistributedDataParallel):
    """Eval bypasses the wrapper; ``state_dict`` is the bare module's (reference dist_utils.py:5-16)."""

    def forward(self, *inputs, **kwargs):
        if self.training:
            return super().forward(*inputs, **kwargs)
        return self.module(*inputs, **kwargs)

    def state_dict(self, *args, **kwargs):
        return self.module.state_dict(*args, **kwargs)

    def load_state_dict(self, *args, **kwargs):
        return self.module.load_state_dict(*args, **kwargs)
