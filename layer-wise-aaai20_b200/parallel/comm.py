"""Process-group plumbing (one process per GPU, ``torch.distributed``).

Covers the reference's helpers (IMAGENET/training/dist_utils.py:21-28: ``reduce_tensor``,
``sum_tensor``, ``env_world_size``, ``env_rank``) and both of its rendezvous styles:
explicit ``tcp://`` address + rank + world size with gloo (CIFAR10/core.py:334) and
``env://`` with NCCL (train_imagenet_nv.py:160-163).
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as dist


def env_world_size() -> int:
    return int(os.environ.get("WORLD_SIZE", 1))


def env_rank() -> int:
    return int(os.environ.get("RANK", 0))


def env_local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", os.environ.get("OMPI_COMM_WORLD_LOCAL_RANK", 0)))


def is_initialized() -> bool:
    return dist.is_available() and dist.is_initialized()


def world_size(group=None) -> int:
    return dist.get_world_size(group) if is_initialized() else 1


def rank(group=None) -> int:
    return dist.get_rank(group) if is_initialized() else 0


def init_process_group(backend: Optional[str] = None, init_method: Optional[str] = None,
                       world_size: Optional[int] = None, rank: Optional[int] = None,
                       device: Optional[torch.device] = None):
    """Initialise the default group; tolerant of a bare ``host[:port]`` master address.

    The reference's default ``--master_address 127.0.0.1`` is not a valid init URL
    (CIFAR10/dawn.py:11); a bare address is completed to ``tcp://addr:29500`` here.
    """
    if is_initialized():
        return
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    kw = {}
    if init_method and init_method != "env://":
        if "://" not in init_method:
            host = init_method if ":" in init_method else f"{init_method}:29500"
            init_method = f"tcp://{host}"
        kw.update(init_method=init_method, world_size=world_size, rank=rank)
    else:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kw.update(init_method="env://",
                  world_size=world_size if world_size is not None else env_world_size(),
                  rank=rank if rank is not None else env_rank())
    if backend == "nccl" and device is not None:
        try:
            kw["device_id"] = torch.device(device)
        except Exception:
            pass
    dist.init_process_group(backend=backend, **kw)


def sum_tensor(t: torch.Tensor, group=None) -> torch.Tensor:
    """All-reduce(SUM) of a clone of ``t`` (identity without a process group)."""
    out = t.clone()
    if is_initialized():
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


def reduce_tensor(t: torch.Tensor, group=None) -> torch.Tensor:
    return sum_tensor(t, group) / world_size(group)


def max_tensor(t: torch.Tensor, group=None) -> torch.Tensor:
    out = t.clone()
    if is_initialized():
        dist.all_reduce(out, op=dist.ReduceOp.MAX, group=group)
    return out


def barrier(group=None):
    if is_initialized():
        if dist.get_backend(group) == "nccl":
            dist.barrier(group=group, device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier(group=group)


def broadcast_coalesced(tensors, src: int = 0, group=None, bucket_bytes: int = 250 << 20):
    """Broadcast many tensors from ``src`` in flat buckets of at most ``bucket_bytes``.

    Stand-in for the removed c10d helper the reference calls at wrap time
    (IMAGENET/training/sparsified_ddp.py:199-203, :366-367).
    """
    if not is_initialized() or world_size(group) == 1:
        return
    by_dtype = {}
    for t in tensors:
        by_dtype.setdefault((t.dtype, t.device), []).append(t)
    for (_, _), ts in by_dtype.items():
        bucket, size = [], 0
        def flush():
            if not bucket:
                return
            flat = torch.cat([t.detach().reshape(-1) for t in bucket])
            dist.broadcast(flat, src, group=group)
            off = 0
            for t in bucket:
                n = t.numel()
                t.detach().copy_(flat[off:off + n].view_as(t))
                off += n
        for t in ts:
            nbytes = t.numel() * t.element_size()
            if bucket and size + nbytes > bucket_bytes:
                flush()
                bucket, size = [], 0
            bucket.append(t)
            size += nbytes
        flush()
