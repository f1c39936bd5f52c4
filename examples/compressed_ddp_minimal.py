#!/usr/bin/env python
"""Smallest end-to-end use of the library API on your own model: compressed data-parallel SGD.

    python examples/compressed_ddp_minimal.py                                   # one process (CPU or one GPU)
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 \
        examples/compressed_ddp_minimal.py --method Topk --ratio 0.01           # 2 ranks: NCCL + peer memory on GPUs,
                                                                                 # gloo + the reference math on CPUs

What the reference spreads over ``core.py:175-301`` (``layerwise_compressed_comm`` called by hand between
``backward()`` and ``optimizer.step()``) is three objects here: a ``CompressionConfig``, a ``GradientExchanger`` and,
optionally, ``FusedSGD``.  On CUDA the exchanger uses the fused peer-memory engine (compress + pack + store into
every rank's slot in one launch sequence, overlapped with backward through gradient hooks); elsewhere the same
arithmetic runs through ``torch.distributed``.
"""
import argparse
import os
import sys

import torch
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops.sgd import FusedSGD  # noqa: E402
from layerwise_b200.parallel import comm  # noqa: E402
from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--method", default="Topk", help="Topk | Randomk | Thresholdv | AdaptiveThreshold | TernGrad | QSGD | none")
    ap.add_argument("--ratio", type=float, default=0.05)
    ap.add_argument("--compress", default="layerwise", choices=["layerwise", "entiremodel", "none"])
    ap.add_argument("--steps", type=int, default=30)
    args = ap.parse_args(argv)

    world = comm.env_world_size()
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(comm.env_local_rank() % torch.cuda.device_count())
    device = torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")
    if world > 1:
        comm.init_process_group(backend="nccl" if cuda else "gloo", init_method="env://",
                                device=device if cuda else None)
    rank = comm.rank()

    torch.manual_seed(0)                                     # identical replicas
    model = nn.Sequential(nn.Linear(64, 256), nn.ReLU(), nn.Linear(256, 256), nn.ReLU(), nn.Linear(256, 10)).to(device)
    cfg = CompressionConfig(compress=args.compress, method=args.method, ratio=args.ratio, qstates=255,
                            error_feedback=args.method not in ("none", "TernGrad", "QSGD"))
    exchanger = GradientExchanger(model.parameters(), cfg, overlap=cuda)        # engine: fused on CUDA, torch elsewhere
    optimizer = FusedSGD(model.parameters(), lr=0.05, momentum=0.9, nesterov=True, weight_decay=1e-4)

    g = torch.Generator().manual_seed(100 + rank)            # every rank sees its own shard
    w_true = torch.randn(64, 10, generator=torch.Generator().manual_seed(7))
    first = last = None
    for step in range(args.steps):
        x = torch.randn(128, 64, generator=g)
        y = (x @ w_true).argmax(1)
        loss = nn.functional.cross_entropy(model(x.to(device)), y.to(device))
        loss.backward()
        exchanger.exchange()                                 # compress -> exchange -> decompress; p.grad = average
        optimizer.step()
        zero = getattr(exchanger.engine, "zero_grad", None)
        zero() if zero else optimizer.zero_grad(set_to_none=False)
        last = float(loss.detach())
        first = last if first is None else first
    health = exchanger.check_health()                        # raises if a peer's flags stopped arriving
    if rank == 0:
        print(f"engine={exchanger.engine_name} world={world} method={args.method} loss {first:.3f} -> {last:.3f} "
              f"health={health}")
    if world > 1:
        torch.distributed.destroy_process_group()
    return first, last


if __name__ == "__main__":
    main()
