#!/usr/bin/env python
"""compress + exchange + decompress ms/step: fused engine vs the reference's own loops over NCCL.

Covers BASELINE.json configs 2, 4 and 5 (VGG-16 layer-wise vs entire-model Random-k, AlexNet QSGD /
TernGrad, ratio sweep).  Gradients are random tensors with the exact parameter shapes of each
model (the exchange does not depend on how the gradient was produced); device-timed with CUDA events,
max over ranks.  Launch with torchrun for W > 1:

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/exchange_sweep.py

Reference side: ``layerwise_compressed_comm`` imported unmodified from ``baseline/_ref`` (layer-wise); the
reference's entire-model function raises TypeError as committed, so its math is run through
this repo's ``TorchEngine`` (same ops, ``.parameters()`` fixed) and labelled ``ref-math``.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import layerwise_b200  # noqa: E402,F401
from layerwise_b200 import models  # noqa: E402
from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger  # noqa: E402


def param_shapes(name):
    with torch.device("meta"):
        m = {"resnet9": lambda: models.build_cifar_model("resnet9"), "alexnet": models.AlexNet,
             "vgg16": models.vgg16, "resnet50": models.resnet50}[name]()
    return [tuple(p.shape) for p in m.parameters()]


class Bag(torch.nn.Module):
    """A module that only holds parameters of the given shapes (what the exchange functions iterate)."""

    def __init__(self, shapes, device):
        super().__init__()
        self.ps = torch.nn.ParameterList([torch.nn.Parameter(torch.zeros(s, device=device)) for s in shapes])


def time_ms(fn, iters, local):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier(device_ids=[local])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", default="resnet9,alexnet,vgg16,resnet50")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--nvls", action="store_true", help="multicast payload stores in the fused engine")
    ap.add_argument("--only-ours", action="store_true")
    ap.add_argument("--out", default="gpurun_out/exchange_sweep.json")
    args = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29577")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    ref_fn = None
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isdir(os.path.join(ref_dir, "CIFAR10")):
        sys.path.insert(0, ref_dir)
        from CIFAR10.core import layerwise_compressed_comm as ref_fn  # noqa: E402
    cases = [("Topk", dict(ratio=0.001)), ("Topk", dict(ratio=0.01)), ("Topk", dict(ratio=0.1)),
             ("Randomk", dict(ratio=0.001)), ("Randomk", dict(ratio=0.01)), ("Randomk", dict(ratio=0.1)),
             ("TernGrad", {}), ("RandomDithering", dict(qstates=255)), ("none", {})]
    if args.quick:
        cases = [("Topk", dict(ratio=0.01)), ("Randomk", dict(ratio=0.01)), ("TernGrad", {}), ("RandomDithering", dict(qstates=255)),
                 ("none", {})]
    rows = []
    for mname in args.models.split(","):
        shapes = param_shapes(mname)
        numel = sum(int(torch.tensor(s).prod()) for s in shapes)
        for compress in ("layerwise", "entiremodel"):
            for method, kw in cases:
                cfg = CompressionConfig(compress=compress, method=method, error_feedback=method not in ("none",), **kw)
                g = torch.Generator(device="cuda").manual_seed(7 + rank)
                # ---- ours --------------------------------------------------------------------------
                bag = Bag(shapes, "cuda")
                ex = GradientExchanger(bag, cfg, engine="fused", nvls=args.nvls)
                for p in bag.parameters():
                    p.grad.copy_(torch.randn(p.shape, device="cuda", generator=g))
                ms_ours = time_ms(ex.exchange, args.iters, local)
                st = ex.engine.stats()
                ex.engine.close()
                del bag, ex
                torch.cuda.empty_cache()
                # ---- reference ---------------------------------------------------------------------
                if args.only_ours:
                    row = {"model": mname, "compress": compress, "method": method, **kw, "world": world, "nvls": args.nvls,
                           "ms_ours": ms_ours, "timeouts": st["timeouts"], "dropped": st["dropped"]}
                    rows.append(row)
                    if rank == 0:
                        print(json.dumps(row), flush=True)
                    continue
                bag = Bag(shapes, "cuda")
                for p in bag.parameters():
                    p.grad = torch.randn(p.shape, device="cuda", generator=g)
                K, V, Q = kw.get("ratio"), 0.001, kw.get("qstates", 255)
                label = "reference"
                if compress == "layerwise" and ref_fn is not None:
                    fn = lambda: ref_fn(bag, float(world), method, K, V, Q)      # noqa: E731
                else:
                    label = "ref-math"
                    rcfg = CompressionConfig(compress=compress, method=method, error_feedback=False, **kw)
                    rex = GradientExchanger(bag, rcfg, engine="torch")
                    fn = rex.exchange
                try:
                    ms_ref = time_ms(fn, max(2, args.iters // 3), local)
                except Exception as e:      # noqa: BLE001
                    ms_ref, label = None, f"failed: {type(e).__name__}"
                del bag
                torch.cuda.empty_cache()
                row = {"model": mname, "tensors": len(shapes), "numel": numel, "compress": compress, "method": method,
                       **kw, "world": world, "ms_ours": ms_ours, "ms_ref": ms_ref, "ref_kind": label,
                       "speedup": (ms_ref / ms_ours) if ms_ref else None,
                       "payload_bytes_per_dst": st["bytes"] // max(args.iters + 3, 1), "dropped": st["dropped"], "timeouts": st["timeouts"]}
                rows.append(row)
                if rank == 0:
                    print(json.dumps(row), flush=True)
    if rank == 0:
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        json.dump(rows, open(args.out, "w"), indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
