#!/usr/bin/env python
"""Multi-GPU check of the whole-step CUDA graph (run under torchrun, >= 2 GPUs).

Two trainers per rank see the same per-rank batches: one steps eagerly, one captures the step after three eager
steps and replays it.  Losses and fp32 master weights must match bit for bit, replicas must agree, and no peer
flag may time out.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import nn as lnn  # noqa: E402
from layerwise_b200.parallel.exchange import CompressionConfig  # noqa: E402
from layerwise_b200.train.imagenet import Trainer  # noqa: E402


class MLP(torch.nn.Module):
    def __init__(self, d_in=768, hidden=256, classes=16):
        super().__init__()
        self.f1, self.f2, self.f3 = lnn.Linear(d_in, hidden), lnn.Linear(hidden, hidden), lnn.Linear(hidden, classes)

    def forward(self, x):
        return self.f3(torch.relu(self.f2(torch.relu(self.f1(x.flatten(1))))))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok, results = True, {}
    for method, ratio in (("Topk", 0.05), ("Randomk", 0.1), ("RandomDithering", 1.0)):
        def make(graph):
            torch.manual_seed(0)        # the model is built here, before Trainer seeds the generator itself
            cfg = CompressionConfig(compress="layerwise", method=method, ratio=ratio, error_feedback=True)
            return Trainer(model=MLP(), dtype="bf16", cfg=cfg, lr=0.01, momentum=0.9, weight_decay=1e-4,
                           device=f"cuda:{local}", overlap=True, engine="fused", no_bn_wd=False, cuda_graph=graph)
        a, b, c = make(False), make(True), make(False)
        g = torch.Generator().manual_seed(100 + rank)
        la, lb, lc = [], [], []
        for i in range(9):
            n = 32 if i == 6 else 64            # one odd-shaped (eager) step in the middle of the replays
            x = torch.randint(0, 256, (n, 3, 16, 16), dtype=torch.uint8, generator=g).cuda()
            y = torch.randint(0, 16, (n,), generator=g).cuda()
            for t, out in ((a, la), (b, lb), (c, lc)):
                t.optimizer.set_lr(0.01 * (1 + i % 3))
                out.append(float(t.train_step(x, y)[1].detach()))
        torch.cuda.synchronize()
        good = b._graph is not None and la == lb
        eager_repeatable = la == lc and all(torch.equal(pa, pc) for pa, pc in zip(a.master_params, c.master_params))
        for pa, pb in zip(a.master_params, b.master_params):
            good = good and torch.equal(pa, pb)
            ref = pb.detach().clone()
            dist.broadcast(ref, 0)
            good = good and torch.equal(ref, pb)
        st = b.exchanger.engine.stats()
        good = good and st["timeouts"] == 0 and a.exchanger.engine.stats()["timeouts"] == 0
        flag = torch.tensor([1.0 if good else 0.0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        results[method] = {"ok": bool(flag.item()), "losses_eager": la, "losses_graph": lb, "losses_eager2": lc,
                           "eager_repeatable": eager_repeatable,
                           "graph_kernels": b._graph_launches}
        ok = ok and bool(flag.item())
        a.exchanger.engine.close()
        c.exchanger.engine.close()
        b.exchanger.engine.close()
    if rank == 0:
        print(json.dumps({"world": world, "ok": ok, "results": results}, indent=1))
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
