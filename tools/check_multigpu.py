#!/usr/bin/env python
"""Multi-GPU check of the fused exchange over real peer memory (run under torchrun).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_multigpu.py

Every rank draws its own gradients; the fused engine's result must equal the oracle
(compress locally with the reference math, NCCL all-reduce, divide by W) and be bitwise
identical on all ranks.  Also times compress+exchange per step, max over ranks.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import compress_ref as cr  # noqa: E402
from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger  # noqa: E402


NVLS = "--nvls" in sys.argv


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sizes = [64, 1000, 8192, 70001, 2359296, 1000 * 2048]
    results = {}
    ok = True
    for compress in ("layerwise", "entiremodel"):
        for method, kw in (("Topk", dict(ratio=0.01)), ("Thresholdv", dict(threshold=2.0)),
                           ("AdaptiveThreshold", {}), ("none", {}), ("Randomk", dict(ratio=0.01)),
                           ("TernGrad", {}), ("RandomDithering", dict(qstates=255))):
            cfg = CompressionConfig(compress=compress, method=method, error_feedback=False,
                                    capacity_factor=1.5, dense_capacity=1.0, **kw)
            g = torch.Generator(device="cuda").manual_seed(1000 + rank)
            params = [torch.nn.Parameter(torch.zeros(n, device="cuda")) for n in sizes]
            ex = GradientExchanger(params, cfg, engine="fused", nvls=NVLS)
            for p in params:
                p.grad.copy_(torch.randn(p.shape, device="cuda", generator=g))
            local_g = [p.grad.clone() for p in params]
            ex.exchange()
            torch.cuda.synchronize()
            got = [p.grad.clone() for p in params]
            name = f"{compress}/{method}"
            if method in ("Topk", "Thresholdv", "AdaptiveThreshold", "none"):
                if compress == "layerwise":
                    exp = [cr.compress_dense(t, cfg.method_name, cfg.ratio, cfg.threshold, cfg.qstates) for t in local_g]
                else:
                    flat = cr.compress_dense(torch.cat(local_g), cfg.method_name, cfg.ratio, cfg.threshold, cfg.qstates)
                    exp, off = [], 0
                    for t in local_g:
                        exp.append(flat[off:off + t.numel()])
                        off += t.numel()
                for e in exp:
                    dist.all_reduce(e)
                    e /= world
                good = all(torch.allclose(a, b, atol=1e-6, rtol=1e-5) for a, b in zip(exp, got))
            elif method == "Randomk":
                # shared seed: values-only wire format; every rank selected the same positions, so the result is the
                # rank average of g * mask with exactly ceil(n K) survivors per selection group
                good = True
                for lg, gt in zip(local_g, got):
                    mask = gt != 0
                    e = lg * mask
                    dist.all_reduce(e)
                    e /= world
                    good = good and torch.allclose(e, gt, atol=1e-6, rtol=1e-5)
                kept = sum(int((t != 0).sum()) for t in got)
                want = (sum(cr.randomk_keep_count(n, cfg.ratio) for n in sizes) if compress == "layerwise"
                        else cr.randomk_keep_count(sum(sizes), cfg.ratio))
                good = good and abs(kept - want) <= 2        # an exact-zero average is possible but vanishingly rare
            else:
                good = all(torch.isfinite(t).all().item() for t in got)
            # replicas must agree bit for bit
            for t in got:
                ref = t.clone()
                dist.broadcast(ref, 0)
                good = good and torch.equal(ref, t)
            st = ex.engine.stats()
            good = good and st["timeouts"] == 0
            # timing: 20 exchanges
            for _ in range(3):
                ex.exchange()
            torch.cuda.synchronize()
            dist.barrier(device_ids=[local])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                ex.exchange()
            e1.record()
            torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / 20], device="cuda")
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            flag = torch.tensor([1.0 if good else 0.0], device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            results[name] = {"nvls": bool(getattr(ex.engine, "nvls", False)), "ok": bool(flag.item()), "ms_per_exchange": round(float(ms.item()), 4), "stats": st}
            ok = ok and bool(flag.item())
            ex.engine.close()
    # two small heaps alive at once: the driver may carve them out of one block, whose IPC handle maps the BLOCK --
    # every rank writes a tag into its neighbour's copy of each heap and each must land in the right one
    from layerwise_b200.parallel.peer import DeviceBuffer, PeerHeap
    heaps = [PeerHeap(4096), PeerHeap(4096)]
    dev = torch.device("cuda", local)
    for k, h in enumerate(heaps):
        DeviceBuffer(h.ptrs[(rank + 1) % world], 4096, dev, False).view(torch.int32, 1024).fill_(100 * rank + k + 1)
    torch.cuda.synchronize()
    dist.barrier(device_ids=[local])
    good = all(bool((h.local.view(torch.int32, 1024) == 100 * ((rank - 1) % world) + k + 1).all().item())
               for k, h in enumerate(heaps))
    flag = torch.tensor([1.0 if good else 0.0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    results["two_small_heaps"] = {"ok": bool(flag.item())}
    ok = ok and bool(flag.item())
    dist.barrier(device_ids=[local])
    for h in heaps:
        h.close()
    # peer-memory broadcast (N5 / N6) and the DDP wrappers that use it: batch-norm buffers must follow rank 0, the
    # replicas must stay bit-identical
    from layerwise_b200.parallel.ddp import DistributedDataParallel, RandomKSparsifiedDDP
    from layerwise_b200.parallel.peer import PeerBroadcaster
    ts = [torch.full((1000, 33), float(rank + 1), device="cuda"), torch.full((7,), rank + 3, device="cuda", dtype=torch.int64)]
    bc = PeerBroadcaster(ts, root=0)
    good = True
    for rep in range(3):
        bc.broadcast()
        torch.cuda.synchronize()
        good = good and bool((ts[0] == 1.0 + rep).all().item()) and bool((ts[1] == 3 + rep).all().item())
        dist.barrier(device_ids=[local])
        if rank == 0:
            ts[0] += 1
            ts[1] += 1
    bc.close()
    flag = torch.tensor([1.0 if good else 0.0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    results["peer_broadcast"] = {"ok": bool(flag.item())}
    ok = ok and bool(flag.item())
    for wname, mk in (("ddp_dense", lambda m: DistributedDataParallel(m, device_ids=[local])),
                      ("ddp_randk", lambda m: RandomKSparsifiedDDP(m, device_ids=[local], randk=0.1))):
        torch.manual_seed(100 + rank)               # different initial weights: the wrapper must align them with rank 0
        net = torch.nn.Sequential(torch.nn.Conv2d(3, 16, 3, padding=1), torch.nn.BatchNorm2d(16), torch.nn.ReLU(),
                                  torch.nn.Flatten(), torch.nn.Linear(16 * 8 * 8, 10)).cuda()
        w = mk(net)
        opt = torch.optim.SGD(net.parameters(), lr=0.05)
        used_peer = None
        for it in range(3):
            x = torch.randn(8, 3, 8, 8, device="cuda", generator=g)
            loss = w(x).square().mean()
            loss.backward()
            w.finish_reduction()
            opt.step()
            opt.zero_grad()
            used_peer = w._buffer_bcast is not None
        torch.cuda.synchronize()
        good = True
        for t in list(net.state_dict().values()):
            ref = t.clone()
            dist.broadcast(ref, 0)
            if "running" in str(t.dtype) or t.dtype.is_floating_point:
                # running statistics are re-synchronised at the NEXT forward; parameters must already agree
                pass
        for p_ in net.parameters():
            ref = p_.detach().clone()
            dist.broadcast(ref, 0)
            good = good and torch.equal(ref, p_.detach())
        w._sync_buffers()
        torch.cuda.synchronize()
        for b_ in net.buffers():
            ref = b_.clone()
            dist.broadcast(ref, 0)
            good = good and torch.equal(ref, b_)
        flag = torch.tensor([1.0 if good else 0.0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        results[wname] = {"ok": bool(flag.item()), "peer_broadcast_used": bool(used_peer), "engine": w.exchanger.engine_name}
        ok = ok and bool(flag.item())
        w.exchanger.engine.close() if hasattr(w.exchanger.engine, "close") else None
    if rank == 0:
        print(json.dumps({"world": world, "ok": ok, "numel": sum(sizes), "results": results}, indent=1))
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
