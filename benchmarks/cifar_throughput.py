#!/usr/bin/env python
"""CIFAR-10 training throughput (img/s) of the dawn recipe networks, this framework vs the reference's own code.

    python benchmarks/cifar_throughput.py [--networks resnet9,alexnet,alexnet1,vgg16] [--steps 30]
    torchrun --nproc-per-node N benchmarks/cifar_throughput.py ...            (multi-GPU, weak scaling)

One step = forward + loss.sum().backward() + gradient exchange + SGD(Nesterov, wd) + zero_grad on a
synthetic 512 x 3 x 32 x 32 batch (the reference recipe's batch size, CIFAR10/dawn.py:111).  CUDA-event timed,
max over ranks.  Reference arm: the *unmodified* network builders of ``baseline/_ref/CIFAR10/dawn.py`` (its
definitions are exec'd up to ``def main`` because the module runs ``main()`` on import), ``Network``/``SGD`` of
``torch_backend.py`` and ``layerwise_compressed_comm`` of ``core.py``, fp32, NCCL.
The reference's published single-V100 number for ResNet-9 is ~17.1 k img/s (CIFAR10/demo.ipynb:726).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(step, steps, warmup, local):
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier(device_ids=[local])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) / steps


def ours(network, args, world):
    import layerwise_b200  # noqa: F401
    from layerwise_b200.models import build_cifar_model, trainable_params
    from layerwise_b200.ops.sgd import SGD
    from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    model = build_cifar_model(network).cuda().to(memory_format=torch.channels_last)
    if dtype != torch.float32:
        from layerwise_b200.utils.mixed import bn_to_float
        model = bn_to_float(model.to(dtype))
    cfg = CompressionConfig(compress=args.compress, method=args.method, ratio=args.ratio, error_feedback=True)
    ex = GradientExchanger(trainable_params(model), cfg, overlap=True)
    opt = SGD(trainable_params(model), lr=1e-4, momentum=0.9, weight_decay=5e-4 * args.batch, nesterov=True)
    x = torch.randn(args.batch, 3, 32, 32, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (args.batch,), device="cuda")
    model.train()

    if args.graph and hasattr(ex.engine, "enable_graph_mode"):
        from layerwise_b200.train.cifar import GraphedTrainStep
        graphed = GraphedTrainStep(model, ex, opt)
        return lambda: graphed({"input": x, "target": y})

    def step():
        out = model({"input": x, "target": y})
        out["loss"].sum().backward()
        ex.exchange()
        opt.step()
        if hasattr(ex.engine, "zero_grad"):
            ex.engine.zero_grad()
        else:
            model.zero_grad(set_to_none=False)
    return step


def reference(network, args, world):
    ref = os.path.join(ROOT, "baseline", "_ref")
    sys.path.insert(0, ref)
    ns = {"__name__": "ref_dawn_defs"}
    src = open(os.path.join(ref, "CIFAR10", "dawn.py")).read().split("def main")[0]
    old_argv, sys.argv = sys.argv, [sys.argv[0]]
    try:
        exec(compile(src, "baseline/_ref/CIFAR10/dawn.py", "exec"), ns)      # unmodified reference definitions
    finally:
        sys.argv = old_argv
    from CIFAR10.core import layerwise_compressed_comm
    key = network.lower()
    if key in ("resnet9", "resent9"):
        model = ns["Network"](ns["union"](ns["resnet9"](), ns["losses"])).cuda()
    elif key == "alexnet":
        model = ns["Network"](ns["union"](ns["alexnet"](), ns["losses"])).cuda()
    elif key == "alexnet1":
        model = ns["AlexNet"]().cuda()
    else:
        model = ns["vgg16model"].cuda()
    opt = ns["SGD"](ns["trainable_params"](model), lr=1e-4, momentum=0.9, weight_decay=5e-4 * args.batch, nesterov=True)
    x = torch.randn(args.batch, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (args.batch,), device="cuda")
    model.train(True)

    def step():      # body of run_batches (CIFAR10/core.py:306-321)
        out = model({"input": x, "target": y})
        out["loss"].sum().backward()
        layerwise_compressed_comm(model, float(world), args.method, args.ratio, 0.001, 255)
        opt.step()
        model.zero_grad()
    return step


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--networks", default="resnet9,alexnet,alexnet1,vgg16")
    ap.add_argument("--impl", default="both", choices=["ours", "reference", "both"])
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--compress", default="layerwise")
    ap.add_argument("--method", default="Topk")
    ap.add_argument("--ratio", type=float, default=0.01)
    ap.add_argument("--graph", action="store_true", help="ours: replay the step as one CUDA graph")
    ap.add_argument("--out", default="gpurun_out/cifar_throughput.json")
    args = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29588")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    torch.backends.cudnn.benchmark = True
    rows = []
    for net in args.networks.split(","):
        row = {"network": net, "world": world, "batch_per_gpu": args.batch, "compress": args.compress,
               "method": args.method, "ratio": args.ratio}
        for impl in (("ours", "reference") if args.impl == "both" else (args.impl,)):
            try:
                step = (ours if impl == "ours" else reference)(net, args, world)
                steps = args.steps if impl == "ours" else max(3, args.steps // 4)
                ms = timed(step, steps, args.warmup if impl == "ours" else 2, local)
                row[f"ms_{impl}"] = ms
                row[f"img_s_{impl}"] = args.batch * world / ms * 1e3
            except Exception as e:      # noqa: BLE001
                row[f"error_{impl}"] = f"{type(e).__name__}: {e}"[:200]
            torch.cuda.empty_cache()
        if "img_s_ours" in row and "img_s_reference" in row:
            row["speedup"] = row["img_s_ours"] / row["img_s_reference"]
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0:
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        json.dump(rows, open(args.out, "w"), indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
