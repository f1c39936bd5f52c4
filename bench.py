#!/usr/bin/env python
"""Headline benchmark: ResNet-50 training img/s with layer-wise Top-k 1 % gradient exchange.

    python bench.py --gpus N --steps K --warmup W            # this framework (fused engine)
    python bench.py --impl reference --gpus N ...            # unmodified reference code path

For N > 1 launch under ``python -m torch.distributed.run --nproc-per-node N`` (one rank per
GPU, NCCL for the plumbing).  Synthetic ImageNet-shaped data (224x224 uint8), random-init
weights, per-GPU batch 224 (the reference recipe's batch at 224 px,
IMAGENET/training/train_imagenet_nv.py:211).  One step = normalise -> forward -> loss ->
backward -> compress/exchange/decompress -> SGD(momentum, nesterov, wd) incl. the fp32-master
update.  Timed with CUDA events between barriers, max over ranks; clocks sampled with
nvidia-smi during the timed region.  Rank 0 prints ONE JSON line.

The reference arm imports the reference's own ``resnet50``, ``fp16util`` and
``layerwise_compressed_comm`` from ``baseline/_ref`` (a verbatim copy of the reference tree;
it has no setup.py, so ``pip install`` is impossible -- see DESIGN.md) and drives them exactly
like its ``train()`` does in ``--fp16`` mode (train_imagenet_nv.py:404-441).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

BASELINE_PUBLISHED = None      # BASELINE.json "published": {} -> vs_baseline is null


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--batch", type=int, default=224, help="per-GPU batch")
    p.add_argument("--size", type=int, default=224)
    p.add_argument("--arch", default="resnet50")
    p.add_argument("--method", default="Topk")
    p.add_argument("--ratio", type=float, default=0.01)
    p.add_argument("--compress", default="layerwise")
    p.add_argument("--no-ef", action="store_true")
    p.add_argument("--no-overlap", action="store_true")
    p.add_argument("--engine", default="auto")
    p.add_argument("--ref-dtype", default="fp16", choices=["fp16", "fp32"])
    p.add_argument("--skip-e2e", action="store_true")
    p.add_argument("--cifar-dtype", default="bf16", choices=["bf16", "fp32"],
                   help="compute dtype of our arm for the CIFAR architectures (the reference's CIFAR code is fp32)")
    p.add_argument("--nvls", action="store_true", help="NVSwitch-multicast payload stores (N > 1)")
    p.add_argument("--phase-times", action="store_true", help="also report compress+comm ms/step")
    p.add_argument("--graph", action="store_true", default=True,
                   help="replay the whole train step as one CUDA graph (default; bitwise identical to eager stepping)")
    p.add_argument("--no-graph", dest="graph", action="store_false", help="step eagerly (~1.5k launches per step)")
    return p.parse_args()


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def setup_dist(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1 or args.impl == "reference":
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    return world, rank, local


def barrier():
    if dist.is_initialized():
        dist.barrier(device_ids=[torch.cuda.current_device()])
    torch.cuda.synchronize()


def max_over_ranks(ms: float) -> float:
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return ms
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed(step_fn, steps, sampler=None):
    barrier()
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step_fn(i)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop() if sampler else None
    barrier()
    return max_over_ranks(e0.elapsed_time(e1)), clocks


# ---------------------------------------------------------------------------------------------------------
def build_ours(args, world, rank):
    import layerwise_b200  # noqa: F401
    from layerwise_b200 import _native
    from layerwise_b200.parallel.exchange import CompressionConfig
    from layerwise_b200.train.imagenet import Trainer
    _native.require()
    cfg = CompressionConfig(compress=args.compress, method=args.method, ratio=args.ratio,
                            error_feedback=not args.no_ef, nvls=args.nvls)
    tr = Trainer(arch=args.arch, dtype="bf16", cfg=cfg, lr=0.1, momentum=0.9, weight_decay=1e-4,
                 overlap=not args.no_overlap, engine=args.engine, device=torch.device("cuda", torch.cuda.current_device()))
    return tr


def build_reference(args, world, rank):
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "IMAGENET", "training")):
        raise FileNotFoundError("baseline/_ref missing (cp -r /root/reference/. baseline/_ref)")
    sys.path.insert(0, ref)
    sys.path.insert(0, os.path.join(ref, "IMAGENET", "training"))
    import resnet as ref_resnet                       # IMAGENET/training/resnet.py (unmodified)
    import fp16util as ref_fp16                        # IMAGENET/training/fp16util.py (unmodified)
    from CIFAR10.core import layerwise_compressed_comm, entiremodel_compressed_comm  # noqa: F401
    torch.backends.cudnn.benchmark = True             # train_imagenet_nv.py:95
    torch.manual_seed(2147483647)                     # :117-118
    model = ref_resnet.resnet50(bn0=True).cuda()
    fp16 = args.ref_dtype == "fp16"
    if fp16:
        model = ref_fp16.network_to_half(model)
        model_params, master_params = ref_fp16.prep_param_lists(model)
    else:
        model_params = master_params = list(model.parameters())
    criterion = torch.nn.CrossEntropyLoss().cuda()
    opt = torch.optim.SGD(master_params, 0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)
    loss_scale = 1024.0
    mean = torch.tensor([0.485 * 255, 0.456 * 255, 0.406 * 255]).cuda().view(1, 3, 1, 1)
    std = torch.tensor([0.229 * 255, 0.224 * 255, 0.225 * 255]).cuda().view(1, 3, 1, 1)
    if fp16:
        mean, std = mean.half(), std.half()
    wsize = float(world)

    def exchange():
        if args.compress == "layerwise":
            layerwise_compressed_comm(model, wsize, args.method, args.ratio, 0.001, 255)
        else:       # --compress none: the reference's plain per-tensor all-reduce (train_imagenet_nv.py:382-386)
            for param in model.parameters():
                dist.all_reduce(param.grad.data)
                param.grad.data /= wsize

    def step(images_u8, target):
        # BatchTransformDataLoader.process_tensors (dataloader.py:88-93)
        x = images_u8.cuda(non_blocking=True)
        x = x.half() if fp16 else x.float()
        x = x.sub_(mean).div_(std)
        t = target.cuda(non_blocking=True)
        output = model(x)
        loss = criterion(output, t)
        if fp16:                                       # train_imagenet_nv.py:411-428
            loss = loss * loss_scale
            model.zero_grad()
            loss.backward()
            exchange()
            ref_fp16.model_grads_to_master_grads(model_params, master_params)
            for p in master_params:
                p.grad.data = p.grad.data / loss_scale
            opt.step()
            ref_fp16.master_params_to_model_params(model_params, master_params)
            loss = loss / loss_scale
        else:                                          # :429-441
            opt.zero_grad()
            loss.backward()
            exchange()
            opt.step()
        return loss
    return step


CIFAR_ARCHS = ("resnet9", "alexnet", "alexnet_graph", "vgg16")


class _CifarStep:
    """Adapter: one CIFAR training step (dawn.py recipe) behind the ``train_step(images, target) -> (out, loss)``
    interface that the timing code drives.  ``images``: float32 NCHW batch (host pinned or device), as the
    reference's ``Batches`` hands them over (CIFAR10/torch_backend.py:57-60)."""

    def __init__(self, args, world, rank):
        import layerwise_b200  # noqa: F401
        from layerwise_b200.models import build_cifar_model
        from layerwise_b200 import _native
        from layerwise_b200.models.graph import trainable_params
        from layerwise_b200.ops.sgd import SGD
        from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger
        from layerwise_b200.train.cifar import GraphedTrainStep
        from layerwise_b200.utils.mixed import bn_to_float
        _native.require()
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.dtype = {"fp32": torch.float32, "bf16": torch.bfloat16}[args.cifar_dtype]
        name = {"alexnet": "alexnet1", "alexnet_graph": "alexnet"}.get(args.arch, args.arch)
        model = build_cifar_model(name).to(self.device).to(memory_format=torch.channels_last)
        if self.dtype != torch.float32:
            model = bn_to_float(model.to(self.dtype))
        self.model = model
        bs = args.batch
        cfg = CompressionConfig(compress=args.compress, method=args.method, ratio=args.ratio, qstates=255,
                                threshold=0.001, error_feedback=not args.no_ef, nvls=args.nvls)
        self.exchanger = GradientExchanger(trainable_params(model), cfg, engine=args.engine)
        self.opt = SGD(trainable_params(model), lr=lambda step: 0.1 / bs, momentum=0.9, weight_decay=5e-4 * bs, nesterov=True)
        self.graphed = GraphedTrainStep(model, self.exchanger, self.opt) if args.graph else None
        self._graph = None
        self.cuda_graph = False
        self.phase_timer = None

    def _batch(self, images, target):
        x = images.to(self.device, non_blocking=True).to(self.dtype).contiguous(memory_format=torch.channels_last)
        return {"input": x, "target": target.to(self.device, non_blocking=True)}

    def train_step(self, images, target):
        from layerwise_b200.train.cifar import _zero_grads
        batch = self._batch(images, target)
        if self.graphed is not None:
            out = self.graphed(batch)
        else:
            self.model.train(True)
            out = self.model(batch)
            out["loss"].sum().backward()
            self.exchanger.exchange()
            self.opt.step()
            _zero_grads(self.model, self.exchanger)
        return None, out["loss"].detach().sum()


def build_reference_cifar(args, world, rank):
    """The reference's own CIFAR-10 code (CIFAR10/core.py run_batches body :303-321 with its models, its
    TorchOptimiser/SGD and its layerwise_compressed_comm), fp32 as committed.  ``entiremodel`` raises TypeError in the
    reference as committed (``for param in model.parameters:``, core.py:230) and is reported as unavailable."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "CIFAR10")):
        raise FileNotFoundError("baseline/_ref missing (cp -r /root/reference/. baseline/_ref)")
    sys.path.insert(0, ref)
    from CIFAR10 import core as rcore
    from CIFAR10 import torch_backend as rtb
    if args.arch == "alexnet":
        from CIFAR10.alexnet import AlexNet
        model = AlexNet().cuda()
    elif args.arch == "vgg16":
        from CIFAR10.vgg16 import vgg16model
        model = vgg16model.cuda()
    else:
        # dawn.py runs main() at import: execute only its definitions (the text above ``def main``), unmodified
        src = open(os.path.join(ref, "CIFAR10", "dawn.py")).read().split("def main")[0]
        ns = {}
        exec(compile(src, "dawn_defs", "exec"), ns)
        builder = ns["resnet9"] if args.arch == "resnet9" else ns["alexnet"]
        model = rtb.Network(rcore.union(builder(), ns["losses"])).cuda()
    bs = args.batch
    opt = rtb.SGD(rtb.trainable_params(model), lr=lambda step: 0.1 / bs, momentum=0.9, weight_decay=5e-4 * bs, nesterov=True)
    wsize = float(world)
    model.train(True)

    def step(images, target):
        batch = {"input": images.cuda(non_blocking=True), "target": target.cuda(non_blocking=True).long()}
        output = model(batch)
        output["loss"].sum().backward()
        if args.compress == "layerwise":
            rcore.layerwise_compressed_comm(model, wsize, args.method, args.ratio, 0.001, 255)
        elif args.compress == "entiremodel":
            rcore.entiremodel_compressed_comm(model, wsize, args.method, args.ratio, 0.001, 255)
        else:
            for layer in model.parameters():
                dist.all_reduce(layer.grad.data)
                layer.grad.data /= wsize
        opt.step()
        model.zero_grad()
        return output["loss"].sum()
    return step


def _phase_summary(tr, batch, steps):
    from layerwise_b200.utils.timing import PhaseTimer
    graph, tr._graph = tr._graph, None
    cg, tr.cuda_graph = tr.cuda_graph, False
    tr.phase_timer = PhaseTimer()
    try:
        for _ in range(steps):
            tr.train_step(batch[0], batch[1])
        return tr.phase_timer.summary()
    finally:
        tr.phase_timer = None
        tr._graph, tr.cuda_graph = graph, cg


def main():
    args = parse()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device"}))
        return 0
    world, rank, local = setup_dist(args)
    cifar = args.arch in CIFAR_ARCHS
    if cifar and args.size == 224:
        args.size = 32
        if args.batch == 224:
            args.batch = 512
    B, S = args.batch, args.size
    g = torch.Generator().manual_seed(1234 + rank)
    nbuf = 4
    if cifar:       # normalised float32 batches, 10 classes (what the reference's Batches yields)
        host_x = [torch.randn(B, 3, S, S, generator=g).pin_memory() for _ in range(nbuf)]
        host_y = [torch.randint(0, 10, (B,), generator=g).pin_memory() for _ in range(nbuf)]
    else:
        host_x = [torch.randint(0, 256, (B, 3, S, S), dtype=torch.uint8, generator=g).pin_memory() for _ in range(nbuf)]
        host_y = [torch.randint(0, 1000, (B,), generator=g).pin_memory() for _ in range(nbuf)]
    dev_x = [t.cuda() for t in host_x]
    dev_y = [t.cuda() for t in host_y]

    launches = {"n": 0}
    if args.impl == "reference":
        try:
            ref_step = build_reference_cifar(args, world, rank) if cifar else build_reference(args, world, rank)
            ref_step(dev_x[0], dev_y[0])               # the reference as committed may raise here (entiremodel: D1)
            torch.cuda.synchronize()
        except Exception as e:      # noqa: BLE001
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]}))
            return 0
        step_dev = lambda i: ref_step(dev_x[i % nbuf], dev_y[i % nbuf])      # noqa: E731
        step_e2e = lambda i: ref_step(host_x[i % nbuf], host_y[i % nbuf])    # noqa: E731
        dtype, par = ("fp32" if cifar else args.ref_dtype), f"dp{world}"
    elif cifar:
        tr = _CifarStep(args, world, rank)
        step_dev = lambda i: tr.train_step(dev_x[i % nbuf], dev_y[i % nbuf])[1]     # noqa: E731
        step_e2e = lambda i: tr.train_step(host_x[i % nbuf], host_y[i % nbuf])[1]   # noqa: E731
        dtype, par = args.cifar_dtype, f"dp{world}"
    else:
        tr = build_ours(args, world, rank)
        args.graph = args.graph and hasattr(tr.exchanger.engine, "enable_graph_mode")
        if args.graph:
            tr.capture_graph(dev_x[0], dev_y[0], warmup=3)
        step_dev = lambda i: tr.train_step(dev_x[i % nbuf], dev_y[i % nbuf])[1]     # noqa: E731
        step_e2e = lambda i: tr.train_step(host_x[i % nbuf], host_y[i % nbuf])[1]   # noqa: E731
        dtype, par = "bf16", f"dp{world}"

    # (the CIFAR step captures its CUDA graph at its 4th call of a shape: keep the capture out of the timed region)
    for i in range(args.warmup + (4 if (cifar and args.impl == "ours" and args.graph) else 0)):
        step_dev(i)
    torch.cuda.synchronize()
    if args.impl == "ours":
        from layerwise_b200 import _native
        _native.counters.reset()
    ms, clocks = timed(step_dev, args.steps, ClockSampler(local))
    if args.impl == "ours":
        launches["n"] = _native.counters.kernels
    imgs = B * world * args.steps
    value = imgs / (ms / 1e3)

    e2e = None
    if not args.skip_e2e:
        sink = []

        if args.impl == "ours":
            # public data path of the package: pinned host batches -> DevicePrefetcher (H2D of batch i+1 on a copy
            # stream while step i computes) -> Trainer.train_step.  Every step still moves its own 33.7 MB batch
            # host->device inside the timed region and reads its loss back.
            from layerwise_b200.data import DevicePrefetcher

            def host_batches(n):
                for i in range(n):
                    yield host_x[i % nbuf], host_y[i % nbuf]
            pf = iter(DevicePrefetcher(host_batches(args.steps + 2), torch.device("cuda", local)))

            def step_e2e(i):      # noqa: F811
                xd, yd = next(pf)
                return tr.train_step(xd, yd)[1]

        reader = None
        if args.impl == "ours":
            # every step's loss is copied device -> host (4 bytes, pinned) and read by the host, one step late: the host
            # never waits for the step it has just launched (utils/hostread.py).  The last value is drained before the
            # closing event, so all `steps` reads are inside the timed region.
            from layerwise_b200.utils import LaggedReader
            reader = LaggedReader(depth=1)

        def e2e_step(i, last=False):
            loss = step_e2e(i)
            if reader is None:
                sink.append(float(loss.detach() if hasattr(loss, 'detach') else loss))    # device -> host read of the step's result
                return
            v = reader.push(loss)
            if v is not None:
                sink.append(v)
            if last:
                sink.extend(reader.drain())
        for i in range(2):
            e2e_step(i, last=(i == 1))
        n_before = len(sink)
        ms2, _ = timed(lambda i: e2e_step(i, last=(i == args.steps - 1)), args.steps)
        assert len(sink) - n_before == args.steps, "every timed step's loss must have been read on the host"
        e2e = {"value": imgs / (ms2 / 1e3), "unit": "img/s",
               "h2d_bytes_per_step": host_x[0].numel() * host_x[0].element_size() + host_y[0].numel() * 8,
               "d2h_bytes_per_step": 4,
               "ms_per_step": ms2 / args.steps}

    extra = {}
    if args.impl == "ours" and hasattr(tr.exchanger.engine, "benchmark_exchange"):
        # compress + exchange + decompress alone (no backward to overlap with), device-timed, max over ranks
        extra["compress_comm_ms_per_step"] = max_over_ranks(tr.exchanger.engine.benchmark_exchange(10))
        extra["exchange_stats"] = tr.exchanger.engine.stats()
    if args.impl == "ours" and args.phase_times and hasattr(tr, "profile_phases"):
        # eager steps timed phase by phase with CUDA events (max over ranks): explains the number, is not the number
        tr.profile_phases(dev_x[0], dev_y[0], steps=2)           # warm the eager path
        extra["phases_ms_per_step"] = {k: v / 5 for k, v in _phase_summary(tr, (dev_x[0], dev_y[0]), 5).items()}
    if rank == 0:
        out = {
            "metric": ("ResNet-50 img/s, layerwise Top-k 1% (train step incl. compress+exchange+SGD)"
                       if (args.arch == "resnet50" and args.method == "Topk" and args.compress == "layerwise") else
                       f"{args.arch} img/s, {args.compress} {args.method} {args.ratio:g} (train step incl. compress+exchange+SGD)"),
            "value": value, "unit": "img/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / BASELINE_PUBLISHED) if BASELINE_PUBLISHED else None,
            "dtype": dtype,
            "data": (f"synthetic (float32 {S}x{S} normalised images, 10 classes, random-init weights)" if cifar
                     else f"synthetic (uint8 {S}x{S} images, random-init weights)"),
            "impl": args.impl,
            "config": {"model": args.arch, "global_batch": B * world, "per_gpu_batch": B, "seq_len": None,
                       "image_size": S, "parallelism": par, "compress": args.compress, "method": args.method,
                       "ratio": args.ratio, "error_feedback": (not args.no_ef) if args.impl == "ours" else False,
                       "l2_policy": "inputs+activations per step (>1 GB) exceed the 126 MB L2; 4 rotating input buffers",
                       "optimizer": ("SGD momentum 0.9 nesterov wd 5e-4*bs, lr/bs (dawn.py recipe)" if cifar
                                     else "SGD momentum 0.9 nesterov wd 1e-4, fp32 masters"),
                       "cuda_graph": bool(args.graph and args.impl == "ours")},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches["n"],
        }
        out.update(extra)
        print(json.dumps(out))
    if dist.is_initialized():
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
