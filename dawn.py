#!/usr/bin/env python
"""CIFAR-10 entry point with the reference's command line (CIFAR10/dawn.py:8-20).

    python dawn.py --master_address tcp://127.0.0.1:29500 --rank 0 --world_size 2 \\
        --network resnet9 --compress layerwise --method Topk --ratio 0.01

Additions: ``--error_feedback``, ``--epochs``, ``--backend`` (gloo|nccl), ``--engine``
(auto|torch|fused), ``--synthetic`` (no dataset on disk needed), ``--max_batches``
(short epochs for smoke runs), ``--dtype``.  Under ``torchrun`` the rank / world size
come from the environment.
"""
import argparse
import os

import torch

import layerwise_b200  # noqa: F401
from layerwise_b200.data import cifar10, cifar_batches, synthetic_cifar10
from layerwise_b200.models import build_cifar_model, trainable_params
from layerwise_b200.ops.sgd import SGD
from layerwise_b200.parallel import comm
from layerwise_b200.train.cifar import train
from layerwise_b200.utils import PiecewiseLinear, TableLogger, Timer, TSVLogger


def get_parser():
    p = argparse.ArgumentParser(description="CIFAR-10 compressed-gradient training (B200)")
    p.add_argument("--data_dir", type=str, default="./data")
    p.add_argument("--log_dir", type=str, default=".")
    p.add_argument("--master_address", "-m", type=str, default="127.0.0.1")
    p.add_argument("--rank", "-r", type=int, default=None)
    p.add_argument("--world_size", "-w", type=int, default=None)
    p.add_argument("--network", "-n", type=str, default="resnet9")
    p.add_argument("--compress", "-c", type=str, default="none")
    p.add_argument("--method", type=str, default="none")
    p.add_argument("--ratio", "-K", type=float, default=0.5)
    p.add_argument("--threshold", "-V", type=float, default=0.001)
    p.add_argument("--qstates", "-Q", type=int, default=255)
    p.add_argument("--momentum", type=float, default=0.0)
    # additions
    p.add_argument("--error_feedback", action="store_true")
    p.add_argument("--epochs", type=int, default=None)
    p.add_argument("--batch_size", type=int, default=512)
    p.add_argument("--backend", type=str, default=None)
    p.add_argument("--engine", type=str, default="auto")
    p.add_argument("--synthetic", action="store_true")
    p.add_argument("--synthetic_size", type=int, nargs=2, default=[4096, 1024], metavar=("TRAIN", "TEST"),
                   help="images in the synthetic train / test sets (50000 10000 = CIFAR-10's sizes)")
    p.add_argument("--max_batches", type=int, default=None)
    p.add_argument("--cuda_graph", action="store_true",
                   help="capture the training step once and replay it as a CUDA graph (fused engine)")
    p.add_argument("--phase_times", action="store_true",
                   help="print per-phase device times (CUDA events + NVTX ranges, max over ranks) after every epoch")
    p.add_argument("--dtype", type=str, default="fp32", choices=["fp32", "bf16", "fp16"])
    p.add_argument("--host_loader", action="store_true", help="DataLoader pipeline instead of device-resident data")
    p.add_argument("--nchw", action="store_true", help="keep NCHW activations (default on CUDA: channels_last, which the fused kernels need)")
    return p


def main(argv=None):
    args = get_parser().parse_args(argv)
    rank = args.rank if args.rank is not None else comm.env_rank()
    world = args.world_size if args.world_size is not None else comm.env_world_size()
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.backends.cudnn.benchmark = True         # reference: CIFAR10/torch_backend.py:7
        torch.cuda.set_device(comm.env_local_rank() % torch.cuda.device_count())
    device = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")

    dataset = synthetic_cifar10(*args.synthetic_size) if args.synthetic else cifar10(args.data_dir)
    epochs = args.epochs or (40 if args.method in ("Randomk", "Thresholdv") else 24)
    lr_schedule = PiecewiseLinear([0, 5, epochs], [0, 0.4, 0])
    bs = args.batch_size

    model = build_cifar_model(args.network).to(device)
    channels_last = use_cuda and not args.nchw
    if channels_last:
        model = model.to(memory_format=torch.channels_last)
    dtype = {"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}[args.dtype]
    if dtype != torch.float32:
        from layerwise_b200.utils.mixed import bn_to_float
        model = bn_to_float(model.to(dtype))

    timer = Timer(synch=torch.cuda.synchronize if use_cuda else None)
    train_batches, test_batches = cifar_batches(dataset, bs, device=device,
                                                on_device=not args.host_loader, dtype=dtype,
                                                channels_last=channels_last)
    lr = lambda step: lr_schedule(step / len(train_batches)) / bs  # noqa: E731
    if args.momentum > 0:
        opt = SGD(trainable_params(model), lr=lr, momentum=args.momentum, weight_decay=5e-4 * bs,
                  nesterov=True)
    else:
        opt = SGD(trainable_params(model), lr=lr, weight_decay=5e-4 * bs)

    tsv = TSVLogger()
    loggers = (TableLogger(), tsv) if rank == 0 else ()
    master = args.master_address if (world > 1 and "MASTER_ADDR" not in os.environ) else \
        ("env://" if world > 1 else None)
    summary = train(model, opt, train_batches, test_batches, epochs, master, world, rank,
                    loggers=loggers, timer=timer, test_time_in_total=False, compress=args.compress,
                    method=args.method, K=args.ratio, V=args.threshold, qstates=args.qstates,
                    error_feedback=args.error_feedback, backend=args.backend, engine=args.engine,
                    max_batches=args.max_batches, cuda_graph=args.cuda_graph, phase_times=args.phase_times)
    if rank == 0:
        os.makedirs(os.path.expanduser(args.log_dir), exist_ok=True)
        with open(os.path.join(os.path.expanduser(args.log_dir), "logs.tsv"), "w") as f:
            f.write(str(tsv))
    return summary


if __name__ == "__main__":
    main()
