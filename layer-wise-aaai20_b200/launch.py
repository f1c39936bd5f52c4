"""Single-node launcher: one process per GPU, NCCL over NVLink/NVSwitch.

Replaces the reference's AWS launcher (IMAGENET/train.py:290-453: ncluster job creation, EBS
mounts, hand-built NCCL ring orders for V100 hybrid cube-mesh :159-203, torch.distributed.launch /
mpirun).  On an NVSwitch box every GPU pair has full bandwidth, so no ring/topology tuning
exists here; multi-machine schedules are expressed with ``--phases`` exactly as before
(``schedules`` below mirrors the per-machine-count tables of IMAGENET/train.py:60-149).

    python -m layerwise_b200.launch --nproc 8 -- train_imagenet.py /data --bf16 --distributed ...
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys


def one_machine(lr=1.0):
    from .train.imagenet import default_phases
    return default_phases(lr)


def scaled(lr, bs128, bs224, bs288, ep_scale=1.0):
    s224, s288 = bs224 / bs128, bs288 / bs128
    return [
        {"ep": 0, "sz": 128, "bs": bs128, "trndir": "-sz/160"},
        {"ep": (0, 5), "lr": (lr, lr * 2)},
        {"ep": 5, "lr": lr},
        {"ep": 14, "sz": 224, "bs": bs224, "lr": lr * s224},
        {"ep": 16, "lr": lr / 10 * s224},
        {"ep": 27, "lr": lr / 100 * s224},
        {"ep": 32, "sz": 288, "bs": bs288, "min_scale": 0.5, "rect_val": True, "lr": lr / 100 * s288},
        {"ep": (33, 35), "lr": lr / 1000 * s288},
    ]


# machines -> phases; per-GPU batch sizes shrink as the global batch grows (reference tables)
schedules = {1: one_machine(), 2: scaled(1.5, 256, 192, 128), 4: scaled(3.0, 256, 192, 128),
             8: scaled(6.0, 128, 128, 64), 16: scaled(12.0, 64, 64, 32)}


def nccl_env(debug=False):
    """Environment for the workers.  NVSwitch: nothing to tune; optional NCCL debug output."""
    env = {"NCCL_DEBUG": "INFO" if debug else "VERSION"}
    return env


def main(argv=None):
    ap = argparse.ArgumentParser(description="single-node multi-GPU launcher")
    ap.add_argument("--nproc", type=int, default=None, help="processes (GPUs); default: all visible")
    ap.add_argument("--master-port", type=int, default=29500)
    ap.add_argument("--nccl-debug", action="store_true")
    ap.add_argument("--cuda-debug", action="store_true", help="CUDA_LAUNCH_BLOCKING=1")
    ap.add_argument("--machines", type=int, default=1, help="pick the phases schedule for this machine count")
    ap.add_argument("--print-phases", action="store_true")
    ap.add_argument("cmd", nargs=argparse.REMAINDER)
    args = ap.parse_args(argv)
    if args.print_phases:
        print(json.dumps(schedules[args.machines]))
        return 0
    cmd = [c for c in args.cmd if c != "--"]
    if not cmd:
        ap.error("no training command given")
    nproc = args.nproc
    if nproc is None:
        import torch
        nproc = max(torch.cuda.device_count(), 1)
    env = dict(os.environ, **nccl_env(args.nccl_debug))
    if args.cuda_debug:
        env["CUDA_LAUNCH_BLOCKING"] = "1"
    full = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
            "--master-addr", "127.0.0.1", "--master-port", str(args.master_port)] + cmd
    return subprocess.call(full, env=env)


if __name__ == "__main__":
    sys.exit(main())
