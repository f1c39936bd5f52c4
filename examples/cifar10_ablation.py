#!/usr/bin/env python
"""Architecture ablations of the ResNet-9 family (the reference's experiments.ipynb study), one short run each.

    python examples/cifar10_ablation.py --variant final --epochs 2 --synthetic
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200 import models  # noqa: E402
from layerwise_b200.data import cifar10, cifar_batches, synthetic_cifar10  # noqa: E402
from layerwise_b200.ops.sgd import SGD  # noqa: E402
from layerwise_b200.train.cifar import train  # noqa: E402
from layerwise_b200.utils import PiecewiseLinear, TableLogger, Timer  # noqa: E402

VARIANTS = {
    "backbone": dict(res_layers=(), extra_layers=()),
    "extra_l1": dict(res_layers=(), extra_layers=("layer1",)),
    "extra_all": dict(res_layers=(), extra_layers=("layer1", "layer2", "layer3")),
    "res_l1": dict(res_layers=("layer1",), extra_layers=()),
    "final": dict(res_layers=("layer1", "layer3"), extra_layers=()),
    "double_width": dict(channels={"prep": 128, "layer1": 256, "layer2": 512, "layer3": 1024}),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="final", choices=sorted(VARIANTS))
    ap.add_argument("--epochs", type=int, default=20)
    ap.add_argument("--batch_size", type=int, default=512)
    ap.add_argument("--synthetic", action="store_true")
    ap.add_argument("--data_dir", default="./data")
    ap.add_argument("--max_batches", type=int, default=None)
    args = ap.parse_args()
    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    net = models.Network(models.union(models.resnet9(**VARIANTS[args.variant]), models.losses())).to(dev)
    if dev.type == "cuda":
        net = net.to(memory_format=torch.channels_last)
    ds = synthetic_cifar10(4096, 1024) if args.synthetic else cifar10(args.data_dir)
    trn, tst = cifar_batches(ds, args.batch_size, device=dev, channels_last=dev.type == "cuda")
    sched = PiecewiseLinear([0, max(1, args.epochs // 5), args.epochs], [0, 0.4, 0])
    opt = SGD(models.trainable_params(net), lr=lambda s: sched(s / len(trn)) / args.batch_size, momentum=0.9,
              weight_decay=5e-4 * args.batch_size, nesterov=True)
    train(net, opt, trn, tst, args.epochs, loggers=(TableLogger(),), timer=Timer(torch.cuda.synchronize if dev.type == "cuda" else None),
          max_batches=args.max_batches)


if __name__ == "__main__":
    main()
