#!/usr/bin/env python
"""ImageNet entry point (flags of the reference's training/train_imagenet_nv.py).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 train_imagenet.py /data/imagenet --bf16 \\
        --distributed --init-bn0 --no-bn-wd --compress layerwise --method Topk -K 0.01 --error_feedback
    python train_imagenet.py x --synthetic 2048 --short-epoch --epochs 1 --bf16   # smoke run, no dataset
"""
import sys
import warnings

import layerwise_b200  # noqa: F401
from layerwise_b200.train.imagenet_main import main

if __name__ == "__main__":
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", category=UserWarning)
        main()
    sys.exit(0)
