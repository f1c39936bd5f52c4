"""CIFAR-sized AlexNet module (dict in, dict out).

Five 3x3 convolutions with bias (3->64 stride 2, 64->192, 192->384, 384->256,
256->256) with ReLU and three 2x2 max-pools, then Dropout-Linear(1024,4096)-ReLU-
Dropout-Linear(4096,4096)-ReLU-Linear(4096,classes).  ``forward`` consumes
``{'input','target'}`` and returns ``{'loss','correct'}`` with an unreduced loss
(reference CIFAR10/alexnet.py:11-57; 16 tensors, 23 272 266 parameters).
"""
from __future__ import annotations

import torch.nn as nn

from ..ops import nn as lnn
from .graph import Correct

NUM_CLASSES = 10


class AlexNet(nn.Module):
    def __init__(self, num_classes: int = NUM_CLASSES, dropout: float = 0.5):
        super().__init__()
        cfg = [(3, 64, 2, True), (64, 192, 1, True), (192, 384, 1, False), (384, 256, 1, False),
               (256, 256, 1, True)]
        feats = []
        for c_in, c_out, stride, pool in cfg:
            conv, relu = lnn.Conv2d(c_in, c_out, kernel_size=3, stride=stride, padding=1), lnn.ReLU(True)
            conv.fuse_relu = True              # bias + ReLU in the convolution epilogue (ops/conv.py)
            relu.already_applied = True
            feats += [conv, relu]
            if pool:
                feats.append(lnn.MaxPool2d(kernel_size=2))
        self.features = nn.Sequential(*feats)
        self.classifier = nn.Sequential(
            lnn.Dropout(dropout), lnn.Linear(256 * 2 * 2, 4096), lnn.ReLU(True),
            lnn.Dropout(dropout), lnn.Linear(4096, 4096), lnn.ReLU(True),
            lnn.Linear(4096, num_classes),
        )
        for i in (1, 4):                 # bias + ReLU in the GEMM epilogue; the ReLU nodes stay for the module names
            self.classifier[i].fuse_relu = True
            self.classifier[i + 1].already_applied = True
        self.criterion = lnn.CrossEntropyLoss(reduction="none")
        self.correct = Correct()

    def logits(self, x):
        x = self.features(x)
        return self.classifier(x.reshape(x.size(0), -1))

    def forward(self, batch):
        target = batch["target"]
        out = self.logits(batch["input"])
        return {"loss": self.criterion(out, target), "correct": self.correct(out, target)}
