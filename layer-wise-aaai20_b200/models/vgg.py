"""VGG family for 32x32 inputs (dict in, dict out), configuration 'D' = VGG-16.

Torchvision-style feature stack (3x3 convs with bias, ReLU, five 2x2 max-pools),
adaptive average pool to 7x7 (a replication from 1x1 for CIFAR), classifier
25088-4096-4096-classes with two dropouts; Kaiming-normal conv init, N(0, 0.01)
linear init (reference CIFAR10/vgg16.py:10-94; VGG-16: 32 tensors, 134 301 514
parameters).  The reference exposes a module-level instance ``vgg16model``; here
``vgg16()`` builds one on demand (a 512 MiB model should not be built at import).
"""
from __future__ import annotations

import torch.nn as nn

from ..ops import nn as lnn
from .graph import Correct

cfgs = {
    "A": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "B": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "D": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "E": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M",
          512, 512, 512, 512, "M"],
}


def make_layers(cfg, batch_norm: bool = False) -> nn.Sequential:
    layers, c_in = [], 3
    for v in cfg:
        if v == "M":
            layers.append(lnn.MaxPool2d(kernel_size=2, stride=2))
            continue
        conv, relu = lnn.Conv2d(c_in, v, kernel_size=3, padding=1), lnn.ReLU(True)
        layers.append(conv)
        if batch_norm:
            conv.emit_bn_stats = True          # (a biased conv does not emit them; kept for bias=False variants)
            layers.append(lnn.BatchNorm2d(v))
        else:
            # bias + ReLU run in the convolution kernel's epilogue; the ReLU node stays for the module names
            conv.fuse_relu = True
            relu.already_applied = True
        layers.append(relu)
        c_in = v
    return nn.Sequential(*layers)


class VGG(nn.Module):
    def __init__(self, features: nn.Module, num_classes: int = 10, init_weights: bool = True,
                 dropout: float = 0.5):
        super().__init__()
        self.features = features
        self.avgpool = lnn.AdaptiveAvgPool2d((7, 7))
        self.classifier = nn.Sequential(
            lnn.Linear(512 * 7 * 7, 4096), lnn.ReLU(True), lnn.Dropout(dropout),
            lnn.Linear(4096, 4096), lnn.ReLU(True), lnn.Dropout(dropout),
            lnn.Linear(4096, num_classes),
        )
        for i in (0, 3):                 # bias + ReLU in the GEMM epilogue; the ReLU nodes stay for the module names
            self.classifier[i].fuse_relu = True
            self.classifier[i + 1].already_applied = True
        self.criterion = lnn.CrossEntropyLoss(reduction="none")
        self.correct = Correct()
        if init_weights:
            self._initialize_weights()

    def logits(self, x):
        x = self.avgpool(self.features(x))
        return self.classifier(x.reshape(x.size(0), -1))

    def forward(self, batch):
        target = batch["target"]
        out = self.logits(batch["input"])
        return {"loss": self.criterion(out, target), "correct": self.correct(out, target)}

    def _initialize_weights(self):
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, nn.modules.batchnorm._BatchNorm):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Linear):
                nn.init.normal_(m.weight, 0, 0.01)
                nn.init.zeros_(m.bias)


def _vgg(cfg: str, batch_norm: bool, **kw) -> VGG:
    return VGG(make_layers(cfgs[cfg], batch_norm=batch_norm), **kw)


def vgg11(**kw):
    return _vgg("A", False, **kw)


def vgg13(**kw):
    return _vgg("B", False, **kw)


def vgg16(**kw):
    return _vgg("D", False, **kw)


def vgg16_bn(**kw):
    return _vgg("D", True, **kw)


def vgg19(**kw):
    return _vgg("E", False, **kw)
