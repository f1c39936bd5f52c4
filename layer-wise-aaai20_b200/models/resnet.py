"""ImageNet ResNet-18/34/50/101/152.

Same topology as the reference's network (IMAGENET/training/resnet.py:1-220):
7x7/2 stem, 3x3/2 max-pool, four stages, global average pool, fc; the bottleneck
places its stride on the 3x3 convolution (resnet.py:63); Kaiming-normal (fan_out)
conv init; with ``bn0=True`` the last batch-norm scale of every residual block
starts at zero and the classifier weights are drawn from N(0, 0.01)
(``init_dist_weights``, resnet.py:154-160).  ResNet-50: 161 parameter tensors,
25 557 032 parameters.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import nn as lnn

__all__ = ["ResNet", "resnet18", "resnet34", "resnet50", "resnet101", "resnet152"]


def _stats_on():
    from ..ops import conv as _conv
    return _conv._FUSE_STATS["on"]


def _bn_fed(conv):
    """Every convolution of these networks feeds a batch norm: its epilogue sums the batch statistics
    (``ops/conv.py``), so the batch norm runs no statistics pass of its own."""
    conv.emit_bn_stats = True
    return conv


def conv3x3(c_in, c_out, stride=1):
    return _bn_fed(lnn.Conv2d(c_in, c_out, kernel_size=3, stride=stride, padding=1, bias=False))


def conv1x1(c_in, c_out, stride=1):
    return _bn_fed(lnn.Conv2d(c_in, c_out, kernel_size=1, stride=stride, bias=False))


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = conv3x3(inplanes, planes, stride)
        self.bn1 = lnn.BatchNorm2d(planes)
        self.relu = lnn.ReLU(inplace=True)
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = lnn.BatchNorm2d(planes)
        self.downsample = downsample
        self.stride = stride

    @property
    def last_bn(self):
        return self.bn2

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x), relu=True, fuse_bwd=True)
        return self.bn2(self.conv2(out), residual=identity, relu=True)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = conv1x1(inplanes, planes)
        self.bn1 = lnn.BatchNorm2d(planes)
        self.conv2 = conv3x3(planes, planes, stride)
        self.bn2 = lnn.BatchNorm2d(planes)
        self.conv3 = conv1x1(planes, planes * self.expansion)
        self.bn3 = lnn.BatchNorm2d(planes * self.expansion)
        self.relu = lnn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    @property
    def last_bn(self):
        return self.bn3

    def forward(self, x):
        # the block input forks into conv1 and the skip connection (plain or through the downsample convolution):
        # conv1's data gradient runs on the tcgen05 kernel whose epilogue adds the skip gradient (ops/conv.py)
        from ..ops.conv import fork_conv1x1
        out, identity = fork_conv1x1(self.conv1, x)
        if self.downsample is not None:
            identity = self.downsample(identity)
        # bn1 / bn2 feed exactly one convolution each: its data gradient sums their backward reductions (fuse_bwd)
        out = self.bn1(out, relu=True, fuse_bwd=True)
        out = self.bn2(self.conv2(out), relu=True, fuse_bwd=True)
        return self.bn3(self.conv3(out), residual=identity, relu=True)


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes=1000, bn0=False):
        super().__init__()
        self.inplanes = 64
        self.conv1 = _bn_fed(lnn.Conv2d(3, 64, kernel_size=7, stride=2, padding=3, bias=False))
        self.bn1 = lnn.BatchNorm2d(64)
        self.relu = lnn.ReLU(inplace=True)
        self.maxpool = lnn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = lnn.AdaptiveAvgPool2d(1)
        self.fc = lnn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.modules.batchnorm._BatchNorm):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
        if bn0:
            self.init_dist_weights()

    def init_dist_weights(self):
        for m in self.modules():
            if isinstance(m, (BasicBlock, Bottleneck)):
                nn.init.zeros_(m.last_bn.weight)
            elif isinstance(m, nn.Linear):
                nn.init.normal_(m.weight, 0, 0.01)

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(conv1x1(self.inplanes, planes * block.expansion, stride),
                                       lnn.BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        layers += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def stem(self, x):
        """7x7/2 stem.  A 16-channel input is the 2x2 space-to-depth form produced by
        ``ops.input.normalize_u8(..., s2d=True)``: the same convolution expressed as a dense 4x4/1 one
        (K = 256 instead of a 3-channel 7x7 window), which maps onto the tensor cores far better."""
        if x.shape[1] == 16:
            from ..ops.input import stem_s2d
            return stem_s2d(x, self.conv1.weight, want_stats=self.training and torch.is_grad_enabled() and _stats_on())
        return self.conv1(x)

    def forward(self, x):
        x = self.maxpool(self.bn1(self.stem(x), relu=True))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = self.avgpool(x)
        return self.fc(x.reshape(x.size(0), -1))


def resnet18(**kw):
    return ResNet(BasicBlock, [2, 2, 2, 2], **kw)


def resnet34(**kw):
    return ResNet(BasicBlock, [3, 4, 6, 3], **kw)


def resnet50(**kw):
    return ResNet(Bottleneck, [3, 4, 6, 3], **kw)


def resnet101(**kw):
    return ResNet(Bottleneck, [3, 4, 23, 3], **kw)


def resnet152(**kw):
    return ResNet(Bottleneck, [3, 8, 36, 3], **kw)
