"""DAWNBench ResNet-9 and the graph-style AlexNet, expressed in the network DSL.

Architecture facts (reference CIFAR10/dawn.py:22-87): conv3x3(no bias)
+ BN + ReLU blocks 3->64, 64->128 (+pool, +residual), 128->256 (+pool), 256->512
(+pool, +residual), maxpool 4, linear 512->10 without bias, logits scaled by 0.125;
25 parameter tensors, 6 573 120 parameters.  The AlexNet graph variant uses a
stride-2 first conv and channels 64/192/384/256/256 (16 tensors, 2 255 296 params).
"""
from __future__ import annotations

import torch.nn as nn

from ..ops import nn as lnn
from .graph import Add, Correct, Flatten, Identity, Mul, Network, rel_path, union


def batch_norm(num_channels, bn_bias_init=None, bn_bias_freeze=False, bn_weight_init=None,
               bn_weight_freeze=False):
    m = lnn.BatchNorm2d(num_channels)
    if bn_bias_init is not None:
        m.bias.data.fill_(bn_bias_init)
    if bn_bias_freeze:
        m.bias.requires_grad = False
    if bn_weight_init is not None:
        m.weight.data.fill_(bn_weight_init)
    if bn_weight_freeze:
        m.weight.requires_grad = False
    return m


def conv_bn(c_in, c_out, bn_weight_init=1.0, stride=1, fuse=True, **kw):
    """conv3x3 -> BN -> ReLU.  With ``fuse`` the ReLU runs inside the batch-norm kernel: the ``relu`` node stays in
    the graph (same node names / state_dict keys as the reference) but is a no-op, and the ``bn`` entry of the
    activation cache then holds the post-ReLU tensor."""
    bn = batch_norm(c_out, bn_weight_init=bn_weight_init, **kw)
    relu = lnn.ReLU(True)
    bn.fuse_relu = bool(fuse)
    relu.already_applied = bool(fuse)
    conv = lnn.Conv2d(c_in, c_out, kernel_size=3, stride=stride, padding=1, bias=False)
    conv.emit_bn_stats = True              # the batch-norm statistics come out of the convolution epilogue
    return {
        "conv": conv,
        "bn": bn,
        "relu": relu,
    }


def conv_bn_stride(c_in, c_out, bn_weight_init=1.0, **kw):
    return conv_bn(c_in, c_out, bn_weight_init=bn_weight_init, stride=2, **kw)


def residual(c, **kw):
    return {
        "in": Identity(),
        "res1": conv_bn(c, c, **kw),
        "res2": conv_bn(c, c, **kw),
        "add": (Add(), [rel_path("in"), rel_path("res2", "relu")]),
    }


def basic_resnet9(channels, weight, pool, **kw):
    return {
        "prep": conv_bn(3, channels["prep"], **kw),
        "layer1": dict(conv_bn(channels["prep"], channels["layer1"], **kw), pool=pool),
        "layer2": dict(conv_bn(channels["layer1"], channels["layer2"], **kw), pool=pool),
        "layer3": dict(conv_bn(channels["layer2"], channels["layer3"], **kw), pool=pool),
        "pool": lnn.MaxPool2d(4),
        "flatten": Flatten(),
        "linear": lnn.Linear(channels["layer3"], 10, bias=False),
        "classifier": Mul(weight),
    }


def resnet9(channels=None, weight=0.125, pool=None, extra_layers=(),
            res_layers=("layer1", "layer3"), **kw):
    channels = channels or {"prep": 64, "layer1": 128, "layer2": 256, "layer3": 512}
    pool = pool if pool is not None else lnn.MaxPool2d(2)
    net = basic_resnet9(channels, weight, pool, **kw)
    for layer in res_layers:
        net[layer]["residual"] = residual(channels[layer], **kw)
    for layer in extra_layers:
        net[layer]["extra"] = conv_bn(channels[layer], channels[layer], **kw)
    return net


def basic_alexnet(channels, weight, pool, **kw):
    return {
        "prep": dict(conv_bn_stride(3, channels["prep"], **kw), pool=pool),
        "layer1": dict(conv_bn(channels["prep"], channels["layer1"], **kw), pool=pool),
        "layer2": conv_bn(channels["layer1"], channels["layer2"], **kw),
        "layer3": conv_bn(channels["layer2"], channels["layer3"], **kw),
        "layer4": dict(conv_bn(channels["layer3"], channels["layer4"], **kw), pool=pool),
        "pool": lnn.MaxPool2d(2),
        "flatten": Flatten(),
        "linear": lnn.Linear(channels["layer4"], 10, bias=False),
        "classifier": Mul(weight),
    }


def alexnet(channels=None, weight=0.125, pool=None, extra_layers=(), **kw):
    channels = channels or {"prep": 64, "layer1": 192, "layer2": 384, "layer3": 256, "layer4": 256}
    pool = pool if pool is not None else lnn.MaxPool2d(2)
    return basic_alexnet(channels, weight, pool, **kw)


def losses():
    """Per-sample cross entropy (``reduction='none'``) and the ``correct`` indicator."""
    return {
        "loss": (lnn.CrossEntropyLoss(reduction="none"), [("classifier",), ("target",)]),
        "correct": (Correct(), [("classifier",), ("target",)]),
    }


def resnet9_network(**kw) -> Network:
    return Network(union(resnet9(**kw), losses()))


def alexnet_graph_network(**kw) -> Network:
    return Network(union(alexnet(**kw), losses()))
