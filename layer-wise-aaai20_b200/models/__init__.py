"""Model zoo: graph DSL, ResNet-9, AlexNet (graph + module), VGG, ResNet-18..152."""
from .graph import (Network, build_graph, rel_path, union, path_iter, Identity, Mul, Flatten,  # noqa: F401
                    Add, Concat, Correct, trainable_params, to_dot, remove_by_type)
from .resnet9 import (resnet9, alexnet, losses, conv_bn, conv_bn_stride, residual, batch_norm,  # noqa: F401
                      resnet9_network, alexnet_graph_network)
from .alexnet import AlexNet  # noqa: F401
from .vgg import VGG, make_layers, cfgs, vgg11, vgg13, vgg16, vgg16_bn, vgg19  # noqa: F401
from .resnet import ResNet, resnet18, resnet34, resnet50, resnet101, resnet152  # noqa: F401


def build_cifar_model(name: str):
    """Model selection of the CIFAR entry point, tolerant of the reference's spellings.

    The reference compares ``--network`` against ``Resent9`` / ``Alexnet`` / ``Alexnet1``
    / ``vgg16`` while defaulting to ``resnet9`` (CIFAR10/dawn.py:14, :115-122).
    """
    key = name.strip().lower()
    if key in ("resnet9", "resent9"):
        return resnet9_network()
    if key in ("alexnet", "alexnet_graph"):
        return alexnet_graph_network()
    if key in ("alexnet1", "alexnet_module"):
        return AlexNet()
    if key in ("vgg16", "vgg"):
        return vgg16()
    raise ValueError(f"unknown --network {name!r} (resnet9 | Alexnet | Alexnet1 | vgg16)")
