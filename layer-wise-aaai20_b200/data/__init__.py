"""Data pipelines: CIFAR-10 (host and device-resident), ImageNet loaders, synthetic sets."""
from .cifar import (cifar10, synthetic_cifar10, normalise, pad, transpose, Crop, FlipLR, Cutout,  # noqa: F401
                    Transform, Batches, DeviceBatches, cifar_batches, cifar10_mean, cifar10_std)
