"""Data pipelines: CIFAR-10 (host and device-resident), ImageNet loaders, synthetic sets."""
from .cifar import (cifar10, synthetic_cifar10, normalise, pad, transpose, Crop, FlipLR, Cutout,  # noqa: F401
                    Transform, Batches, DeviceBatches, cifar_batches, cifar10_mean, cifar10_std)
from .imagenet import (get_loaders, fast_collate, BatchTransformLoader, DistValSampler, sort_ar,  # noqa: F401,E402
                       map_idx2ar, rect_crop_size, CropArTfm, ValDataset, SyntheticImageNet, DevicePrefetcher)
