"""CIFAR-10 data path: preprocessing, augmentation, batching, synthetic stand-in.

Capabilities of the reference pipeline (CIFAR10/core.py:43-114 and
CIFAR10/torch_backend.py:36-63): per-channel normalisation with the CIFAR mean/std
scaled to 0..255, 4-pixel reflect padding, NHWC->NCHW, and the Crop / FlipLR /
Cutout augmentations whose random choices are drawn once per epoch for the whole
dataset.  Two batchers are provided:

* ``Batches`` -- host pipeline over a ``DataLoader`` (pinned memory, H2D per batch),
  the reference's structure.
* ``DeviceBatches`` -- the B200-first path: the whole (padded, normalised) dataset
  lives in HBM (50k x 3 x 40 x 40 fp16 = 480 MB) and crop / flip / cutout are batched
  gather ops on the device, so an epoch issues no host->device traffic at all.

There is no network in this environment, so ``cifar10()`` falls back to a synthetic
dataset of the same shape when the files are absent (``synthetic_cifar10``).
"""
from __future__ import annotations

import os
from collections import namedtuple
from typing import Dict, Optional

import numpy as np
import torch

cifar10_mean = (0.4914, 0.4822, 0.4465)
cifar10_std = (0.2471, 0.2435, 0.2616)


def normalise(x, mean=cifar10_mean, std=cifar10_std):
    x = np.array(x, np.float32)
    mean = np.array(mean, np.float32) * 255.0
    inv = 1.0 / (255.0 * np.array(std, np.float32))
    return (x - mean) * inv


def pad(x, border=4):
    return np.pad(x, [(0, 0), (border, border), (border, border), (0, 0)], mode="reflect")


def transpose(x, source="NHWC", target="NCHW"):
    return x.transpose([source.index(d) for d in target])


# ---- augmentations (host side, per sample) ------------------------------------------------------
class Crop(namedtuple("Crop", ("h", "w"))):
    def __call__(self, x, x0, y0):
        return x[:, y0:y0 + self.h, x0:x0 + self.w]

    def options(self, x_shape):
        _, H, W = x_shape
        return {"x0": range(W + 1 - self.w), "y0": range(H + 1 - self.h)}

    def output_shape(self, x_shape):
        return (x_shape[0], self.h, self.w)


class FlipLR(namedtuple("FlipLR", ())):
    def __call__(self, x, choice):
        return x[:, :, ::-1].copy() if choice else x

    def options(self, x_shape):
        return {"choice": [True, False]}


class Cutout(namedtuple("Cutout", ("h", "w"))):
    def __call__(self, x, x0, y0):
        x = x.copy()
        x[:, y0:y0 + self.h, x0:x0 + self.w] = 0.0
        return x

    def options(self, x_shape):
        _, H, W = x_shape
        return {"x0": range(W + 1 - self.w), "y0": range(H + 1 - self.h)}


class Transform:
    """Dataset wrapper applying transforms with per-epoch pre-drawn random choices."""

    def __init__(self, dataset, transforms):
        self.dataset, self.transforms = dataset, transforms
        self.choices = None

    def __len__(self):
        return len(self.dataset)

    def __getitem__(self, index):
        data, label = self.dataset[index]
        for choices, f in zip(self.choices, self.transforms):
            data = f(data, **{k: v[index] for k, v in choices.items()})
        return data, label

    def set_random_choices(self, rng: Optional[np.random.RandomState] = None):
        rng = rng or np.random
        self.choices = []
        shape = self.dataset[0][0].shape
        n = len(self)
        for t in self.transforms:
            opts = t.options(shape)
            shape = t.output_shape(shape) if hasattr(t, "output_shape") else shape
            self.choices.append({k: rng.choice(v, size=n) for k, v in opts.items()})


# ---- datasets -----------------------------------------------------------------------------------------
def synthetic_cifar10(n_train=50000, n_test=10000, seed=0) -> Dict:
    """Random uint8 images / labels with CIFAR-10's shapes (NHWC)."""
    rng = np.random.RandomState(seed)
    return {
        "train": {"data": rng.randint(0, 256, (n_train, 32, 32, 3), dtype=np.uint8),
                  "labels": rng.randint(0, 10, n_train).tolist()},
        "test": {"data": rng.randint(0, 256, (n_test, 32, 32, 3), dtype=np.uint8),
                 "labels": rng.randint(0, 10, n_test).tolist()},
    }


def cifar10(root: str, download: bool = False, synthetic_fallback: bool = True) -> Dict:
    """CIFAR-10 as numpy NHWC uint8 + label lists (torchvision reader when files exist)."""
    try:
        import torchvision
        have = os.path.isdir(os.path.join(root, "cifar-10-batches-py"))
        if have or download:
            tr = torchvision.datasets.CIFAR10(root=root, train=True, download=download)
            te = torchvision.datasets.CIFAR10(root=root, train=False, download=download)
            return {"train": {"data": tr.data, "labels": tr.targets},
                    "test": {"data": te.data, "labels": te.targets}}
    except Exception:
        if not synthetic_fallback:
            raise
    if not synthetic_fallback:
        raise FileNotFoundError(f"CIFAR-10 not found under {root}")
    return synthetic_cifar10()


# ---- batchers -----------------------------------------------------------------------------------------
class Batches:
    """Host ``DataLoader`` batches moved to ``device`` as ``{'input','target'}`` dicts."""

    def __init__(self, dataset, batch_size, shuffle, set_random_choices=False, num_workers=0,
                 drop_last=False, device=None, dtype=None):
        self.dataset = dataset
        self.batch_size = batch_size
        self.set_random_choices = set_random_choices
        self.device = device or torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.dtype = dtype
        self.dataloader = torch.utils.data.DataLoader(
            dataset, batch_size=batch_size, num_workers=num_workers,
            pin_memory=torch.cuda.is_available(), shuffle=shuffle, drop_last=drop_last)

    def __iter__(self):
        if self.set_random_choices:
            self.dataset.set_random_choices()
        for x, y in self.dataloader:
            x = x.to(self.device, non_blocking=True)
            if self.dtype is not None:
                x = x.to(self.dtype)
            yield {"input": x, "target": y.to(self.device, non_blocking=True).long()}

    def __len__(self):
        return len(self.dataloader)


class DeviceBatches:
    """Whole dataset resident on the device; augmentation as batched gathers.

    ``data`` is NCHW float (already normalised and, for training, reflect-padded).
    ``crop`` = output H=W; ``flip`` and ``cutout`` (square size or 0) toggle the other
    two augmentations.  Choices are drawn once per epoch like ``Transform``.
    """

    def __init__(self, data, labels, batch_size, shuffle, crop=None, flip=False, cutout=0,
                 drop_last=False, device=None, dtype=torch.float32, seed=0,
                 channels_last=False):
        self.device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
        self.data = torch.as_tensor(np.ascontiguousarray(data)).to(self.device, dtype)
        self.labels = torch.as_tensor(np.asarray(labels)).to(self.device).long()
        self.batch_size, self.shuffle, self.drop_last = batch_size, shuffle, drop_last
        self.crop, self.flip, self.cutout = crop, flip, cutout
        self.channels_last = channels_last
        self.gen = torch.Generator(device=self.device)
        self.gen.manual_seed(seed)

    def __len__(self):
        n = self.data.size(0)
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)

    def _rand(self, hi, n):
        return torch.randint(0, hi, (n,), device=self.device, generator=self.gen)

    def __iter__(self):
        N, C, H, W = self.data.shape
        order = (torch.randperm(N, device=self.device, generator=self.gen) if self.shuffle
                 else torch.arange(N, device=self.device))
        h = w = self.crop or H
        if self.crop:
            cx, cy = self._rand(W + 1 - w, N), self._rand(H + 1 - h, N)
        if self.flip:
            fl = self._rand(2, N).bool()
        if self.cutout:
            ux, uy = self._rand(w + 1 - self.cutout, N), self._rand(h + 1 - self.cutout, N)
        ar_h = torch.arange(h, device=self.device)
        ar_w = torch.arange(w, device=self.device)
        for b in range(len(self)):
            idx = order[b * self.batch_size:(b + 1) * self.batch_size]
            x = self.data[idx]
            if self.crop or self.flip:
                rows = (cy[idx] if self.crop else 0 * idx)[:, None] + ar_h[None, :]         # B,h
                cols = ar_w[None, :].expand(idx.numel(), w)
                if self.flip:
                    cols = torch.where(fl[idx][:, None], (w - 1) - cols, cols)
                cols = cols + (cx[idx] if self.crop else 0 * idx)[:, None]                    # B,w
                bi = torch.arange(idx.numel(), device=self.device)[:, None, None]
                x = x.permute(0, 2, 3, 1)[bi, rows[:, :, None], cols[:, None, :]].permute(0, 3, 1, 2)
            if self.cutout:
                yy = ar_h[None, :, None]
                xx = ar_w[None, None, :]
                y0, x0 = uy[idx][:, None, None], ux[idx][:, None, None]
                hole = (yy >= y0) & (yy < y0 + self.cutout) & (xx >= x0) & (xx < x0 + self.cutout)
                x = x.masked_fill(hole[:, None, :, :], 0.0)
            x = x.contiguous(memory_format=torch.channels_last) if self.channels_last else x.contiguous()
            yield {"input": x, "target": self.labels[idx]}


def cifar_batches(dataset: Dict, batch_size=512, device=None, on_device=True, dtype=torch.float32,
                  channels_last=False, seed=0):
    """Train / test batchers with the reference recipe's transforms (dawn.py:112, :132-141)."""
    train_x = transpose(normalise(pad(dataset["train"]["data"], 4)))
    test_x = transpose(normalise(dataset["test"]["data"]))
    if on_device:
        train = DeviceBatches(train_x, dataset["train"]["labels"], batch_size, shuffle=True, crop=32,
                              flip=True, cutout=8, drop_last=True, device=device, dtype=dtype,
                              seed=seed, channels_last=channels_last)
        test = DeviceBatches(test_x, dataset["test"]["labels"], batch_size, shuffle=False,
                             drop_last=False, device=device, dtype=dtype, channels_last=channels_last)
        return train, test
    train_set = list(zip(train_x, dataset["train"]["labels"]))
    test_set = list(zip(test_x, dataset["test"]["labels"]))
    tfm = Transform(train_set, [Crop(32, 32), FlipLR(), Cutout(8, 8)])
    train = Batches(tfm, batch_size, shuffle=True, set_random_choices=True, drop_last=True,
                    device=device, dtype=dtype)
    test = Batches(test_set, batch_size, shuffle=False, drop_last=False, device=device, dtype=dtype)
    return train, test
