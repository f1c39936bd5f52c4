"""ImageNet data path: folder datasets, uint8 collation, device-side normalisation,
distributed train / validation sharding, rectangular validation, and a synthetic stand-in.

Capabilities of the reference pipeline (IMAGENET/training/dataloader.py:26-201):

* ``get_loaders`` -- RandomResizedCrop(min_scale..1)+flip for training, Resize(1.14x)+CenterCrop
  (or aspect-ratio bucketed rectangular crops) for validation, ``DistributedSampler`` for training,
  equal-batch-count validation sharding;
* ``fast_collate`` -- batches stay uint8 NCHW on the host (4x less H2D traffic than float);
* ``BatchTransformLoader`` -- H2D copy + ``(x - mean) / std`` on the device in the compute dtype;
* ``DistValSampler`` -- every rank sees the same number of batches, trailing ones may be empty;
* ``sort_ar`` / ``map_idx2ar`` / ``CropArTfm`` / ``ValDataset`` -- rectangular validation: images are
  sorted by aspect ratio, each batch is cropped to its mean ratio (sides multiples of 8).

``SyntheticImageNet`` yields random uint8 images of a fixed size so that every code path above (and
the benchmarks) runs without the dataset; there is no network or ImageNet copy in this environment.
"""
from __future__ import annotations

import math
import os
import pickle
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, Sampler
from torch.utils.data.distributed import DistributedSampler

from ..parallel.comm import env_rank, env_world_size
from ..utils.misc import is_set

MEAN = (0.485 * 255, 0.456 * 255, 0.406 * 255)
STD = (0.229 * 255, 0.224 * 255, 0.225 * 255)


def fast_collate(batch):
    """List of (PIL image | HWC/CHW uint8 array, label) -> (uint8 NCHW tensor, int64 labels)."""
    if not batch:
        return torch.tensor([]), torch.tensor([])
    targets = torch.tensor([b[1] for b in batch], dtype=torch.int64)
    first = np.asarray(batch[0][0])
    if first.ndim == 3 and first.shape[0] == 3 and first.shape[-1] != 3:      # already CHW
        h, w = first.shape[1:]
        chw = True
    else:
        h, w = first.shape[:2]
        chw = False
    out = torch.zeros((len(batch), 3, h, w), dtype=torch.uint8)
    for i, (img, _) in enumerate(batch):
        arr = np.asarray(img, dtype=np.uint8)
        if arr.ndim < 3:
            arr = np.repeat(arr[..., None], 3, axis=-1)
        if not chw:
            arr = np.transpose(arr, (2, 0, 1))
        out[i] = torch.from_numpy(np.ascontiguousarray(arr))
    return out, targets


class BatchTransformLoader:
    """Wraps a loader: async H2D of the uint8 batch, then normalisation on the device."""

    def __init__(self, loader, dtype=torch.float32, device=None, channels_last: bool = True):
        self.loader = loader
        self.device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
        self.dtype = dtype
        self.channels_last = channels_last and self.device.type == "cuda"
        self.mean = torch.tensor(MEAN, device=self.device).view(1, 3, 1, 1)
        self.std = torch.tensor(STD, device=self.device).view(1, 3, 1, 1)

    def __len__(self):
        return len(self.loader)

    def process(self, x, y):
        x = x.to(self.device, non_blocking=True)
        y = y.to(self.device, non_blocking=True)
        if x.dim() < 3:          # an empty validation batch
            return x.to(self.dtype), y
        x = x.float().sub_(self.mean).div_(self.std).to(self.dtype)
        if self.channels_last:
            x = x.contiguous(memory_format=torch.channels_last)
        return x, y

    def update_batch_size(self, bs: int):
        self.loader.batch_sampler.batch_size = bs

    def __iter__(self):
        return (self.process(x, y) for x, y in self.loader)


class DistValSampler(Sampler):
    """Contiguous shard per rank with an identical number of (possibly empty) batches."""

    def __init__(self, indices: Sequence[int], batch_size: int, distributed: bool = True,
                 world_size: Optional[int] = None, rank: Optional[int] = None):
        self.indices = list(indices)
        self.batch_size = batch_size
        self.world_size = (world_size if world_size is not None else env_world_size()) if distributed else 1
        self.global_rank = (rank if rank is not None else env_rank()) if distributed else 0
        self.expected_num_batches = math.ceil(len(self.indices) / self.world_size / self.batch_size)
        self.num_samples = self.expected_num_batches * self.batch_size

    def __iter__(self):
        start = self.num_samples * self.global_rank
        mine = self.indices[start:start + self.num_samples]
        for b in range(self.expected_num_batches):
            yield mine[b * self.batch_size:(b + 1) * self.batch_size]

    def __len__(self):
        return self.expected_num_batches

    def set_epoch(self, epoch):
        return


# ---- rectangular validation ------------------------------------------------------------------------------
def sort_ar(valdir: str, sizes: Optional[List[Tuple[int, int]]] = None):
    """[(index, aspect ratio w/h)] sorted by ratio; cached next to the validation folder."""
    cache = os.path.join(valdir, "..", "sorted_idxar.p")
    if sizes is None and os.path.isfile(cache):
        with open(cache, "rb") as f:
            return pickle.load(f)
    if sizes is None:
        from torchvision import datasets
        ds = datasets.ImageFolder(valdir)
        from PIL import Image
        sizes = []
        for path, _ in ds.imgs:
            with Image.open(path) as im:
                sizes.append(im.size)
    idx_ar = sorted(((i, round(w / h, 5)) for i, (w, h) in enumerate(sizes)), key=lambda t: t[1])
    try:
        with open(cache, "wb") as f:
            pickle.dump(idx_ar, f)
    except OSError:
        pass
    return idx_ar


def map_idx2ar(idx_ar_sorted, batch_size: int) -> Dict[int, float]:
    """Every image gets the mean aspect ratio of the batch it falls into."""
    out = {}
    for s in range(0, len(idx_ar_sorted), max(1, batch_size)):
        chunk = idx_ar_sorted[s:s + batch_size]
        mean = round(float(np.mean([ar for _, ar in chunk])), 5)
        for i, _ in chunk:
            out[i] = mean
    return out


def rect_crop_size(target_ar: float, target_size: int) -> Tuple[int, int]:
    """(height, width) of the centre crop for a batch aspect ratio; the long side is a multiple of 8."""
    if target_ar < 1:
        return (int(target_size / target_ar) // 8 * 8, target_size)
    return (target_size, int(target_size * target_ar) // 8 * 8)


class CropArTfm:
    def __init__(self, idx2ar: Dict[int, float], target_size: int):
        self.idx2ar, self.target_size = idx2ar, target_size

    def __call__(self, img, idx):
        import torchvision.transforms.functional as TF
        return TF.center_crop(img, rect_crop_size(self.idx2ar[idx], self.target_size))


def _image_folder():
    from torchvision import datasets
    return datasets.ImageFolder


class ValDataset(Dataset):
    """ImageFolder whose transform list may contain index-aware transforms (``CropArTfm``)."""

    def __init__(self, root, transform=None):
        self.inner = _image_folder()(root)
        self.transform = transform or []

    def __len__(self):
        return len(self.inner)

    def __getitem__(self, index):
        path, target = self.inner.imgs[index]
        sample = self.inner.loader(path)
        for tfm in self.transform:
            sample = tfm(sample, index) if isinstance(tfm, CropArTfm) else tfm(sample)
        return sample, target


class SyntheticImageNet(Dataset):
    """Deterministic random uint8 CHW images with ImageNet-like labels."""

    def __init__(self, length: int, size: int = 224, num_classes: int = 1000, seed: int = 0, pool: int = 64):
        g = torch.Generator().manual_seed(seed)
        self.length, self.size = length, size
        self.pool = torch.randint(0, 256, (pool, 3, size, size), dtype=torch.uint8, generator=g)
        self.labels = torch.randint(0, num_classes, (length,), generator=g)

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        return self.pool[i % self.pool.size(0)].numpy(), int(self.labels[i])


def create_validation_set(valdir, batch_size, target_size, rect_val, distributed):
    from torchvision import transforms
    if rect_val:
        idx_ar_sorted = sort_ar(valdir)
        idx_sorted = [i for i, _ in idx_ar_sorted]
        idx2ar = map_idx2ar(idx_ar_sorted, batch_size)
        tfms = [transforms.Resize(int(target_size * 1.14)), CropArTfm(idx2ar, target_size)]
        return ValDataset(valdir, tfms), DistValSampler(idx_sorted, batch_size, distributed)
    tfms = transforms.Compose([transforms.Resize(int(target_size * 1.14)), transforms.CenterCrop(target_size)])
    ds = _image_folder()(valdir, tfms)
    return ds, DistValSampler(list(range(len(ds))), batch_size, distributed)


def get_loaders(traindir, valdir, sz, bs, dtype=torch.float32, val_bs=None, workers=8, rect_val=False,
                min_scale=0.08, distributed=False, device=None, synthetic: Optional[int] = None,
                channels_last=True, **_unused):
    """Returns ``(train_loader, val_loader, train_sampler, val_sampler)``.

    ``synthetic=N`` replaces both folders by ``SyntheticImageNet`` with N training images.
    """
    val_bs = val_bs or bs
    if synthetic:
        train_ds = SyntheticImageNet(synthetic, sz)
        val_ds = SyntheticImageNet(max(synthetic // 10, val_bs), sz, seed=1)
        val_sampler = DistValSampler(list(range(len(val_ds))), val_bs, distributed)
    else:
        from torchvision import transforms
        tf = transforms.Compose([transforms.RandomResizedCrop(sz, scale=(min_scale, 1.0)),
                                 transforms.RandomHorizontalFlip()])
        train_ds = _image_folder()(traindir, tf)
        val_ds, val_sampler = create_validation_set(valdir, val_bs, sz, rect_val, distributed)
    train_sampler = DistributedSampler(train_ds, num_replicas=env_world_size(), rank=env_rank()) if distributed else None
    kw = dict(num_workers=workers, pin_memory=torch.cuda.is_available(), collate_fn=fast_collate)
    if workers > 0 and is_set("PYTORCH_USE_SPAWN"):
        kw["multiprocessing_context"] = "spawn"
    train_loader = DataLoader(train_ds, batch_size=bs, shuffle=train_sampler is None, sampler=train_sampler,
                              drop_last=False, **kw)
    val_loader = DataLoader(val_ds, batch_sampler=val_sampler, **kw)
    return (BatchTransformLoader(train_loader, dtype, device, channels_last),
            BatchTransformLoader(val_loader, dtype, device, channels_last), train_sampler, val_sampler)


class DevicePrefetcher:
    """Iterates ``(images, target)`` host batches (ideally pinned) and hands out device tensors, copying batch
    i+1 on a side stream while batch i is being consumed -- the host->device transfer (33.7 MB per 224-image uint8
    batch, ~1.3 ms over PCIe) then overlaps the previous step's kernels instead of sitting in front of them.
    The reference copies with ``non_blocking=True`` on the compute stream (dataloader.py:88-99), which serialises."""

    def __init__(self, loader, device=None):
        self.loader = loader
        self.device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
        self.stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None

    def __len__(self):
        return len(self.loader)

    def _stage(self, batch):
        x, y = batch
        if self.stream is None:
            return x.to(self.device), y.to(self.device), None
        with torch.cuda.stream(self.stream):
            xd = x.to(self.device, non_blocking=True)
            yd = y.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return xd, yd, ev

    def __iter__(self):
        it = iter(self.loader)
        try:
            staged = self._stage(next(it))
        except StopIteration:
            return
        while staged is not None:
            xd, yd, ev = staged
            try:
                nxt = self._stage(next(it))      # launch the next copy before the consumer enqueues its compute
            except StopIteration:
                nxt = None
            if ev is not None:
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(ev)
                xd.record_stream(cur)
                yd.record_stream(cur)
            yield xd, yd
            staged = nxt
