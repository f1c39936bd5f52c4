import numpy as np
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200 import data


def test_cifar_preprocessing_and_transforms():
    ds = data.synthetic_cifar10(64, 16)
    x = data.transpose(data.normalise(data.pad(ds["train"]["data"], 4)))
    assert x.shape == (64, 3, 40, 40) and x.dtype == np.float32
    tfm = data.Transform(list(zip(x, ds["train"]["labels"])), [data.Crop(32, 32), data.FlipLR(), data.Cutout(8, 8)])
    tfm.set_random_choices(np.random.RandomState(0))
    img, lab = tfm[3]
    assert img.shape == (3, 32, 32) and (img == 0).sum() >= 3 * 64
    b = data.Batches(tfm, 16, shuffle=True, set_random_choices=True, drop_last=True, device="cpu")
    batch = next(iter(b))
    assert batch["input"].shape == (16, 3, 32, 32) and batch["target"].dtype == torch.int64 and len(b) == 4


def test_device_batches_cpu_path_matches_shapes_and_cutout():
    ds = data.synthetic_cifar10(40, 8)
    train, test = data.cifar_batches(ds, 16, device="cpu", on_device=True)
    seen = 0
    for batch in train:
        assert batch["input"].shape == (16, 3, 32, 32)
        holes = (batch["input"] == 0).all(dim=1).flatten(1).sum(1)
        assert (holes >= 64).all()            # every sample carries its 8x8 cut-out
        seen += 1
    assert seen == len(train) == 2
    assert sum(b["input"].size(0) for b in test) == 8


def test_dist_val_sampler_equal_batches():
    s0 = data.DistValSampler(list(range(10)), 4, distributed=True, world_size=2, rank=0)
    s1 = data.DistValSampler(list(range(10)), 4, distributed=True, world_size=2, rank=1)
    b0, b1 = list(s0), list(s1)
    assert len(b0) == len(b1) == 2
    assert b0 == [[0, 1, 2, 3], [4, 5, 6, 7]] and b1 == [[8, 9], []]


def test_rect_val_helpers_and_collate():
    sizes = [(500, 375), (375, 500), (400, 400), (640, 480)]
    idx_ar = data.sort_ar("/nonexistent/validation", sizes=sizes)
    assert [i for i, _ in idx_ar] == [1, 2, 0, 3]
    m = data.map_idx2ar(idx_ar, 2)
    assert m[1] == m[2] and m[0] == m[3]
    assert data.rect_crop_size(0.75, 128) == (168, 128) and data.rect_crop_size(1.33333, 128) == (128, 168)
    batch = [(np.zeros((8, 6, 3), np.uint8), 1), (np.ones((8, 6), np.uint8), 2)]
    x, y = data.fast_collate(batch)
    assert x.shape == (2, 3, 8, 6) and x.dtype == torch.uint8 and y.tolist() == [1, 2] and int(x[1].sum()) == 3 * 48


def test_synthetic_imagenet_loaders():
    trn, val, trn_smp, val_smp = data.get_loaders("x", "y", sz=32, bs=4, workers=0, synthetic=16, device="cpu")
    x, y = next(iter(trn))
    assert x.shape == (4, 3, 32, 32) and x.dtype == torch.float32 and abs(float(x.mean())) < 1.0
    assert len(trn) == 4 and len(val) >= 1


def test_device_prefetcher_cpu_passthrough():
    batches = [(torch.full((2, 3), i, dtype=torch.uint8), torch.tensor([i, i])) for i in range(4)]
    out = list(data.DevicePrefetcher(batches, device="cpu"))
    assert len(out) == 4 and all(int(x[0, 0]) == i and y.tolist() == [i, i] for i, (x, y) in enumerate(out))
