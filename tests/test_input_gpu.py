import pytest
import torch

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_fused_normalize_s2d_matches_reference(dtype):
    from layerwise_b200.ops.input import IMAGENET_MEAN, IMAGENET_STD, normalize_u8, space_to_depth_reference
    x = torch.randint(0, 256, (3, 3, 64, 96), dtype=torch.uint8, device="cuda")
    got = normalize_u8(x, dtype, s2d=True)
    mean = torch.tensor(IMAGENET_MEAN, device="cuda").view(1, 3, 1, 1)
    std = torch.tensor(IMAGENET_STD, device="cuda").view(1, 3, 1, 1)
    ref = space_to_depth_reference((x.float() - mean) / std)
    assert got.shape == ref.shape and got.is_contiguous(memory_format=torch.channels_last)
    assert torch.allclose(got.float(), ref, atol=2e-2 if dtype == torch.bfloat16 else 1e-5)


def test_resnet_s2d_stem_forward_backward_matches_plain_path():
    from layerwise_b200 import models
    from layerwise_b200.ops.input import normalize_u8
    torch.manual_seed(0)
    m = models.resnet18(num_classes=10).cuda().to(memory_format=torch.channels_last)
    x = torch.randint(0, 256, (4, 3, 64, 64), dtype=torch.uint8, device="cuda")
    xs = normalize_u8(x, torch.float32, s2d=True, keep_plain=True)
    y1 = m.stem(xs)
    dy = torch.randn_like(y1)
    y1.backward(dy)
    g1 = m.conv1.weight.grad.clone()
    m.zero_grad()
    y2 = m.stem(normalize_u8(x, torch.float32).contiguous(memory_format=torch.channels_last))
    y2.backward(dy)
    g2 = m.conv1.weight.grad
    assert y1.shape == y2.shape == (4, 64, 32, 32)
    assert (y1 - y2).abs().max() <= 1e-2 * y2.abs().max()        # TF32 convolutions on both paths
    assert (g1 - g2).abs().max() <= 1e-2 * g2.abs().max()
    # and the whole network accepts the space-to-depth input
    out = m(xs)
    assert out.shape == (4, 10) and torch.isfinite(out).all()
