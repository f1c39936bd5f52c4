"""tcgen05 stem weight-gradient kernel (csrc/stem.cu) against a plain fp32 PyTorch convolution backward."""
import pytest
import torch
import torch.nn.functional as F

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


def _reference(x_u8, dy):
    from layerwise_b200.ops.input import IMAGENET_MEAN, IMAGENET_STD
    mean = torch.tensor(IMAGENET_MEAN, device="cuda").view(1, 3, 1, 1)
    std = torch.tensor(IMAGENET_STD, device="cuda").view(1, 3, 1, 1)
    xn = ((x_u8.float() - mean) / std).to(torch.bfloat16).float()          # what the kernel sees, in fp32
    w = torch.zeros(64, 3, 7, 7, device="cuda", requires_grad=True)
    y = F.conv2d(xn, w, None, stride=2, padding=3)
    y.backward(dy.float())
    return w.grad


@pytest.mark.parametrize("n,size", [(2, 32), (3, 64), (1, 96), (5, 224), (2, 288), (7, 128)])
def test_stem_wgrad_matches_fp32_reference(n, size):
    from layerwise_b200.ops.input import normalize_u8, stem_wgrad, stem_wgrad_native_ok
    torch.manual_seed(n * 1000 + size)
    assert stem_wgrad_native_ok(size // 2, size // 2, torch.bfloat16)
    x = torch.randint(0, 256, (n, 3, size, size), dtype=torch.uint8, device="cuda")
    xs = normalize_u8(x, torch.bfloat16, s2d=True)
    dy = torch.randn(n, 64, size // 2, size // 2, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    got = stem_wgrad(dy, xs, torch.float32)
    torch.cuda.synchronize()
    ref = _reference(x, dy)
    assert got.shape == ref.shape
    scale = float(ref.abs().max())
    err = float((got - ref).abs().max())
    assert err <= 2e-3 * scale + 1e-3, (err, scale)
    # bf16 output + scale
    got16 = stem_wgrad(dy, xs, torch.bfloat16, scale=0.5)
    assert float((got16.float() - 0.5 * ref).abs().max()) <= 1e-2 * scale + 1e-3


def test_stem_module_path_uses_native_wgrad():
    """End to end through the model's stem: gradient of conv1.weight equals the stock 7x7 convolution's."""
    from layerwise_b200.ops import input as inp
    torch.manual_seed(0)
    x = torch.randint(0, 256, (4, 3, 64, 64), dtype=torch.uint8, device="cuda")
    w7 = (torch.randn(64, 3, 7, 7, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w7.requires_grad_(True)
    xs = inp.normalize_u8(x, torch.bfloat16, s2d=True, keep_plain=True)
    assert getattr(xs, "plain3", None) is None            # no 3-channel copy needed any more
    y = inp.stem_s2d(xs, w7)
    dy = torch.randn_like(y)
    y.backward(dy)
    ref = _reference(x, dy)
    scale = float(ref.abs().max())
    assert float((w7.grad.float() - ref).abs().max()) <= 1e-2 * scale + 1e-3
