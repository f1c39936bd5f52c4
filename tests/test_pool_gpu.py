import pytest
import torch
import torch.nn.functional as F

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("k,s,p,shape", [(3, 2, 1, (4, 64, 37, 41)), (2, 2, 0, (8, 128, 16, 16)),
                                        (4, 4, 0, (8, 512, 4, 4)), (3, 2, 1, (2, 64, 112, 112)), (3, 1, 1, (2, 8, 9, 7)),
                                        (3, 2, 1, (3, 16, 6, 10)), (3, 2, 1, (2, 8, 2, 2))])
def test_maxpool_matches_torch(dtype, k, s, p, shape):
    from layerwise_b200.ops import nn as lnn
    torch.manual_seed(0)
    x = torch.randn(shape, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last)
    x1 = x.clone().requires_grad_(True)
    x2 = x.clone().requires_grad_(True)
    pool = lnn.MaxPool2d(k, s, p)
    y1 = pool(x1)
    y2 = F.max_pool2d(x2, k, s, p)
    assert y1.shape == y2.shape and torch.equal(y1, y2)
    dy = torch.randn_like(y2)
    y1.backward(dy)
    y2.backward(dy)
    assert torch.allclose(x1.grad.float(), x2.grad.float(), atol=1e-2 if dtype == torch.bfloat16 else 1e-6)
