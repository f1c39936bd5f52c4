import pytest
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200 import models


@pytest.mark.parametrize("name,tensors,params", [("resnet9", 25, 6573120), ("Resent9", 25, 6573120),
                                                  ("Alexnet", 16, 2255296), ("Alexnet1", 16, 23272266)])
def test_cifar_zoo_counts(name, tensors, params):
    m = models.build_cifar_model(name)
    ps = list(m.parameters())
    assert len(ps) == tensors and sum(p.numel() for p in ps) == params


def test_resnet50_counts_and_bn0():
    m = models.resnet50(bn0=True)
    ps = list(m.parameters())
    assert len(ps) == 161 and sum(p.numel() for p in ps) == 25557032
    assert float(m.layer1[0].bn3.weight.abs().sum()) == 0.0
    assert float(m.layer1[0].bn1.weight.sum()) == 64.0


def test_resnet18_forward():
    m = models.resnet18(num_classes=10).eval()
    y = m(torch.randn(2, 3, 64, 64))
    assert y.shape == (2, 10)


def test_vgg16_structure():
    feats = models.make_layers(models.cfgs["D"])
    convs = [l for l in feats if isinstance(l, torch.nn.Conv2d)]
    assert len(convs) == 13 and all(c.bias is not None for c in convs)


def test_graph_network_forward_and_outputs():
    m = models.build_cifar_model("resnet9")
    batch = {"input": torch.randn(4, 3, 32, 32), "target": torch.randint(0, 10, (4,))}
    out = m(batch)
    assert out["loss"].shape == (4,) and out["correct"].shape == (4,)
    assert "layer1_residual_add" in out and out["classifier"].shape == (4, 10)
    out["loss"].sum().backward()
    assert all(p.grad is not None for p in m.parameters())


def test_alexnet_module_dict_io():
    m = models.AlexNet()
    out = m({"input": torch.randn(2, 3, 32, 32), "target": torch.tensor([1, 2])})
    assert set(out) == {"loss", "correct"} and out["loss"].shape == (2,)


def test_dot_export_and_remove_by_type():
    net = models.union(models.resnet9(), models.losses())
    g = models.build_graph(net)
    dot = models.to_dot(g)
    assert dot.startswith("digraph") and '"prep_conv" -> "prep_bn"' in dot
    g2 = models.remove_by_type(g, models.Identity)
    assert "layer1_residual_in" not in g2
    assert g2["layer1_residual_res1_conv"][1] == ["layer1_pool"]


def test_unknown_network_raises():
    with pytest.raises(ValueError):
        models.build_cifar_model("lenet")


def test_space_to_depth_stem_is_the_same_convolution():
    import torch.nn.functional as F
    from layerwise_b200.ops.input import normalize_u8, s2d_stem_weight, space_to_depth_reference
    torch.manual_seed(0)
    x = torch.randint(0, 256, (2, 3, 32, 48), dtype=torch.uint8)
    w7 = torch.randn(8, 3, 7, 7)
    ref = F.conv2d(normalize_u8(x, torch.float32), w7, stride=2, padding=3)
    xs = normalize_u8(x, torch.float32, s2d=True)
    assert xs.shape == (2, 16, 19, 27)
    got = F.conv2d(xs, s2d_stem_weight(w7))
    assert got.shape == ref.shape and torch.allclose(got, ref, atol=1e-3)
    m = models.resnet18(num_classes=4).eval()
    xn = normalize_u8(x, torch.float32)
    assert torch.allclose(m(xn), m(space_to_depth_reference(xn)), atol=1e-3)
