"""tcgen05 tap-table convolution kernels (csrc/convx.cu, csrc/conv.cu) vs an fp32 PyTorch reference."""
import os

import pytest
import torch
import torch.nn.functional as F

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


def _mk(n, ci, co, h, w, k):
    torch.manual_seed(0)
    x = torch.randn(n, ci, h, w, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(co, ci, k, k, device="cuda") / (ci * k * k) ** 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    return x, wt


@pytest.mark.parametrize("n,ci,co,h,w", [(2, 64, 64, 8, 8), (3, 128, 256, 14, 14), (5, 256, 64, 56, 56), (1, 64, 192, 7, 9)])
def test_conv1x1_forward_stats_and_backward(n, ci, co, h, w):
    from layerwise_b200.ops.conv import conv2d
    x, wt = _mk(n, ci, co, h, w, 1)
    x1, w1 = x.clone().requires_grad_(True), wt.clone().requires_grad_(True)
    y, stats = conv2d(x1, w1, None, want_stats=True)
    ref_in, ref_w = x.float().requires_grad_(True), wt.float().requires_grad_(True)
    yr = F.conv2d(ref_in, ref_w)
    assert y.shape == yr.shape and y.is_contiguous(memory_format=torch.channels_last)
    assert (y.float() - yr).abs().max() <= 2e-2 * yr.abs().max()
    # epilogue statistics = column sums of the rounded output
    s = stats.sums()
    yf = y.float()
    assert torch.allclose(s[0], yf.sum((0, 2, 3)), rtol=1e-3, atol=1e-2)
    assert torch.allclose(s[1], (yf * yf).sum((0, 2, 3)), rtol=1e-3, atol=1e-2)
    dy = torch.randn_like(y)
    y.backward(dy)
    yr.backward(dy.float())
    assert (x1.grad.float() - ref_in.grad).abs().max() <= 2e-2 * ref_in.grad.abs().max()
    assert (w1.grad.float() - ref_w.grad).abs().max() <= 2e-2 * ref_w.grad.abs().max()


@pytest.mark.parametrize("n,ci,co,h,w", [(2, 64, 64, 8, 8), (3, 64, 128, 32, 32), (4, 128, 128, 28, 28), (3, 256, 256, 14, 14),
                                         (5, 512, 512, 7, 7), (2, 64, 64, 56, 56), (16, 64, 64, 4, 4)])
def test_conv3x3_forward_stats_and_dgrad(n, ci, co, h, w):
    from layerwise_b200.ops.conv import conv2d
    x, wt = _mk(n, ci, co, h, w, 3)
    x1, w1 = x.clone().requires_grad_(True), wt.clone().requires_grad_(True)
    y, stats = conv2d(x1, w1, None, want_stats=True)
    ref_in, ref_w = x.float().requires_grad_(True), wt.float().requires_grad_(True)
    yr = F.conv2d(ref_in, ref_w, padding=1)
    assert (y.float() - yr).abs().max() <= 2e-2 * yr.abs().max()
    s = stats.sums()
    yf = y.float()
    assert torch.allclose(s[0], yf.sum((0, 2, 3)), rtol=1e-3, atol=2e-2)
    assert torch.allclose(s[1], (yf * yf).sum((0, 2, 3)), rtol=1e-3, atol=2e-2)
    dy = torch.randn_like(y)
    y.backward(dy)
    yr.backward(dy.float())
    assert (x1.grad.float() - ref_in.grad).abs().max() <= 2e-2 * ref_in.grad.abs().max()
    assert (w1.grad.float() - ref_w.grad).abs().max() <= 3e-2 * ref_w.grad.abs().max()


def test_fork_conv1x1_fused_residual_gradient():
    from layerwise_b200.ops.conv import fork_conv1x1
    torch.manual_seed(0)
    conv = torch.nn.Conv2d(256, 64, 1, bias=False).cuda().to(torch.bfloat16).to(memory_format=torch.channels_last)
    x = torch.randn(4, 256, 14, 14, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    x1 = x.clone().requires_grad_(True)
    y, ident = fork_conv1x1(conv, x1)
    gy, gi = torch.randn_like(y), torch.randn_like(ident)
    torch.autograd.backward([y, ident], [gy, gi])
    gw1 = conv.weight.grad.clone()
    conv.weight.grad = None
    x2 = x.float().requires_grad_(True)
    w2 = conv.weight.detach().float().requires_grad_(True)
    y2 = F.conv2d(x2, w2)
    torch.autograd.backward([y2, x2 * 1.0], [gy.float(), gi.float()])
    assert (y.float() - y2).abs().max() <= 2e-2 * y2.abs().max()
    assert (x1.grad.float() - x2.grad).abs().max() <= 2e-2 * x2.grad.abs().max()
    assert (gw1.float() - w2.grad).abs().max() <= 3e-2 * w2.grad.abs().max()


def test_fork_premasks_gradient_for_the_producing_batchnorm():
    """Two bottlenecks in a row: the second block's fork multiplies its input gradient with the first block's ReLU mask
    (kernel epilogue) and the first block's bn3 backward hands dy on without masking or writing a residual gradient.
    Gradients must match the stock-op pipeline."""
    import copy
    from layerwise_b200.models import resnet as zoo
    from layerwise_b200.ops import nn as lnn
    torch.manual_seed(0)
    blocks = torch.nn.Sequential(zoo.Bottleneck(256, 64), zoo.Bottleneck(256, 64), zoo.Bottleneck(256, 64)).cuda()
    blocks = blocks.to(memory_format=torch.channels_last).to(torch.bfloat16)
    for m in blocks.modules():
        if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
            m.float()
            torch.nn.init.uniform_(m.weight, 0.5, 1.5)
            torch.nn.init.normal_(m.bias, 0.0, 0.5)
    ref = copy.deepcopy(blocks)
    x0 = torch.randn(8, 256, 12, 12, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    g = torch.randn(8, 256, 12, 12, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)

    def run(net, native):
        lnn.set_backend("auto" if native else "torch")
        try:
            x = x0.clone().requires_grad_(True)
            y1 = net[0](x)
            if native:
                assert y1.relu_mask_taken[0] is False
            y2 = net[1](y1)
            if native and not os.environ.get("LWC_NO_PREMASK"):
                assert y1.relu_mask_taken[0] is True and y2.relu_mask_taken[0] is False
            y3 = net[2](y2)
            (y3 * g).sum().backward()
            return x.grad.float(), [p.grad.float() for p in net.parameters()]
        finally:
            lnn.set_backend("auto")

    # (1) pre-masking is an exact re-association of the same arithmetic: bit-identical to the path that masks inside
    #     the batch-norm backward and writes a separate residual gradient
    gx_a, gp_a = run(blocks, True)
    os.environ["LWC_NO_PREMASK"] = "1"
    try:
        for p_ in blocks.parameters():
            p_.grad = None
        gx_c, gp_c = run(blocks, True)
    finally:
        del os.environ["LWC_NO_PREMASK"]
    #     (with the fork kernel that also sums bn3's backward reductions the sums are taken in another order: then the
    #     two agree to rounding, not to the bit)
    from layerwise_b200.ops import conv as C_

    def same(a, c):
        if not (C_._FORK_NEW["on"] and C_._FORK_LINK["on"] and C_._FUSE_BNB["on"]):
            return torch.equal(a, c)
        return float((a - c).norm() / (c.norm() + 1e-12)) < 5e-3
    assert same(gx_a, gx_c)
    for a, c in zip(gp_a, gp_c):
        assert same(a, c)
    # (2) against the stock-op pipeline: a ReLU whose bf16 pre-activation rounds to the other side of zero flips a whole
    #     gradient path, so compare in relative L2 and loosely (nine batch-norms deep, bf16)
    gx_b, gp_b = run(ref, False)

    def rel(a, b):
        return float((a - b).norm() / (b.norm() + 1e-12))
    assert rel(gx_a, gx_b) < 0.25, rel(gx_a, gx_b)
    for a, b in zip(gp_a, gp_b):
        assert rel(a, b) < 0.25, (a.shape, rel(a, b))


def _rel(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12))


@pytest.mark.parametrize("n,ci,co,h,w,k,stride,pad", [
    (4, 128, 128, 56, 56, 3, 2, 1),      # ResNet-50 stage transitions: 3x3 / 2
    (3, 256, 256, 28, 28, 3, 2, 1),
    (5, 512, 512, 14, 14, 3, 2, 1),
    (4, 256, 512, 56, 56, 1, 2, 0),      # 1x1 / 2 downsample
    (3, 1024, 2048, 14, 14, 1, 2, 0),
    (2, 64, 128, 15, 17, 3, 2, 1),       # odd extent
    (2, 192, 384, 8, 8, 3, 1, 1),        # AlexNet channel counts
    (2, 64, 192, 16, 16, 3, 1, 1),
    (6, 64, 64, 2, 2, 3, 1, 1),          # tiny late VGG maps
])
def test_convx_strided_and_odd_shapes(n, ci, co, h, w, k, stride, pad):
    from layerwise_b200.ops.conv import conv2d
    x, wt = _mk(n, ci, co, h, w, k)
    x1, w1 = x.clone().requires_grad_(True), wt.clone().requires_grad_(True)
    y, stats = conv2d(x1, w1, None, want_stats=True, stride=stride, padding=pad)
    ref_in, ref_w = x.float().requires_grad_(True), wt.float().requires_grad_(True)
    yr = F.conv2d(ref_in, ref_w, stride=stride, padding=pad)
    assert y.shape == yr.shape and y.is_contiguous(memory_format=torch.channels_last)
    assert _rel(y, yr) <= 2e-2
    s = stats.sums()
    yf = y.float()
    assert torch.allclose(s[0], yf.sum((0, 2, 3)), rtol=1e-3, atol=3e-2)
    assert torch.allclose(s[1], (yf * yf).sum((0, 2, 3)), rtol=1e-3, atol=3e-2)
    dy = torch.randn_like(y)
    y.backward(dy)
    yr.backward(dy.float())
    assert _rel(x1.grad, ref_in.grad) <= 2e-2
    assert _rel(w1.grad, ref_w.grad) <= 3e-2


@pytest.mark.parametrize("bn", [64, 128, 256])
@pytest.mark.parametrize("k", [1, 3])
def test_convx_every_tile_width(bn, k):
    from layerwise_b200.ops.conv import conv_forward, conv_dgrad, conv_wgrad
    x, wt = _mk(3, 256, 256, 14, 14, k)
    y, st = conv_forward(x, wt, want_stats=True, bn_hint=bn)
    yr = F.conv2d(x.float(), wt.float(), padding=k // 2)
    assert _rel(y, yr) <= 2e-2
    assert torch.allclose(st.sums()[0], y.float().sum((0, 2, 3)), rtol=1e-3, atol=3e-2)
    dy = torch.randn_like(y)
    xi, wi = x.float().requires_grad_(True), wt.float().requires_grad_(True)
    F.conv2d(xi, wi, padding=k // 2).backward(dy.float())
    assert _rel(conv_dgrad(dy, wt, (14, 14), 1, k // 2, bn_hint=bn), xi.grad) <= 2e-2
    assert _rel(conv_wgrad(x, dy, wt.shape, 1, k // 2, bn_hint=bn), wi.grad) <= 3e-2


@pytest.mark.parametrize("n,co,h,w,k,stride,pad", [(8, 64, 32, 32, 3, 1, 1), (4, 64, 32, 32, 3, 2, 1), (2, 128, 9, 11, 3, 1, 1)])
def test_convx_three_channel_input_bias_relu(n, co, h, w, k, stride, pad):
    """First layers: 3 input channels (stored as 8, padded to the 64-wide K chunk by the TMA descriptor), bias + ReLU
    fused in the epilogue (CIFAR10/alexnet.py:15-16, CIFAR10/vgg16.py:69-82, CIFAR10/dawn.py:23-28)."""
    from layerwise_b200.ops.conv import conv2d
    torch.manual_seed(1)
    x = torch.randn(n, 3, h, w, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(co, 3, k, k, device="cuda") / 5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    b = torch.randn(co, device="cuda").to(torch.bfloat16)
    w1, b1 = wt.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y, _ = conv2d(x, w1, b1, stride=stride, padding=pad, relu=True)
    wr, br = wt.float().requires_grad_(True), b.float().requires_grad_(True)
    yr = F.relu(F.conv2d(x.float(), wr, br, stride=stride, padding=pad))
    assert _rel(y, yr) <= 2e-2
    dy = torch.randn_like(y)
    y.backward(dy)
    (F.conv2d(x.float(), wr, br, stride=stride, padding=pad) * (dy.float() * (y.float() > 0))).sum().backward()
    assert _rel(w1.grad, wr.grad) <= 3e-2
    assert _rel(b1.grad, br.grad) <= 3e-2


@pytest.mark.parametrize("n,hw", [(4, 64), (2, 224), (3, 96)])
def test_convx_stem_forward_matches_7x7(n, hw):
    """The 7x7/2 stem through the overlapping-window tensor map (IMAGENET/training/resnet.py:100-101)."""
    from layerwise_b200.ops.input import normalize_u8, stem_forward
    torch.manual_seed(0)
    xu = torch.randint(0, 256, (n, 3, hw, hw), dtype=torch.uint8, device="cuda")
    w7 = (torch.randn(64, 3, 7, 7, device="cuda") / 12).to(torch.bfloat16)
    xs = normalize_u8(xu, s2d=True)
    y = stem_forward(xs, w7, want_stats=False)
    xn = normalize_u8(xu, s2d=False).float()
    yr = F.conv2d(xn.to(torch.bfloat16).float(), w7.float(), stride=2, padding=3)
    assert y.shape == yr.shape
    assert _rel(y, yr) <= 2e-2


def test_conv_emitted_statistics_feed_the_batch_norm():
    """conv -> bn with the statistics summed in the convolution epilogue equals conv -> bn with the statistics pass."""
    from layerwise_b200.ops import nn as lnn
    torch.manual_seed(0)
    conv = lnn.Conv2d(128, 256, 3, padding=1, bias=False).cuda().to(torch.bfloat16).to(memory_format=torch.channels_last)
    bn = lnn.BatchNorm2d(256).cuda()
    x = torch.randn(6, 128, 14, 14, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    outs = []
    for fused in (True, False):
        conv.emit_bn_stats = fused
        bn.reset_running_stats()
        xi = x.clone().requires_grad_(True)
        h = conv(xi)
        assert (getattr(h, "bn_stats", None) is not None) == fused
        y = bn(h, relu=True)
        y.float().square().sum().backward()
        outs.append((y.float(), xi.grad.float(), bn.running_mean.clone(), bn.running_var.clone(), int(bn.num_batches_tracked)))
        conv.weight.grad = None
        bn.weight.grad = bn.bias.grad = None
    (ya, ga, ma, va, na), (yb, gb, mb, vb, nb) = outs
    assert na == nb == 1
    assert torch.allclose(ma, mb, rtol=1e-4, atol=1e-5) and torch.allclose(va, vb, rtol=1e-3, atol=1e-5)
    assert _rel(ya, yb) <= 1e-2 and _rel(ga, gb) <= 2e-2


@pytest.mark.parametrize("n,ci,co,hw,k", [(4, 64, 64, 56, 3), (6, 128, 512, 28, 1), (3, 256, 256, 14, 3), (5, 512, 128, 7, 1), (2, 64, 256, 9, 1)])
@pytest.mark.parametrize("relu", [True, False])
def test_dgrad_epilogue_sums_batch_norm_backward_reductions(n, ci, co, hw, k, relu):
    """lwc_convx_dgrad_bnstats: dX and, from the same launch, S1 = sum d and S2 = sum d * x with d = dX * [bn(x) > 0]."""
    from layerwise_b200.ops.conv import conv_dgrad, conv_dgrad_bnstats
    torch.manual_seed(0)
    _, wt = _mk(n, ci, co, hw, hw, k)
    dy = torch.randn(n, co, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xb = torch.randn(n, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gamma, beta = torch.rand(ci, device="cuda") + 0.5, torch.randn(ci, device="cuda") * 0.3
    mean, invstd = torch.randn(ci, device="cuda") * 0.2, torch.rand(ci, device="cuda") + 0.5
    link = dict(x=xb, gamma=gamma, beta=beta, mean=mean, invstd=invstd, relu=relu)
    dx = conv_dgrad_bnstats(dy, wt, (hw, hw), k // 2, link)
    ref = conv_dgrad(dy, wt, (hw, hw), 1, k // 2)
    assert torch.equal(dx, ref)                                   # same kernel, same arithmetic
    part, nparts = link["partials"]
    s = part[:nparts].sum(0)
    a = gamma * invstd
    b = beta - mean * a
    d = dx.float()
    if relu:
        d = d * (torch.addcmul(b.view(1, -1, 1, 1), xb.float(), a.view(1, -1, 1, 1)) > 0)
    s1, s2 = d.sum((0, 2, 3)), (d * xb.float()).sum((0, 2, 3))
    scale = float(d.abs().sum((0, 2, 3)).max())
    assert (s[0] - s1).abs().max() <= 2e-3 * scale and (s[1] - s2).abs().max() <= 4e-3 * scale


def test_bottleneck_with_fused_batch_norm_backward_matches_unfused():
    """Two bottlenecks: bn1 / bn2 get their backward sums from the data-gradient epilogue of conv2 / conv3; gradients
    must agree with the path that runs the separate reduction kernel."""
    import copy
    from layerwise_b200.models import resnet as zoo
    from layerwise_b200.ops import conv as C_
    torch.manual_seed(0)
    blocks = torch.nn.Sequential(zoo.Bottleneck(256, 64), zoo.Bottleneck(256, 64)).cuda().to(memory_format=torch.channels_last).to(torch.bfloat16)
    for m in blocks.modules():
        if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
            m.float()
            torch.nn.init.uniform_(m.weight, 0.5, 1.5)
            torch.nn.init.normal_(m.bias, 0.0, 0.5)
    x0 = torch.randn(8, 256, 14, 14, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    g = torch.randn(8, 256, 14, 14, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res = []
    for fused in (True, False):
        C_.fuse_bn_backward(fused)
        try:
            net = copy.deepcopy(blocks)
            x = x0.clone().requires_grad_(True)
            (net(x) * g).sum().backward()
            res.append((x.grad.float(), [p.grad.float() for p in net.parameters()]))
        finally:
            C_.fuse_bn_backward(True)
    (gx_a, gp_a), (gx_b, gp_b) = res

    def relerr(a, b):
        return float((a - b).norm() / (b.norm() + 1e-12))
    assert relerr(gx_a, gx_b) < 2e-2
    for a, b in zip(gp_a, gp_b):
        assert relerr(a, b) < 2e-2, (a.shape, relerr(a, b))


def _ref_conv(x, wt, k, relu=False):
    return F.conv2d(x.float(), wt.float(), padding=k // 2)


@pytest.mark.parametrize("n,ci,co,h,w", [(4, 64, 64, 56, 56), (3, 64, 128, 32, 32), (2, 64, 64, 30, 17), (5, 64, 64, 7, 9), (2, 3, 64, 32, 32)])
@pytest.mark.parametrize("bn", [64, 128])
def test_halo_variant_forward_statistics_and_dgrad(n, ci, co, h, w, bn):
    """Hint 1000 + BN: resident weights, one input box per tile, taps read from shifted start addresses.  Forward (+ batch
    statistics) and data gradient against fp32, and bit-identical to the per-tap variant of the same tile width."""
    from layerwise_b200.ops.conv import conv_forward, conv_dgrad
    if co % bn:
        pytest.skip("tile wider than the layer")
    torch.manual_seed(0)
    x, wt = _mk(n, ci, co, h, w, 3)
    y, st = conv_forward(x, wt, want_stats=True, bn_hint=1000 + bn)
    ref = _ref_conv(x, wt, 3)
    assert _rel(y.float(), ref) <= 6e-3
    y2, _ = conv_forward(x, wt, want_stats=True, bn_hint=bn)
    assert float((y != y2).float().mean()) < 1e-3
    s = st.partials[:st.nparts].sum(0)
    yf = y.float()
    scale = float(yf.abs().sum((0, 2, 3)).max())
    assert (s[0] - yf.sum((0, 2, 3))).abs().max() <= 2e-3 * scale
    assert (s[1] - yf.square().sum((0, 2, 3))).abs().max() <= 2e-3 * float(yf.square().sum((0, 2, 3)).max())
    if ci % 64 == 0 and co == 64:
        # data gradient of a layer with 64 OUTPUT channels: K = 64
        dy = torch.randn(n, co, h, w, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        if ci % bn == 0:
            dx = conv_dgrad(dy, wt, (h, w), 1, 1, bn_hint=1000 + bn)
            dref = torch.ops.aten.convolution_backward(dy.float(), x.float(), wt.float(), None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1,
                                                       [True, False, False])[0]
            assert _rel(dx.float(), dref) <= 6e-3
            assert float((dx != conv_dgrad(dy, wt, (h, w), 1, 1, bn_hint=bn)).float().mean()) < 1e-3


@pytest.mark.parametrize("n,ci,hw", [(4, 64, 56), (2, 64, 13)])
def test_halo_variant_sums_batch_norm_backward_reductions(n, ci, hw):
    from layerwise_b200.ops.conv import conv_dgrad_bnstats
    torch.manual_seed(0)
    _, wt = _mk(n, ci, 64, hw, hw, 3)
    dy = torch.randn(n, 64, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xb = torch.randn(n, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gamma, beta = torch.rand(ci, device="cuda") + 0.5, torch.randn(ci, device="cuda") * 0.3
    mean, invstd = torch.randn(ci, device="cuda") * 0.2, torch.rand(ci, device="cuda") + 0.5
    outs = []
    for hint in (64, 1064):
        link = dict(x=xb, gamma=gamma, beta=beta, mean=mean, invstd=invstd, relu=True)
        dx = conv_dgrad_bnstats(dy, wt, (hw, hw), 1, link, bn_hint=hint)
        part, nparts = link["partials"]
        outs.append((dx, part[:nparts].sum(0)))
    assert float((outs[0][0] != outs[1][0]).float().mean()) < 1e-3
    scale = float(outs[0][1].abs().max())
    assert (outs[0][1] - outs[1][1]).abs().max() <= 2e-3 * scale


@pytest.mark.parametrize("n,co,ci,hw", [(4, 64, 256, 56), (3, 128, 512, 28), (2, 512, 2048, 7), (5, 64, 64, 9)])
@pytest.mark.parametrize("masked", [True, False])
@pytest.mark.parametrize("linked", [True, False])
def test_fork_kernel_adds_masks_and_sums(n, co, ci, hw, masked, linked):
    """lwc_convx_dgrad1x1_fork: dX = mask * (dY W + addend), and with a link the batch-norm backward sums of dX."""
    from layerwise_b200.ops import conv as C_
    if linked and not masked:
        pytest.skip("the link rides on the mask")
    torch.manual_seed(0)
    wt = (torch.randn(co, ci, 1, 1, device="cuda") * (1.0 / co) ** 0.5).to(torch.bfloat16)
    dy = torch.randn(n, co, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    add = torch.randn(n, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xb = torch.randn(n, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    keep = torch.rand(n, hw, hw, ci, device="cuda") > 0.4                      # channels-last order
    mask = None
    if masked:
        bits = keep.reshape(-1, 8).to(torch.uint8)
        mask = (bits << torch.arange(8, device="cuda", dtype=torch.uint8)).sum(1).to(torch.uint8).contiguous()
    link = dict(x=xb, mean=torch.randn(ci, device="cuda"), invstd=torch.rand(ci, device="cuda") + 0.5) if linked else None
    was, C_._FORK_NEW["on"] = C_._FORK_NEW["on"], True
    was_l, C_._FORK_LINK["on"] = C_._FORK_LINK["on"], True

    def request_restore():
        C_._FORK_NEW["on"], C_._FORK_LINK["on"] = was, was_l
    for hint in [b for b in (64, 128) if ci % b == 0]:
        try:
            dx = C_.conv1x1_dgrad_add(dy, wt, add, mask, link, bn_hint=hint)
        except Exception:
            request_restore()
            raise
        ref = torch.einsum("nohw,oi->nihw", dy.float(), wt.float().reshape(co, ci)) + add.float()
        if masked:
            ref = ref * keep.permute(0, 3, 1, 2)
        assert _rel(dx.float(), ref) <= 6e-3
        if masked:
            assert bool((dx.float()[~keep.permute(0, 3, 1, 2)] == 0).all())
        if linked:
            part, nparts = link.pop("partials")
            s = part[:nparts].sum(0)
            d = dx.float()
            s1, s2 = d.sum((0, 2, 3)), (d * xb.float()).sum((0, 2, 3))
            scale = float(d.abs().sum((0, 2, 3)).max())
            assert (s[0] - s1).abs().max() <= 2e-3 * scale and (s[1] - s2).abs().max() <= 4e-3 * scale
    request_restore()


def test_bottleneck_chain_new_fork_kernel_matches_old_path():
    """Three bottlenecks (the second and third fork from a fused batch-norm + residual + ReLU output): the fork kernel with
    its fused bn3 backward sums against the older kernel + separate reduction."""
    import copy
    from layerwise_b200.models import resnet as zoo
    from layerwise_b200.ops import conv as C_
    torch.manual_seed(0)
    blocks = torch.nn.Sequential(*[zoo.Bottleneck(256, 64) for _ in range(3)]).cuda().to(memory_format=torch.channels_last).to(torch.bfloat16)
    for m in blocks.modules():
        if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
            m.float()
            torch.nn.init.uniform_(m.weight, 0.5, 1.5)
            torch.nn.init.normal_(m.bias, 0.0, 0.5)
    x0 = torch.randn(8, 256, 14, 14, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    g = torch.randn(8, 256, 14, 14, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res = []
    was, was_l = C_._FORK_NEW["on"], C_._FORK_LINK["on"]
    for new in (True, False):
        C_._FORK_NEW["on"] = C_._FORK_LINK["on"] = new
        try:
            net = copy.deepcopy(blocks)
            x = x0.clone().requires_grad_(True)
            (net(x) * g).sum().backward()
            res.append((x.grad.float(), [p.grad.float() for p in net.parameters()]))
        finally:
            C_._FORK_NEW["on"], C_._FORK_LINK["on"] = was, was_l
    (gx_a, gp_a), (gx_b, gp_b) = res

    def relerr(a, b):
        return float((a - b).norm() / (b.norm() + 1e-12))
    assert relerr(gx_a, gx_b) < 2e-2
    for a, b in zip(gp_a, gp_b):
        assert relerr(a, b) < 2e-2, (a.shape, relerr(a, b))


@pytest.mark.parametrize("n,ci,co,h,w,k", [(8, 64, 256, 56, 56, 1), (4, 256, 64, 28, 28, 1), (3, 128, 128, 28, 28, 3), (2, 64, 64, 9, 11, 3)])
@pytest.mark.parametrize("bn", [64, 128])
def test_two_ctas_per_sm_variant(n, ci, co, h, w, k, bn):
    """Hint 2000 + BN: the variant compiled for two resident CTAs per SM (half the shared-memory plan each)."""
    from layerwise_b200.ops.conv import conv_forward, conv_dgrad
    if co % bn:
        pytest.skip("tile wider than the layer")
    x, wt = _mk(n, ci, co, h, w, k)
    ref, _ = conv_forward(x, wt, bn_hint=bn)
    y, _ = conv_forward(x, wt, bn_hint=2000 + bn)
    assert torch.equal(y, ref)
    if True:
        y2, st = conv_forward(x, wt, want_stats=True, bn_hint=2000 + bn)
        _, st_ref = conv_forward(x, wt, want_stats=True, bn_hint=bn)
        assert torch.equal(y2, ref)
        a, b = st.partials[:st.nparts].sum(0), st_ref.partials[:st_ref.nparts].sum(0)
        assert (a - b).abs().max() <= 1e-3 * float(b.abs().max())
    if ci % bn == 0:
        dy = torch.randn(n, co, h, w, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        assert torch.equal(conv_dgrad(dy, wt, (h, w), 1, k // 2, bn_hint=2000 + bn), conv_dgrad(dy, wt, (h, w), 1, k // 2, bn_hint=bn))


@pytest.mark.parametrize("n,ci,co,hw,k", [(4, 64, 256, 56, 1), (3, 128, 128, 28, 3), (2, 256, 64, 9, 1)])
def test_two_ctas_per_sm_with_auxiliary_ring(n, ci, co, hw, k):
    """Hint 2064 on the launches that prefetch a second tile per sub-tile (batch-norm backward sums; residual fork):
    two-slot rings, prefetch distance one."""
    from layerwise_b200.ops import conv as C_
    torch.manual_seed(0)
    _, wt = _mk(n, ci, co, hw, hw, k)
    dy = torch.randn(n, co, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xb = torch.randn(n, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gamma, beta = torch.rand(ci, device="cuda") + 0.5, torch.randn(ci, device="cuda") * 0.3
    mean, invstd = torch.randn(ci, device="cuda") * 0.2, torch.rand(ci, device="cuda") + 0.5
    outs = []
    for hint in (64, 2064):
        link = dict(x=xb, gamma=gamma, beta=beta, mean=mean, invstd=invstd, relu=True)
        dx = C_.conv_dgrad_bnstats(dy, wt, (hw, hw), k // 2, link, bn_hint=hint)
        part, nparts = link["partials"]
        outs.append((dx, part[:nparts].sum(0)))
    assert torch.equal(outs[0][0], outs[1][0])
    assert (outs[0][1] - outs[1][1]).abs().max() <= 2e-3 * float(outs[0][1].abs().max())
    if k == 1:
        add = torch.randn(n, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        mask = torch.randint(0, 256, (n * hw * hw * ci // 8,), device="cuda", dtype=torch.uint8)
        was, C_._FORK_NEW["on"] = C_._FORK_NEW["on"], True
        try:
            a = C_.conv1x1_dgrad_add(dy, wt, add, mask, None, bn_hint=64)
            b = C_.conv1x1_dgrad_add(dy, wt, add, mask, None, bn_hint=2064)
        finally:
            C_._FORK_NEW["on"] = was
        assert torch.equal(a, b)


@pytest.mark.parametrize("n,hw", [(4, 64), (2, 224), (3, 100)])
def test_stem_halo_variant_matches_the_per_tap_one(n, hw):
    """Hint 1064 on the stem: the four kernel-row taps read one 16 x 11-window box, the weights stay resident."""
    from layerwise_b200.ops import input as I
    torch.manual_seed(0)
    xu = torch.randint(0, 256, (n, 3, hw, hw), dtype=torch.uint8, device="cuda")
    w7 = (torch.randn(64, 3, 7, 7, device="cuda") / 12).to(torch.bfloat16)
    xs = I.normalize_u8(xu, s2d=True)
    a = I.stem_forward(xs, w7, want_stats=True, bn_hint=64)
    sa = I._LAST_STATS["v"]
    ta = sa.partials[:sa.nparts].sum(0)
    for hint in (1064, 3064):                      # halo; halo + two CTAs per SM
        b = I.stem_forward(xs, w7, want_stats=True, bn_hint=hint)
        sb = I._LAST_STATS["v"]
        assert float((a != b).float().mean()) < 1e-3
        tb = sb.partials[:sb.nparts].sum(0)
        assert (ta - tb).abs().max() <= 1e-3 * float(ta.abs().max())
