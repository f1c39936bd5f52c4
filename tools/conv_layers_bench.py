#!/usr/bin/env python
"""Per-layer cost of ResNet-50's convolutions through the stock library path (cuDNN / cuBLAS via torch), bf16
channels-last, batch 224: forward, dgrad, wgrad times (CUDA events, L2 flushed), FLOPs and minimum HBM bytes,
so that each layer's distance from its roofline is visible.  Used to decide which shapes deserve hand-written
tcgen05 kernels."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200 import models  # noqa: E402


def timeit(fn, flush, iters=5):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    torch.backends.cudnn.benchmark = True
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 224
    net = models.resnet50()
    shapes = {}
    hooks = []

    def mk(name):
        def hook(m, inp, out):
            key = (m.in_channels, m.out_channels, m.kernel_size[0], m.stride[0], inp[0].shape[2])
            shapes.setdefault(key, []).append(name)
        return hook
    for n, m in net.named_modules():
        if isinstance(m, torch.nn.Conv2d):
            hooks.append(m.register_forward_hook(mk(n)))
    with torch.no_grad():
        net.eval()(torch.zeros(1, 3, 224, 224))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    tot = {"fwd": 0.0, "dgrad": 0.0, "wgrad": 0.0, "roof": 0.0}
    peak_tf, peak_bw = 1632.0, 6540.5
    for (cin, cout, k, s, hw), names in sorted(shapes.items(), key=lambda kv: -len(kv[1])):
        x = torch.randn(B, cin, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = torch.randn(cout, cin, k, k, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        pad = k // 2
        y = F.conv2d(x, w, None, s, pad)
        dy = torch.randn_like(y)
        t_f = timeit(lambda: F.conv2d(x, w, None, s, pad), flush)
        t_d = timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [s, s], [pad, pad], [1, 1], False, [0, 0], 1, [True, False, False]), flush) if cin > 3 else 0.0
        t_w = timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [s, s], [pad, pad], [1, 1], False, [0, 0], 1, [False, True, False]), flush)
        flops = 2.0 * y.numel() * cin * k * k
        byts = 2.0 * (x.numel() + y.numel() + w.numel())
        roof = max(flops / peak_tf / 1e9, byts / peak_bw / 1e6)      # ms, per pass
        n = len(names)
        rows.append({"cin": cin, "cout": cout, "k": k, "stride": s, "hw": hw, "count": n, "ms_fwd": t_f, "ms_dgrad": t_d,
                     "ms_wgrad": t_w, "gflop": flops / 1e9, "mbytes": byts / 1e6, "roof_ms": roof})
        tot["fwd"] += n * t_f; tot["dgrad"] += n * t_d; tot["wgrad"] += n * t_w; tot["roof"] += n * roof * (3 if cin > 3 else 2)
        print(f"{n}x conv{k}x{k}/{s} {cin:4d}->{cout:4d} @{hw:3d}  fwd {t_f:6.3f}  dgrad {t_d:6.3f}  wgrad {t_w:6.3f} ms | "
              f"roof/pass {roof:6.3f} ms ({flops / 1e9:7.1f} GF, {byts / 1e6:6.1f} MB)", flush=True)
    print(f"TOTAL per step: fwd {tot['fwd']:.2f}  dgrad {tot['dgrad']:.2f}  wgrad {tot['wgrad']:.2f}  = {tot['fwd'] + tot['dgrad'] + tot['wgrad']:.2f} ms; "
          f"sum of rooflines {tot['roof']:.2f} ms")
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"batch": B, "layers": rows, "totals": tot}, open("gpurun_out/conv_layers_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
