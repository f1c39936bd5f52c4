#!/usr/bin/env python
"""Native persistent tcgen05 conv (csrc/conv.cu) vs cuDNN on ResNet-50's stride-1 1x1 / 3x3 layers (fwd and dgrad)."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops.conv import conv_forward, flipped_transposed  # noqa: E402

LAYERS = [(64, 256, 1, 56), (256, 64, 1, 56), (64, 64, 1, 56), (256, 128, 1, 56), (128, 512, 1, 28), (512, 128, 1, 28),
          (512, 256, 1, 28), (256, 1024, 1, 14), (1024, 256, 1, 14), (1024, 512, 1, 14), (512, 2048, 1, 7), (2048, 512, 1, 7),
          (64, 64, 3, 56), (128, 128, 3, 28), (256, 256, 3, 14), (512, 512, 3, 7)]


def timeit(fn, flush, iters=7):
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
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    for ci, co, k, hw in LAYERS:
        x = torch.randn(B, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(co, ci, k, k, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        pad = k // 2
        y = F.conv2d(x, w, None, 1, pad)
        dy = torch.randn_like(y)
        mine, _ = conv_forward(x, w, want_stats=True)
        err = float((mine.float() - y.float()).abs().max() / y.float().abs().max())
        t_ref = timeit(lambda: F.conv2d(x, w, None, 1, pad), flush)
        t_me = timeit(lambda: conv_forward(x, w, want_stats=True), flush)
        t_me0 = timeit(lambda: conv_forward(x, w, want_stats=False), flush)
        t_dref = timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [1, 1], [pad, pad], [1, 1], False, [0, 0], 1, [True, False, False]), flush)
        if k == 1:
            t_dme = timeit(lambda: conv_forward(dy, w, b_mn=True, out_channels=ci), flush)
        else:
            wd = flipped_transposed(w)
            t_dme = timeit(lambda: conv_forward(dy, wd), flush)
        byts = 2.0 * (x.numel() + y.numel())
        roof = max(2.0 * y.numel() * ci * k * k / 1632e9, byts / 6540.5e6)
        rows.append({"cin": ci, "cout": co, "k": k, "hw": hw, "fwd_ours_ms": t_me, "fwd_ours_nostats_ms": t_me0, "fwd_cudnn_ms": t_ref, "dgrad_ours_ms": t_dme,
                     "dgrad_cudnn_ms": t_dref, "roof_ms": roof, "rel_err": err})
        print(f"conv{k}x{k} {ci:4d}->{co:4d} @{hw:2d}: fwd+stats ours {t_me:6.3f} (no stats {t_me0:6.3f}) cudnn {t_ref:6.3f} | dgrad ours {t_dme:6.3f} cudnn {t_dref:6.3f} | roof {roof:6.3f} ms | err {err:.4f}", flush=True)
    json.dump(rows, open("gpurun_out/conv_native_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
