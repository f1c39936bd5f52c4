#!/usr/bin/env python
"""Device time of the fused residual-fork data gradient  dX = dY * W + d_identity  (csrc/conv.cu, lwc_conv1x1_add) on
ResNet-50's sixteen bottleneck shapes vs the bytes it must move (CUPTI kernel durations, L2 flushed)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops.conv import conv1x1_dgrad_add  # noqa: E402

# (H, Cin of the block, planes, count)
SHAPES = [(56, 64, 64, 1), (56, 256, 64, 2), (56, 256, 128, 1), (28, 512, 128, 3), (28, 512, 256, 1), (14, 1024, 256, 5),
          (14, 1024, 512, 1), (7, 2048, 512, 2)]
PEAK = 6540.5


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 224
    from torch.profiler import ProfilerActivity, profile
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    tot_us = tot_roof = 0.0
    rows = []
    for H, cin, planes, cnt in SHAPES:
        dy = torch.randn(N, planes, H, H, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(planes, cin, 1, 1, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        add = torch.randn(N, cin, H, H, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        for _ in range(2):
            out = conv1x1_dgrad_add(dy, w, add)
        ref = torch.nn.functional.conv_transpose2d(dy.float(), w.float()) + add.float()
        err = float((out.float() - ref).abs().max() / ref.abs().max())
        torch.cuda.synchronize()
        reps = 5
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(reps):
                flush.zero_()
                torch.cuda.synchronize()
                conv1x1_dgrad_add(dy, w, add)
                torch.cuda.synchronize()
        us = sum(ev.device_time for ev in prof.events()
                 if ev.device_type == torch.autograd.DeviceType.CUDA and "k_conv" in ev.name) / reps
        mb = (dy.numel() + 2 * add.numel()) * 2 / 1e6
        roof = mb / 1e6 / PEAK * 1e9          # us at the measured copy bandwidth
        print(f"{H:3d}x{H:<3d} Cin {cin:5d} planes {planes:4d} x{cnt}: {us:7.1f} us  {mb / 1e3 / (us / 1e6) / 1e3:5.2f} TB/s "
              f"({100 * roof / us:3.0f}% of copy peak)  err {err:.4f}")
        tot_us += us * cnt
        tot_roof += roof * cnt
        rows.append({"H": H, "cin": cin, "planes": planes, "count": cnt, "us": us, "mbytes": mb, "err": err})
    print(f"per-step total {tot_us:.0f} us, at copy peak {tot_roof:.0f} us")
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"batch": N, "rows": rows, "total_us": tot_us, "roofline_us": tot_roof}, open("gpurun_out/fork_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
