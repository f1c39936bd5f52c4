#!/usr/bin/env python
"""Per-shape device time and achieved bandwidth of the fused batch-norm kernels on ResNet-50's layer shapes
(CUPTI kernel durations via torch.profiler, L2 flushed between calls; analysis tool)."""
import collections
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import bn as bnops  # noqa: E402

# (H, C, residual, count in ResNet-50)
SHAPES = [(112, 64, False, 1), (56, 64, False, 6), (56, 256, False, 1), (56, 256, True, 3), (56, 128, False, 1),
          (28, 128, False, 7), (28, 512, False, 1), (28, 512, True, 4), (28, 256, False, 1), (14, 256, False, 11),
          (14, 1024, False, 1), (14, 1024, True, 6), (14, 512, False, 1), (7, 512, False, 5), (7, 2048, False, 1),
          (7, 2048, True, 3)]
PEAK = 6540.5


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 224
    dev = torch.device("cuda:0")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    from torch.profiler import ProfilerActivity, profile
    rows, total = [], collections.defaultdict(float)
    only = os.environ.get("BN_ONLY")            # e.g. "56,256,0": one shape (for ncu captures)
    shapes = SHAPES if not only else [s for s in SHAPES if ",".join(map(str, (s[0], s[1], int(s[2])))) == only]
    for H, C, res, cnt in shapes:
        x = torch.randn(N, C, H, H, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        x.requires_grad_(True)
        r = torch.randn_like(x) if res else None
        w = torch.ones(C, device=dev, requires_grad=True)
        b = torch.zeros(C, device=dev, requires_grad=True)
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        dy = torch.randn_like(x)
        relu = True

        def once():
            flush.zero_()
            y = bnops.batch_norm_act(x, w, b, rm, rv, r, True, 0.1, 1e-5, relu)
            flush.zero_()
            y.backward(dy)
            x.grad = None
        for _ in range(2):
            once()
        torch.cuda.synchronize()
        reps = 5
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(reps):
                once()
            torch.cuda.synchronize()
        t = collections.defaultdict(float)
        for ev in prof.events():
            if ev.device_type == torch.autograd.DeviceType.CUDA and "lwc::bn" in ev.name:
                t[ev.name.split("lwc::bn::")[1].split("<")[0].split("(")[0]] += ev.device_time / reps
        A = N * H * H * C * 2 / 1e6       # MB of one activation pass
        traffic = {"k_bn_stats": A, "k_bn_apply": (3 if res else 2) * A, "k_bn_bwd_reduce": 2 * A,
                   "k_bn_bwd_apply": (4 if res else 3) * A}
        line = f"{H:4d}x{H:<4d} C={C:5d} res={int(res)} A={A:7.1f} MB |"
        for k in ("k_bn_stats", "k_bn_finalize", "k_bn_apply", "k_bn_bwd_reduce", "k_bn_bwd_finalize", "k_bn_bwd_apply"):
            us = t.get(k, 0.0)
            total[k] += us * cnt
            if k in traffic:
                gbs = traffic[k] / 1e3 / (us / 1e6) if us else 0
                line += f" {k[5:]:>10s} {us:6.1f}us {100 * gbs / PEAK:3.0f}%"
            else:
                line += f" {k[5:]:>12s} {us:4.1f}us"
        print(line)
        rows.append({"H": H, "C": C, "res": res, "count": cnt, "us": dict(t), "act_mb": A})
    print("per-step totals (us):", {k: round(v, 1) for k, v in total.items()}, "sum", round(sum(total.values()), 1))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"batch": N, "rows": rows, "totals_us": total}, open("gpurun_out/bn_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
