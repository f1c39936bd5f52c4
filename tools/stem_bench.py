#!/usr/bin/env python
"""Device time of the stem weight-gradient kernel (csrc/stem.cu) vs cuDNN's wgrad on the plain 7x7 formulation."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import input as inp  # noqa: E402


def timed(fn, iters=10):
    """Sum of the device durations of the kernels one call launches (CUPTI; the call's host overhead exceeds
    its device time, so event timing would measure Python), L2 flushed before each call."""
    from torch.profiler import ProfilerActivity, profile
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(iters):
            flush.zero_()
            torch.cuda.synchronize()
            fn()
            torch.cuda.synchronize()
    us = 0.0
    names = {}
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA and "Memset" not in ev.name and "fill" not in ev.name.lower():
            us += ev.device_time
            names[ev.name[:60]] = names.get(ev.name[:60], 0.0) + ev.device_time / iters
    print({k: round(v, 1) for k, v in names.items()}, file=sys.stderr)
    return us / iters / 1e3


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 224
    size = int(sys.argv[2]) if len(sys.argv) > 2 else 224
    x = torch.randint(0, 256, (n, 3, size, size), dtype=torch.uint8, device="cuda")
    xs = inp.normalize_u8(x, torch.bfloat16, s2d=True)
    dy = torch.randn(n, 64, size // 2, size // 2, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w7 = torch.zeros(64, 3, 7, 7, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    x3 = torch.randn(n, 3, size, size, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    torch.backends.cudnn.benchmark = True
    ours = timed(lambda: inp.stem_wgrad(dy, xs, torch.bfloat16))
    cudnn = timed(lambda: torch.ops.aten.convolution_backward(dy, x3, w7, None, [2, 2], [3, 3], [1, 1], False, [0, 0], 1,
                                                               [False, True, False]))
    if int(os.environ.get("LWC_STEM_DEBUG", "0")) & 16:
        ws = list(inp._WS.values())[0]
        st = ws.view(torch.int64)[-32:].cpu().tolist()
        tail = ws.numel() * 4 - 256
        st = ws.view(torch.uint8)[tail:tail + 72].view(torch.int64).cpu().tolist()
        print("issuer clocks: wait_b %d wait_pair %d issue %d commit %d" % tuple(st[5:9]), file=sys.stderr)
        print("stamps (us): producer_done %.1f issuer_done %.1f epilogue_start %.1f end %.1f" %
              tuple((t - st[0]) / 1e3 for t in st[1:5]), file=sys.stderr)
    mb = (dy.numel() * 2 + xs.numel() * 2) / 1e6
    out = {"batch": n, "size": size, "ours_ms": ours, "cudnn_7x7_ms": cudnn, "bytes_mb": mb,
           "ours_tb_s": mb / 1e6 / (ours / 1e3), "roofline_ms_at_6.54TBs": mb / 1e6 / 6.5405 * 1e3}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
