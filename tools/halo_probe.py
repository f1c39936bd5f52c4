#!/usr/bin/env python
"""Halo-variant probe: numerics with and without the descriptor base-offset field, then timings of the halo variant
against the per-tap variants (forward / data gradient) and of the fork kernel against the older one.  One JSON line."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import conv as C_  # noqa: E402


def ms(fn, it=20):
    for _ in range(3):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    tot = 0.0
    for _ in range(it):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / it


def main():
    out = {}
    torch.manual_seed(0)
    n, c, hw = 32, 64, 56
    x = torch.randn(n, c, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(64, c, 3, 3, device="cuda") / 24.0).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ref = F.conv2d(x.float(), w.float(), padding=1)
    for bo in ("1", "0"):
        os.environ["LWC_HALO_BO"] = bo
        y, _ = C_.conv_forward(x, w, bn_hint=1064)
        torch.cuda.synchronize()
        out[f"halo_fwd_relerr_bo{bo}"] = float((y.float() - ref).abs().max() / ref.abs().max())
    good = min(("1", "0"), key=lambda b: out[f"halo_fwd_relerr_bo{b}"])
    os.environ["LWC_HALO_BO"] = good
    out["base_offset_used"] = good
    n = 224
    x = torch.randn(n, c, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(n, 64, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    for hint in (64, 1064):
        out[f"fwd_64x64_56_bs224_hint{hint}_ms"] = ms(lambda: C_.conv_forward(x, w, bn_hint=hint))
        out[f"fwd_stats_64x64_56_bs224_hint{hint}_ms"] = ms(lambda: C_.conv_forward(x, w, want_stats=True, bn_hint=hint))
        out[f"dgrad_64x64_56_bs224_hint{hint}_ms"] = ms(lambda: C_.conv_dgrad(dy, w, (hw, hw), 1, 1, bn_hint=hint))
    x1 = torch.randn(n, 64, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w1 = (torch.randn(256, 64, 1, 1, device="cuda") / 8.0).to(torch.bfloat16)
    x2 = torch.randn(n, 256, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w2 = (torch.randn(64, 256, 1, 1, device="cuda") / 16.0).to(torch.bfloat16)
    for hint in (64, 128, 256, 2064, 2128):
        out[f"fwd_64to256_56_hint{hint}_ms"] = ms(lambda: C_.conv_forward(x1, w1, bn_hint=hint))
        if True:
            out[f"fwd_stats_64to256_56_hint{hint}_ms"] = ms(lambda: C_.conv_forward(x1, w1, want_stats=True, bn_hint=hint))
        if hint != 256:
            out[f"dgrad_256to64_56_hint{hint}_ms"] = ms(lambda: C_.conv_dgrad(x1, w2.reshape(64, 256, 1, 1), (hw, hw), 1, 0, bn_hint=hint))
    for hint in (64, 2064):
        out[f"fwd_256to64_56_hint{hint}_ms"] = ms(lambda: C_.conv_forward(x2, w2, bn_hint=hint))
        out[f"fwd_stats_256to64_56_hint{hint}_ms"] = ms(lambda: C_.conv_forward(x2, w2, want_stats=True, bn_hint=hint))
    out["cudnn_fwd_64x64_56_bs224_ms"] = ms(lambda: F.conv2d(x, w, padding=1))
    # fork: layer1 (64 -> 256 @ 56) and layer3 (256 -> 1024 @ 14)
    C_._FORK_NEW["on"] = True
    for co, ci, s in ((64, 256, 56), (128, 512, 28), (256, 1024, 14), (512, 2048, 7)):
        wt = (torch.randn(co, ci, 1, 1, device="cuda") / co ** 0.5).to(torch.bfloat16)
        d = torch.randn(n, co, s, s, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        add = torch.randn(n, ci, s, s, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        xb = torch.randn(n, ci, s, s, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        mask = torch.randint(0, 256, (n * s * s * ci // 8,), device="cuda", dtype=torch.uint8)
        link = dict(x=xb, mean=torch.randn(ci, device="cuda"), invstd=torch.rand(ci, device="cuda") + 0.5)
        out[f"fork_{co}to{ci}_{s}_new_hint2064_ms"] = ms(lambda: C_.conv1x1_dgrad_add(d, wt, add, mask, None, bn_hint=2064))
        lk = dict(x=xb, gamma=None, beta=None, mean=link["mean"], invstd=link["invstd"], relu=True)
        for hint in [b_ for b_ in (64, 128, 2064) if ci % (b_ % 1000) == 0]:
            # data gradient of the block's LAST 1x1 (dy has ci channels, dx co channels) + bn2's backward sums
            lk2 = dict(x=d, gamma=None, beta=None, mean=link["mean"][:co], invstd=link["invstd"][:co], relu=True)
            wl = (torch.randn(ci, co, 1, 1, device="cuda") / ci ** 0.5).to(torch.bfloat16)
            if co % (hint % 1000) == 0:
                out[f"dgrad_bnb_{ci}to{co}_{s}_hint{hint}_ms"] = ms(lambda: C_.conv_dgrad_bnstats(add, wl, (s, s), 0, lk2, bn_hint=hint))
        for hint in (64, 128):
            out[f"fork_{co}to{ci}_{s}_new_link_hint{hint}_ms"] = ms(lambda: C_.conv1x1_dgrad_add(d, wt, add, mask, link, bn_hint=hint))
            out[f"fork_{co}to{ci}_{s}_new_hint{hint}_ms"] = ms(lambda: C_.conv1x1_dgrad_add(d, wt, add, mask, None, bn_hint=hint))
        C_._FORK_NEW["on"] = False
        out[f"fork_{co}to{ci}_{s}_old_ms"] = ms(lambda: C_.conv1x1_dgrad_add(d, wt, add, mask))
        C_._FORK_NEW["on"] = True
    print(json.dumps(out))


if __name__ == "__main__":
    main()
