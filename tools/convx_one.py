#!/usr/bin/env python
"""Run ONE native convolution configuration a few times (for ncu / sanitizer captures).

usage: convx_one.py kind cin cout k stride hw [batch] [hint] [stats]     kind in fwd | dgrad | wgrad
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import conv as cv  # noqa: E402


def main():
    kind = sys.argv[1]
    ci, co, k, s, hw = (int(a) for a in sys.argv[2:7])
    B = int(sys.argv[7]) if len(sys.argv) > 7 else 224
    hint = int(sys.argv[8]) if len(sys.argv) > 8 else 0
    stats = len(sys.argv) > 9 and sys.argv[9] == "1"
    pad = k // 2
    x = torch.randn(B, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(co, ci, k, k, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    oh = (hw + 2 * pad - k) // s + 1
    dy = torch.randn(B, co, oh, oh, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    for _ in range(3):
        if kind == "fwd":
            cv.conv_forward(x, w, stride=s, padding=pad, bn_hint=hint, want_stats=stats)
        elif kind == "dgrad":
            cv.conv_dgrad(dy, w, (hw, hw), s, pad, bn_hint=hint)
        else:
            cv.conv_wgrad(x, dy, w.shape, s, pad, bn_hint=hint)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
