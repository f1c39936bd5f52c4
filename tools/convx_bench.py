#!/usr/bin/env python
"""Native tcgen05 convolution kernels (csrc/convx.cu) vs cuDNN on all 23 convolution configurations of ResNet-50
(forward, data gradient, weight gradient), per output-tile width.  L2 is flushed before every timed call.

usage: convx_bench.py [batch] [out.json] [--quick]
"""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import conv as cv  # noqa: E402

# (cin, cout, k, stride, input hw, count in ResNet-50)
LAYERS = [(64, 256, 1, 1, 56, 4), (256, 64, 1, 1, 56, 2), (64, 64, 1, 1, 56, 1), (64, 64, 3, 1, 56, 3),
          (256, 128, 1, 1, 56, 1), (128, 128, 3, 2, 56, 1), (256, 512, 1, 2, 56, 1), (128, 512, 1, 1, 28, 4),
          (512, 128, 1, 1, 28, 3), (128, 128, 3, 1, 28, 3), (512, 256, 1, 1, 28, 1), (256, 256, 3, 2, 28, 1),
          (512, 1024, 1, 2, 28, 1), (256, 1024, 1, 1, 14, 6), (1024, 256, 1, 1, 14, 5), (256, 256, 3, 1, 14, 5),
          (1024, 512, 1, 1, 14, 1), (512, 512, 3, 2, 14, 1), (1024, 2048, 1, 2, 14, 1), (512, 2048, 1, 1, 7, 3),
          (2048, 512, 1, 1, 7, 2), (512, 512, 3, 1, 7, 2)]


def timeit(fn, flush, iters=5):
    """Median device time of ``fn`` replayed as a CUDA graph (no host launch overhead in the number), L2 flushed
    before every replay."""
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    del g
    ts.sort()
    return ts[len(ts) // 2]


def rel(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12))


def main():
    torch.backends.cudnn.benchmark = True
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    B = int(args[0]) if args else 224
    out = args[1] if len(args) > 1 else "gpurun_out/convx_bench.json"
    quick = "--quick" in sys.argv
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    tot = {"cudnn": [0.0, 0.0, 0.0], "ours": [0.0, 0.0, 0.0]}
    for ci, co, k, s, hw, cnt in (LAYERS[:6] if quick else LAYERS):
        pad = k // 2
        x = torch.randn(B, ci, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(co, ci, k, k, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        y = F.conv2d(x, w, None, s, pad)
        dy = torch.randn_like(y)
        gx_ref, gw_ref, _ = torch.ops.aten.convolution_backward(dy, x, w, None, [s, s], [pad, pad], [1, 1], False, [0, 0], 1,
                                                                [True, True, False])
        row = {"cin": ci, "cout": co, "k": k, "stride": s, "hw": hw, "count": cnt}
        row["cudnn"] = {
            "fwd": timeit(lambda: F.conv2d(x, w, None, s, pad), flush),
            "dgrad": timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [s, s], [pad, pad], [1, 1], False, [0, 0], 1,
                                                                        [True, False, False]), flush),
            "wgrad": timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [s, s], [pad, pad], [1, 1], False, [0, 0], 1,
                                                                        [False, True, False]), flush)}
        best = {"fwd": (1e9, 0), "fwd_stats": (1e9, 0), "dgrad": (1e9, 0), "wgrad": (1e9, 0)}
        row["ours"] = {}
        for bn in (64, 128, 256):
            r = {}
            try:
                if co % bn == 0:
                    yo, _ = cv.conv_forward(x, w, stride=s, padding=pad, bn_hint=bn)
                    r["fwd_err"] = rel(yo, y)
                    r["fwd"] = timeit(lambda: cv.conv_forward(x, w, stride=s, padding=pad, bn_hint=bn), flush)
                    r["fwd_stats"] = timeit(lambda: cv.conv_forward(x, w, stride=s, padding=pad, bn_hint=bn, want_stats=True), flush)
                    gw = cv.conv_wgrad(x, dy, w.shape, s, pad, bn_hint=bn)
                    r["wgrad_err"] = rel(gw, gw_ref)
                    r["wgrad"] = timeit(lambda: cv.conv_wgrad(x, dy, w.shape, s, pad, bn_hint=bn), flush)
                    for cap in (1, 2):
                        hint = cap * 1000 + bn
                        r[f"wgrad_cap{cap}"] = timeit(lambda: cv.conv_wgrad(x, dy, w.shape, s, pad, bn_hint=hint), flush)
                        r["wgrad"] = min(r["wgrad"], r[f"wgrad_cap{cap}"])
                if ci % bn == 0:
                    gx = cv.conv_dgrad(dy, w, (hw, hw), s, pad, bn_hint=bn)
                    r["dgrad_err"] = rel(gx, gx_ref)
                    r["dgrad"] = timeit(lambda: cv.conv_dgrad(dy, w, (hw, hw), s, pad, bn_hint=bn), flush)
            except Exception as e:  # noqa: BLE001
                r["error"] = repr(e)[:200]
            row["ours"][str(bn)] = r
            for key in best:
                if key in r and r[key] < best[key][0]:
                    best[key] = (r[key], bn)
        row["best"] = {kk: {"ms": v[0], "bn": v[1]} for kk, v in best.items()}
        byts = 2.0 * (x.numel() + y.numel())
        row["roof_ms"] = max(2.0 * y.numel() * ci * k * k / 1632e9, byts / 6540.5e6)
        rows.append(row)
        for i, kk in enumerate(("fwd", "dgrad", "wgrad")):
            tot["cudnn"][i] += cnt * row["cudnn"][kk]
            tot["ours"][i] += cnt * best[kk][0]
        errs = [v for r in row["ours"].values() for kk, v in r.items() if kk.endswith("_err")]
        print(f"{ci:5d}->{co:5d} k{k} s{s} @{hw:3d} x{cnt}: fwd {best['fwd'][0]:.3f}/{best['fwd_stats'][0]:.3f}(stats) bn{best['fwd'][1]} vs "
              f"{row['cudnn']['fwd']:.3f} | dgrad {best['dgrad'][0]:.3f} bn{best['dgrad'][1]} vs {row['cudnn']['dgrad']:.3f} | wgrad "
              f"{best['wgrad'][0]:.3f} bn{best['wgrad'][1]} vs {row['cudnn']['wgrad']:.3f} | roof {row['roof_ms']:.3f} | max err "
              f"{max(errs) if errs else -1:.4f}", flush=True)
    print("totals per step (ms): cudnn fwd/dgrad/wgrad", [round(v, 3) for v in tot["cudnn"]], "ours", [round(v, 3) for v in tot["ours"]])
    os.makedirs(os.path.dirname(out) or ".", exist_ok=True)
    with open(out, "w") as f:
        json.dump({"batch": B, "layers": rows, "totals": tot}, f, indent=1)


if __name__ == "__main__":
    main()
