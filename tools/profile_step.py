#!/usr/bin/env python
"""Per-kernel device-time breakdown of one training step (CUPTI via torch.profiler; analysis only --
numbers quoted as results always come from bench.py's CUDA-event timing, never from a profiler)."""
import argparse
import collections
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.parallel.exchange import CompressionConfig  # noqa: E402
from layerwise_b200.train.imagenet import Trainer  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=224)
    ap.add_argument("--size", type=int, default=224)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--arch", default="resnet50")
    ap.add_argument("--top", type=int, default=45)
    ap.add_argument("--out", default="gpurun_out/step_profile.json")
    ap.add_argument("--timeline", default="", help="also write the kernels of the last profiled step in launch order (TSV)")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.01, error_feedback=True)
    tr = Trainer(arch=args.arch, dtype="bf16", cfg=cfg, lr=0.1, device="cuda:0")
    x = torch.randint(0, 256, (args.batch, 3, args.size, args.size), dtype=torch.uint8, device="cuda")
    y = torch.randint(0, 1000, (args.batch,), device="cuda")
    for _ in range(4):
        tr.train_step(x, y)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(args.steps):
            tr.train_step(x, y)
        torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0, 0.0])
    if args.timeline:
        evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda e: e.time_range.start)
        starts = [i for i, e in enumerate(evs) if "k_normalize" in e.name]
        last = evs[starts[-1]:] if starts else evs
        t0 = last[0].time_range.start
        with open(args.timeline, "w") as f:
            for e in last:
                dur = e.device_time if hasattr(e, "device_time") else e.cuda_time
                f.write(f"{e.time_range.start - t0:10.1f}\t{dur:8.1f}\t{e.name[:110]}\n")
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            agg[ev.name][0] += 1
            agg[ev.name][1] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
    tot = sum(v[1] for v in agg.values())
    rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
    print(f"total device time per step: {tot / args.steps / 1e3:.3f} ms over {sum(v[0] for v in agg.values()) / args.steps:.0f} kernels")
    cats = collections.OrderedDict([("bn (ours)", ("lwc::bn", "k_stats_finalize")),
                                    ("conv fwd/dgrad (ours, tcgen05)", ("lwc::convx::k_fx", "lwc::conv::k_conv")),
                                    ("conv wgrad (ours, tcgen05)", ("lwc::convx::k_wx", "lwc::stem")),
                                    ("exchange (ours)", "lwc::k_"), ("sgd (ours)", "lwc::k_sgd"),
                                    ("pool (ours)", "lwc::pool"), ("gemm (ours)", "lwc::gemm"),
                                    ("input/loss/misc (ours)", ("lwc::input", "lwc::loss", "lwc::k_ce", "lwc::")),
                                    ("memset/memcpy", ("Memset", "Memcpy")),
                                    ("conv/gemm (cudnn/cublas)", ("cutlass", "nvjet", "xmma", "cudnn", "implicit_gemm", "nhwc")),
                                    ("elementwise/other (aten)", "at::native")])
    csum = collections.defaultdict(float)
    for name, (c, t) in rows:
        for cat, pat in cats.items():
            pats = pat if isinstance(pat, tuple) else (pat,)
            if cat == "exchange (ours)" and "k_sgd" in name:
                continue
            if any(x in name for x in pats):
                csum[cat] += t
                break
        else:
            csum["unclassified"] += t
    for cat, t in csum.items():
        print(f"  [{cat:28s}] {t / args.steps / 1e3:8.3f} ms/step {100 * t / tot:5.1f}%")
    out = []
    for name, (c, t) in rows[:args.top]:
        print(f"{t / args.steps:10.1f} us/step {c / args.steps:7.1f} calls {100 * t / tot:5.1f}%  {name[:100]}")
        out.append({"kernel": name, "us_per_step": t / args.steps, "calls_per_step": c / args.steps})
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump({"total_ms_per_step": tot / args.steps / 1e3, "kernels": out}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
