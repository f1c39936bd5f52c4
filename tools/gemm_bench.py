#!/usr/bin/env python
"""tcgen05 GEMM (csrc/gemm.cu) vs cuBLAS (torch.matmul), CUDA-event timed, L2 flushed between iterations."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops.gemm import gemm  # noqa: E402

SHAPES = [  # (M, N, K, a_mn, b_mn, what)
    (512, 4096, 25088, False, False, "VGG fc1 fwd (bs 512)"),
    (512, 4096, 4096, False, False, "VGG/AlexNet fc2 fwd"),
    (512, 25088, 4096, False, True, "VGG fc1 dgrad"),
    (4096, 25088, 512, True, True, "VGG fc1 wgrad"),
    (224 * 56 * 56, 64, 256, False, False, "ResNet-50 1x1 256->64 @56 fwd"),
    (224 * 56 * 56, 256, 64, False, False, "ResNet-50 1x1 64->256 @56 fwd"),
    (224 * 14 * 14, 1024, 256, False, False, "ResNet-50 1x1 256->1024 @14 fwd"),
    (224 * 14 * 14, 256, 1024, False, True, "ResNet-50 1x1 dgrad @14"),
    (256, 1024, 224 * 14 * 14, True, True, "ResNet-50 1x1 wgrad @14 (split-K)"),
    (8192, 8192, 8192, False, False, "square 8192"),
]


def timeit(fn, flush, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()                      # write > L2 between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--one", action="store_true", help="run a single shape once (for ncu)")
    ap.add_argument("--out", default="gpurun_out/gemm_bench.json")
    args = ap.parse_args()
    torch.manual_seed(0)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = peaks.get("bf16_tflops", 1590.0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    for (M, N, K, a_mn, b_mn, what) in (SHAPES[-1:] if args.one else SHAPES):
        a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(torch.bfloat16)
        b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(torch.bfloat16)
        out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        if args.one:
            gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out)
            torch.cuda.synchronize()
            print("ran", M, N, K)
            return
        A = a.t() if a_mn else a
        Bt = b if b_mn else b.t()
        ref = (A.float() @ Bt.float()) if M * N * K < 2e11 else None
        mine = gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out)
        err = float((mine.float() - ref).abs().max() / ref.abs().max()) if ref is not None else None
        t_mine = timeit(lambda: gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out), flush)
        t_ref = timeit(lambda: torch.matmul(A, Bt, out=out), flush)
        fl = 2.0 * M * N * K
        row = {"shape": [M, N, K], "a_mn": a_mn, "b_mn": b_mn, "what": what, "ms_ours": t_mine, "ms_cublas": t_ref,
               "tflops_ours": fl / t_mine / 1e9, "tflops_cublas": fl / t_ref / 1e9,
               "frac_of_measured_peak": fl / t_mine / 1e9 / peak, "rel_err": err}
        rows.append(row)
        print(f"{what:38s} M={M:7d} N={N:6d} K={K:7d}  ours {t_mine:8.3f} ms {row['tflops_ours']:7.1f} TF/s "
              f"({100 * row['frac_of_measured_peak']:4.1f}% of measured)  cuBLAS {t_ref:8.3f} ms {row['tflops_cublas']:7.1f} TF/s  err {err}")
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
