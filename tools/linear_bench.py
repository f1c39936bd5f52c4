#!/usr/bin/env python
"""Linear-layer GEMMs of the model zoo: cuBLAS (F.linear) vs the tcgen05 GEMM kernels (csrc/gemm.cu) vs the tap-table
convolution kernels run as 1x1 convolutions (csrc/convx.cu).  Graph-replay device times, L2 flushed."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import conv as cv  # noqa: E402
from layerwise_b200.ops.gemm import gemm  # noqa: E402
from convx_bench import timeit, rel  # noqa: E402

# (M = batch, K = in features, N = out features, where)
SHAPES = [(512, 25088, 4096, "VGG-16 fc1"), (512, 4096, 4096, "VGG-16 / AlexNet fc2"), (512, 1024, 4096, "AlexNet fc1"),
          (224, 2048, 1000, "ResNet-50 fc"), (512, 4096, 64, "10-class head padded to 64"), (512, 512, 64, "ResNet-9 head padded")]


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/linear_bench.json"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    for M, K, N, name in SHAPES:
        x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(torch.bfloat16)
        dy = torch.randn(M, N, device="cuda").to(torch.bfloat16)
        x4 = x.view(M, K, 1, 1).contiguous(memory_format=torch.channels_last)
        w4 = w.view(N, K, 1, 1).contiguous(memory_format=torch.channels_last)
        dy4 = dy.view(M, N, 1, 1).contiguous(memory_format=torch.channels_last)
        r = {"M": M, "K": K, "N": N, "layer": name}
        yr, dxr, dwr = x.float() @ w.float().t(), dy.float() @ w.float(), dy.float().t() @ x.float()
        r["cublas"] = {"fwd": timeit(lambda: F.linear(x, w), flush), "dgrad": timeit(lambda: dy @ w, flush),
                       "wgrad": timeit(lambda: dy.t() @ x, flush)}
        g = {}
        try:
            if N % 8 == 0:
                g["fwd"] = timeit(lambda: gemm(x, w), flush)
                g["fwd_err"] = rel(gemm(x, w), yr)
                g["dgrad"] = timeit(lambda: gemm(dy, w, b_mn=True), flush)
                g["wgrad"] = timeit(lambda: gemm(dy, x, a_mn=True, b_mn=True), flush)
        except Exception as e:  # noqa: BLE001
            g["error"] = repr(e)[:200]
        r["gemm"] = g
        c = {}
        for bn in (64, 128, 256):
            if N % bn == 0:
                try:
                    y, _ = cv.conv_forward(x4, w4, stride=1, padding=0, bn_hint=bn)
                    c[f"fwd{bn}"] = timeit(lambda: cv.conv_forward(x4, w4, stride=1, padding=0, bn_hint=bn), flush)
                    c["fwd_err"] = rel(y.view(M, N), yr)
                    c[f"wgrad{bn}"] = timeit(lambda: cv.conv_wgrad(x4, dy4, w4.shape, 1, 0, bn_hint=bn), flush)
                    c["wgrad_err"] = rel(cv.conv_wgrad(x4, dy4, w4.shape, 1, 0, bn_hint=bn).reshape(N, K), dwr)
                except Exception as e:  # noqa: BLE001
                    c[f"error{bn}"] = repr(e)[:200]
            if K % bn == 0:
                try:
                    dx = cv.conv_dgrad(dy4, w4, (1, 1), 1, 0, bn_hint=bn)
                    c[f"dgrad{bn}"] = timeit(lambda: cv.conv_dgrad(dy4, w4, (1, 1), 1, 0, bn_hint=bn), flush)
                    c["dgrad_err"] = rel(dx.view(M, K), dxr)
                except Exception as e:  # noqa: BLE001
                    c[f"derror{bn}"] = repr(e)[:200]
        r["convx"] = c
        rows.append(r)
        fmt = lambda d: {k: (round(v, 4) if isinstance(v, float) else v) for k, v in d.items()}  # noqa: E731
        print(name, (M, K, N), "\n   cublas", fmt(r["cublas"]), "\n   gemm  ", fmt(g), "\n   convx ", fmt(c), flush=True)
    with open(out, "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
