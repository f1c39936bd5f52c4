#!/usr/bin/env python
"""Launch ONE named hot op a few times (the target of an `ncu --set full` capture, see profiles/README.md).

    python tools/ncu_target.py fx_fwd_stats         conv 64 -> 256 1x1 @ 56, bs 224, + batch statistics   (k_fx<128,false>)
    python tools/ncu_target.py fx_halo              conv 64 -> 64 3x3 @ 56, halo variant                   (k_fx<64,false>)
    python tools/ncu_target.py fx_dgrad_bnb         dgrad 64 -> 256 1x1 @ 56 + bn backward sums            (k_fx<64,true>)
    python tools/ncu_target.py fork_new | fork_old  residual fork 64 -> 256 @ 56
    python tools/ncu_target.py wx                   wgrad 128 -> 128 3x3 @ 28                              (k_wx)
    python tools/ncu_target.py bn_apply | bn_bwd    batch norm apply / backward apply on 256 ch @ 56
    python tools/ncu_target.py exchange             layer-wise Top-k 1 % send + receive over ResNet-50-sized tensors, one rank
    python tools/ncu_target.py step                 one whole ResNet-50 training step, batch 64 (all remaining kernel families)
    python tools/ncu_target.py misc                 SGD / pool / loss / input / stem / head / quantiser / dense all-reduce / small ops
    python tools/ncu_target.py all [reps]           every target in turn
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.ops import conv as C_  # noqa: E402


def t(*shape):
    x = torch.randn(*shape, device="cuda").to(torch.bfloat16)
    return x.contiguous(memory_format=torch.channels_last) if len(shape) == 4 else x


ALL = ["fx_fwd_stats", "fx_halo", "fx_dgrad_bnb", "fork_new", "fork_old", "wx", "bn_bwd", "exchange"]


def build(what):
    n = 224
    torch.manual_seed(0)
    if what == "fx_fwd_stats":
        x, w = t(n, 64, 56, 56), t(256, 64, 1, 1)
        fn = lambda: C_.conv_forward(x, w, want_stats=True, bn_hint=128)
    elif what == "fx_halo":
        x, w = t(n, 64, 56, 56), t(64, 64, 3, 3)
        fn = lambda: C_.conv_forward(x, w, want_stats=True, bn_hint=1064)
    elif what == "fx_dgrad_bnb":
        dy, w, xb = t(n, 256, 56, 56), t(256, 64, 1, 1), t(n, 64, 56, 56)
        link = dict(x=xb, gamma=torch.ones(64, device="cuda"), beta=torch.zeros(64, device="cuda"),
                    mean=torch.zeros(64, device="cuda"), invstd=torch.ones(64, device="cuda"), relu=True)
        fn = lambda: C_.conv_dgrad_bnstats(dy, w, (56, 56), 0, link, bn_hint=64)
    elif what in ("fork_new", "fork_old"):
        C_._FORK_NEW["on"] = what == "fork_new"
        d, w, add, xb = t(n, 64, 56, 56), t(64, 256, 1, 1), t(n, 256, 56, 56), t(n, 256, 56, 56)
        mask = torch.randint(0, 256, (n * 56 * 56 * 256 // 8,), device="cuda", dtype=torch.uint8)
        link = dict(x=xb, mean=torch.zeros(256, device="cuda"), invstd=torch.ones(256, device="cuda"))
        fn = lambda: C_.conv1x1_dgrad_add(d, w, add, mask, link if what == "fork_new" else None, bn_hint=64)
    elif what == "wx":
        x, dy = t(n, 128, 28, 28), t(n, 128, 28, 28)
        fn = lambda: C_.conv_wgrad(x, dy, (128, 128, 3, 3), 1, 1, out_dtype=torch.bfloat16)
    elif what == "wx1x1":
        # the 1x1 weight gradients of layer1: 256 -> 64 and 64 -> 256 @ 56 (k_wx<64> / k_wx<256>), fixed tile hints
        x1, d1 = t(n, 256, 56, 56), t(n, 64, 56, 56)

        def fn():
            C_.conv_wgrad(x1, d1, (64, 256, 1, 1), 1, 0, out_dtype=torch.bfloat16, bn_hint=64)
            C_.conv_wgrad(d1, x1, (256, 64, 1, 1), 1, 0, out_dtype=torch.bfloat16, bn_hint=256)
    elif what in ("bn_apply", "bn_bwd"):
        from layerwise_b200.ops.bn import batch_norm_act
        x = t(n, 256, 56, 56).requires_grad_(True)
        g, b = torch.ones(256, device="cuda", requires_grad=True), torch.zeros(256, device="cuda", requires_grad=True)
        rm, rv = torch.zeros(256, device="cuda"), torch.ones(256, device="cuda")
        dy = t(n, 256, 56, 56)

        def fn():
            y = batch_norm_act(x, g, b, rm, rv, training=True, relu=True)
            if what == "bn_bwd":
                y.backward(dy)
    elif what == "exchange":
        # one rank's side of the headline exchange: layer-wise Top-k 1 % over ResNet-50-sized tensors, error feedback on
        from layerwise_b200.parallel.fused import ExchangePlan, bind_local_world
        from layerwise_b200.parallel.exchange import CompressionConfig
        sizes = [2048 * 1000, 512 * 2048, 512 * 512 * 9, 2048 * 512] * 4 + [64 * 9 * 64, 256, 64] * 16
        cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.01, error_feedback=True)
        grads = [torch.randn(k, device="cuda").to(torch.bfloat16) for k in sizes]
        outs = [torch.empty(k, device="cuda", dtype=torch.bfloat16) for k in sizes]
        resid = [torch.zeros(k, device="cuda") for k in sizes]
        plan = ExchangePlan(grads, outs, resid, cfg, 1, 0, False)
        keep = bind_local_world([plan])
        step = [0]

        def fn():
            plan.send(step[0])
            plan.recv(step[0])
            step[0] += 1
    elif what == "step":
        # one whole ResNet-50 training step (batch 64 @ 224): every remaining kernel family -- input normalise, stem, max
        # pool, batch norm, adaptive average pool, GEMM head, cross entropy, exchange, SGD
        from layerwise_b200.parallel.exchange import CompressionConfig
        from layerwise_b200.train.imagenet import Trainer
        from layerwise_b200.ops import conv as C2
        C2._TUNE_ON["on"] = False                       # no autotune sweeps under the profiler
        cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.01, error_feedback=True)
        tr = Trainer(arch="resnet50", dtype="bf16", cfg=cfg, lr=0.1, device="cuda:0", overlap=True)
        x = torch.randint(0, 256, (64, 3, 224, 224), dtype=torch.uint8, device="cuda")
        y = torch.randint(0, 1000, (64,), device="cuda")
        fn = lambda: tr.train_step(x, y)
    elif what == "misc":
        # the kernel families without a capture so far: SGD, pools, cross entropy, input normalise, stem weight gradient,
        # GEMM head (one Top-k training step, batch 64), the quantiser send / dense receive, the dense two-shot all-reduce
        # (a `--compress none` step), dropout / ReLU / accuracy counters
        from layerwise_b200.parallel.exchange import CompressionConfig
        from layerwise_b200.parallel.fused import ExchangePlan, bind_local_world
        from layerwise_b200.train.imagenet import Trainer
        from layerwise_b200.ops import conv as C2, misc as M
        C2._TUNE_ON["on"] = False
        x = torch.randint(0, 256, (64, 3, 224, 224), dtype=torch.uint8, device="cuda")
        y = torch.randint(0, 1000, (64,), device="cuda")
        trs = [Trainer(arch="resnet50", dtype="bf16", device="cuda:0", lr=0.1, overlap=True,
                       cfg=CompressionConfig(compress=c, method="Topk", ratio=0.01, error_feedback=(c != "none")))
               for c in ("layerwise", "none")]
        sizes = [2048 * 1000, 512 * 2048, 512 * 512 * 9, 2048 * 512]
        plans = []
        for method in ("RandomDithering", "TernGrad"):
            cfg = CompressionConfig(compress="layerwise", method=method, qstates=255)
            grads = [torch.randn(k, device="cuda").to(torch.bfloat16) for k in sizes]
            outs = [torch.empty(k, device="cuda", dtype=torch.bfloat16) for k in sizes]
            pl = ExchangePlan(grads, outs, None, cfg, 1, 0, False)
            plans.append((pl, bind_local_world([pl]), grads, outs))
        act = torch.randn(64, 4096, device="cuda").to(torch.bfloat16).requires_grad_(True)
        logits = torch.randn(256, 1000, device="cuda").to(torch.bfloat16)
        step = [0]

        def fn():
            for tr in trs:
                tr.train_step(x, y)
            for pl, _, _, _ in plans:
                pl.send(step[0])
                pl.recv(step[0])
            step[0] += 1
            M.dropout(M.relu(act), 0.5).sum().backward()
            M.topk_hits(logits, y.repeat(4))
    else:
        raise SystemExit(__doc__)
    return fn


def main():
    what = sys.argv[1]
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    for w in (ALL if what == "all" else [what]):
        fn = build(w)
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        del fn
        torch.cuda.empty_cache()
        print("ok", w)


if __name__ == "__main__":
    main()
