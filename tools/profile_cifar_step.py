#!/usr/bin/env python
"""Kernel-time sum vs wall time of one CIFAR ResNet-9 training step (is the step launch/CPU bound?)."""
import collections
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200.models import build_cifar_model, trainable_params  # noqa: E402
from layerwise_b200.ops.sgd import SGD  # noqa: E402
from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger  # noqa: E402
from layerwise_b200.utils.mixed import bn_to_float  # noqa: E402

net = sys.argv[1] if len(sys.argv) > 1 else "resnet9"
model = bn_to_float(build_cifar_model(net).cuda().to(memory_format=torch.channels_last).to(torch.bfloat16))
cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.01, error_feedback=True)
ex = GradientExchanger(trainable_params(model), cfg, overlap=True)
opt = SGD(trainable_params(model), lr=1e-4, momentum=0.9, weight_decay=5e-4 * 512, nesterov=True)
x = torch.randn(512, 3, 32, 32, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 10, (512,), device="cuda")


def step():
    out = model({"input": x, "target": y})
    out["loss"].sum().backward()
    ex.exchange()
    opt.step()
    ex.engine.zero_grad()


for _ in range(5):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    step()
torch.cuda.synchronize()
wall = (time.perf_counter() - t0) / 20 * 1e3
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        agg[ev.name][0] += 1
        agg[ev.name][1] += ev.device_time
tot = sum(v[1] for v in agg.values()) / 3 / 1e3
n = sum(v[0] for v in agg.values()) / 3
print(f"{net}: wall {wall:.3f} ms/step, kernel-time sum {tot:.3f} ms over {n:.0f} kernels -> GPU busy {100 * tot / wall:.0f}%")
for name, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:12]:
    print(f"  {t / 3:8.1f} us {c / 3:5.1f}x {name[:90]}")
