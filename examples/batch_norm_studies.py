#!/usr/bin/env python
"""Script counterparts of the experiments in the reference's batch-norm notebook (CIFAR10/batch_norm_post.ipynb).

    lr-stress       "Experiment: LR stress test" (cells 17-19): train the same small conv net with and without batch
                    norm over a range of learning rates, report where training diverges.
    channel-stats   "Deep nets and constant functions" / "Per channel histograms" (cells 24-41): per-channel mean and
                    spread of the activations of a deep ReLU net at initialisation, with and without batch norm, and
                    the singular values of its layer Jacobian proxies (the SVD aside, cell 37).
    hessian         "Computing the Hessian" (cells 107-115): leading eigenvalues of the loss Hessian by Lanczos on
                    Hessian-vector products, with and without batch norm.
    lars            the notebook's inlined LARS / Nesterov update rules (its private library cells) as one optimiser
                    step function, compared against plain SGD on the same problem.

Everything runs on synthetic CIFAR-shaped data on CPU or GPU in seconds (`--steps`, `--width` scale it up); with
`--data_dir` the real CIFAR-10 pipeline of this package is used instead.  None of these studies touch the compressed
exchange -- exactly like the notebook.
"""
from __future__ import annotations

import argparse
import math
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401


def make_net(width=32, depth=4, bn=True, bn_after_relu=False, classes=10):
    """conv3x3 (-> BN) -> ReLU blocks with two 2x2 pools, global max pool, linear head (the notebook's simple net)."""
    layers, c_in = [], 3
    for d in range(depth):
        layers.append(nn.Conv2d(c_in, width, 3, padding=1, bias=not bn))
        if bn and not bn_after_relu:
            layers.append(nn.BatchNorm2d(width))
        layers.append(nn.ReLU())
        if bn and bn_after_relu:
            layers.append(nn.BatchNorm2d(width))
        if d in (1, 3):
            layers.append(nn.MaxPool2d(2))
        c_in = width
    layers += [nn.AdaptiveMaxPool2d(1), nn.Flatten(), nn.Linear(width, classes)]
    return nn.Sequential(*layers)


def synthetic(n, device, seed=0):
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, 10, (n,), generator=g)
    # class-dependent mean image + noise: learnable, so "diverged" vs "trained" is visible in a few steps
    proto = torch.randn(10, 3, 32, 32, generator=g)
    x = proto[y] * 0.5 + torch.randn(n, 3, 32, 32, generator=g)
    return x.to(device), y.to(device)


def train(net, x, y, lr, steps, bs=128, momentum=0.9, update=None):
    params = [p for p in net.parameters() if p.requires_grad]
    bufs = [torch.zeros_like(p) for p in params]
    losses = []
    for s in range(steps):
        i = (s * bs) % (x.shape[0] - bs + 1)
        loss = F.cross_entropy(net(x[i:i + bs]), y[i:i + bs])
        if not torch.isfinite(loss):
            return losses + [float("inf")]
        grads = torch.autograd.grad(loss, params)
        with torch.no_grad():
            for p, g, b in zip(params, grads, bufs):
                if update is None:
                    b.mul_(momentum).add_(g)
                    p.add_(g + momentum * b, alpha=-lr)          # Nesterov SGD, the dawn.py recipe
                else:
                    update(p, g, b, lr, momentum)
        losses.append(float(loss))
    return losses


# ---- LR stress test -------------------------------------------------------------------------------------------------
def lr_stress(args):
    x, y = synthetic(args.samples, args.device)
    print(f"{'lr':>8s} {'with BN':>12s} {'without BN':>12s}   (final loss after {args.steps} steps; inf = diverged)")
    out = []
    for lr in [10 ** e for e in (-3, -2.5, -2, -1.5, -1, -0.5, 0)]:
        row = []
        for bn in (True, False):
            torch.manual_seed(0)
            net = make_net(args.width, args.depth, bn).to(args.device)
            row.append(train(net, x, y, lr, args.steps)[-1])
        out.append((lr, *row))
        print(f"{lr:8.4f} {row[0]:12.4f} {row[1]:12.4f}")
    return out


# ---- activations of a deep ReLU net at initialisation --------------------------------------------------------------
def channel_stats(args):
    x, _ = synthetic(512, args.device)
    res = {}
    for bn in (False, True):
        torch.manual_seed(0)
        net = make_net(args.width, args.depth, bn).to(args.device).train()
        h, stats = x, []
        with torch.no_grad():
            for m in net:
                h = m(h)
                if isinstance(m, nn.ReLU):
                    per_ch_mean = h.mean((0, 2, 3))
                    # how much of each channel's output varies with the INPUT: spread over images of the spatial mean
                    over_inputs = h.mean((2, 3)).std(0)
                    stats.append((float(per_ch_mean.mean()), float(over_inputs.mean() / (per_ch_mean.abs().mean() + 1e-12))))
        res[bn] = stats
        print(f"batch norm = {bn}")
        for d, (mu, rel) in enumerate(stats):
            print(f"   after ReLU {d}: mean activation {mu:8.4f}   input-dependent spread / mean {rel:6.3f}"
                  + ("   <- nearly a constant function" if rel < 0.1 else ""))
    # the SVD aside: singular values of a random conv layer seen as a (Cout) x (Cin * 9) matrix, and of ReLU(W x)
    torch.manual_seed(0)
    w = nn.Conv2d(args.width, args.width, 3).weight.detach().reshape(args.width, -1)
    sv = torch.linalg.svdvals(w)
    z = torch.randn(4096, w.shape[1])
    sv_relu = torch.linalg.svdvals(F.relu(z @ w.t()) / math.sqrt(z.shape[0]))
    print(f"singular values of a He-initialised layer: max {float(sv[0]):.3f} min {float(sv[-1]):.3f};  of ReLU(Wx) over random "
          f"inputs: max {float(sv_relu[0]):.3f} (the mean direction) second {float(sv_relu[1]):.3f}")
    return res, (sv, sv_relu)


# ---- leading Hessian eigenvalues --------------------------------------------------------------------------------------
def hessian_top_k(net, x, y, k=3, iters=30):
    """Top-k eigenvalues of the loss Hessian: Lanczos on Hessian-vector products (double backward)."""
    params = [p for p in net.parameters() if p.requires_grad]
    n = sum(p.numel() for p in params)
    loss = F.cross_entropy(net(x), y)
    grads = torch.autograd.grad(loss, params, create_graph=True)
    flat_g = torch.cat([g.reshape(-1) for g in grads])

    def hvp(v):
        hv = torch.autograd.grad(flat_g, params, v, retain_graph=True)
        return torch.cat([h.reshape(-1) for h in hv]).detach()

    m = min(iters, n)
    q = torch.randn(n, device=x.device)
    q /= q.norm()
    Q, alpha, beta = [q], [], []
    for j in range(m):
        w = hvp(Q[-1])
        a = torch.dot(w, Q[-1])
        w = w - a * Q[-1] - (beta[-1] * Q[-2] if beta else 0)
        for qq in Q:                       # full re-orthogonalisation: m is small
            w = w - torch.dot(w, qq) * qq
        b = w.norm()
        alpha.append(float(a))
        if j == m - 1 or float(b) < 1e-8:
            break
        beta.append(b)
        Q.append(w / b)
    T = torch.diag(torch.tensor(alpha))
    for i, b in enumerate(beta[:len(alpha) - 1]):
        T[i, i + 1] = T[i + 1, i] = float(b)
    ev = torch.linalg.eigvalsh(T)
    return ev.flip(0)[:k].tolist()


def hessian(args):
    x, y = synthetic(256, args.device)
    out = {}
    for bn in (True, False):
        torch.manual_seed(0)
        net = make_net(args.width, args.depth, bn).to(args.device).train()
        ev = hessian_top_k(net, x, y, k=3, iters=args.lanczos)
        out[bn] = ev
        print(f"batch norm = {bn}: leading Hessian eigenvalues at initialisation {['%.3f' % e for e in ev]}"
              f"  -> plain gradient descent is stable for lr < {2.0 / max(ev[0], 1e-12):.4f}")
    return out


# ---- LARS --------------------------------------------------------------------------------------------------------------
def lars_update(p, g, buf, lr, momentum, trust=0.02, wd=5e-4, eps=1e-9):
    """Layer-wise adaptive rate scaling with Nesterov momentum: the local rate is trust * |w| / (|g| + wd |w|)."""
    wn, gn = p.norm(), g.norm()
    local = trust * wn / (gn + wd * wn + eps) if float(wn) > 0 and float(gn) > 0 else 1.0
    d = (g + wd * p) * local
    buf.mul_(momentum).add_(d)
    p.add_(d + momentum * buf, alpha=-lr)


def lars(args):
    x, y = synthetic(args.samples, args.device)
    out = {}
    for name, upd, lr in (("nesterov SGD", None, 0.05), ("LARS + nesterov", lars_update, 1.0)):
        torch.manual_seed(0)
        net = make_net(args.width, args.depth, True).to(args.device)
        losses = train(net, x, y, lr, args.steps, update=upd)
        out[name] = losses
        print(f"{name:18s} lr {lr:5.2f}: loss {losses[0]:.3f} -> {losses[-1]:.3f}")
    return out


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("study", choices=["lr-stress", "channel-stats", "hessian", "lars", "all"])
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--samples", type=int, default=2048)
    ap.add_argument("--width", type=int, default=32)
    ap.add_argument("--depth", type=int, default=4)
    ap.add_argument("--lanczos", type=int, default=20)
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    args = ap.parse_args(argv)
    studies = {"lr-stress": lr_stress, "channel-stats": channel_stats, "hessian": hessian, "lars": lars}
    res = {}
    for name in (studies if args.study == "all" else [args.study]):
        print(f"==== {name}")
        res[name] = studies[name](args)
    return res


if __name__ == "__main__":
    main()
