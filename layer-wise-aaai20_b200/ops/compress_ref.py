"""Reference-math compressors in plain PyTorch (CPU or GPU tensors).

This module is the numerical oracle of the framework and the implementation of
the CPU/gloo path.  Every function returns a *dense* tensor of the input's
shape, exactly like the study code does before its ``dist.all_reduce``
(reference: CIFAR10/core.py:175-225 layer-wise, :227-301 entire-model; the same
math is duplicated in IMAGENET/training/train_imagenet_nv.py:255-380).

Semantics kept from the reference:

* ``Topk``: threshold = the ceil(n*(1-K))-th smallest |g| (1-indexed); every
  element with |g| >= threshold survives, so ties all survive and without ties
  n - ceil(n*(1-K)) + 1 elements survive (core.py:178-183).
* ``Randomk``: a uniformly random subset of exactly ceil(n*K) elements, no
  1/K rescale (core.py:184-188).
* ``Thresholdv``: keep |g| >= V (core.py:189-193).
* ``AdaptiveThreshold``: keep |2g| >= max|g| (core.py:194-199).
* ``TernGrad``: sign(g) * max|g| * Bernoulli(|g|/max|g|) (core.py:200-206).
* ``RandomDithering`` (a.k.a. QSGD): sign(g) * ||g||_2 * floor(|g|/||g||*q + u)/q
  (core.py:207-213).
* K, V, qstates equal to 0/None and unknown method names mean "no compression"
  (core.py:214-215).

Deliberate deviations (documented in DESIGN.md):

* ``QSGD`` is accepted as an alias of ``RandomDithering`` (the reference README
  uses that name but the code silently ignores it).
* A zero-norm (all-zero) tensor quantises to zeros instead of NaN.
  Pass ``zero_guard=False`` to reproduce the reference's 0/0.
* ``Topk`` with K >= 1 keeps everything instead of raising.

Stochastic compressors accept an injected random buffer so that a CUDA kernel
and this oracle can consume identical random numbers in tests.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

# canonical names (the strings the reference CLI accepts, dawn.py:15-19)
TOPK = "Topk"
RANDOMK = "Randomk"
THRESHOLDV = "Thresholdv"
ADAPTIVE = "AdaptiveThreshold"
TERNGRAD = "TernGrad"
DITHERING = "RandomDithering"
NONE = "none"

METHODS = (TOPK, RANDOMK, THRESHOLDV, ADAPTIVE, TERNGRAD, DITHERING, NONE)
SPARSIFIERS = (TOPK, RANDOMK, THRESHOLDV, ADAPTIVE)
QUANTIZERS = (TERNGRAD, DITHERING)

_ALIASES = {
    "topk": TOPK, "top-k": TOPK, "top_k": TOPK,
    "randomk": RANDOMK, "randk": RANDOMK, "random-k": RANDOMK, "random_k": RANDOMK,
    "thresholdv": THRESHOLDV, "threshold-v": THRESHOLDV, "threshold": THRESHOLDV,
    "adaptivethreshold": ADAPTIVE, "adaptive": ADAPTIVE, "adaptive_threshold": ADAPTIVE,
    "terngrad": TERNGRAD,
    "randomdithering": DITHERING, "qsgd": DITHERING, "dithering": DITHERING,
    "none": NONE, "": NONE, "identity": NONE,
}


def canonical_method(name: Optional[str]) -> str:
    """Map a user-supplied method string onto one of :data:`METHODS`.

    Unknown strings map to ``"none"`` like the reference's final ``else``
    (core.py:214-215); ``QSGD`` maps to ``RandomDithering``.
    """
    if name is None:
        return NONE
    return _ALIASES.get(str(name).strip().lower(), NONE)


def effective_method(method: Optional[str], K=None, V=None, qstates=None) -> str:
    """Method after the reference's truthiness tests on K / V / qstates."""
    m = canonical_method(method)
    if m in (TOPK, RANDOMK) and not K:
        return NONE
    if m == THRESHOLDV and not V:
        return NONE
    if m == DITHERING and not qstates:
        return NONE
    return m


def topk_rank(n: int, K: float) -> int:
    """1-indexed ascending rank of the Top-k threshold, ceil(n*(1-K))."""
    return int(math.ceil(n * (1.0 - K)))


def topk_keep_count(n: int, K: float) -> int:
    """Number of survivors of ``Topk`` when no two |g| are equal."""
    r = topk_rank(n, K)
    if r < 1:
        return n
    return n - r + 1


def randomk_keep_count(n: int, K: float) -> int:
    """Number of survivors of ``Randomk``: #{p in 0..n-1 : p < n*K}."""
    return int(min(n, math.ceil(n * K)))


def _topk(g: torch.Tensor, K: float) -> torch.Tensor:
    n = g.numel()
    r = topk_rank(n, K)
    if r < 1:  # K >= 1: everything survives (the reference raises here)
        return g.clone()
    mag = g.abs()
    thres = mag.float().kthvalue(r).values.to(mag.dtype)
    out = g.clone()
    out[mag < thres] = 0
    return out


def _randomk(g: torch.Tensor, K: float, keys: Optional[torch.Tensor], generator) -> torch.Tensor:
    n = g.numel()
    if keys is None:
        perm = torch.randperm(n, device=g.device, generator=generator)
        mask = perm < n * K
    else:
        # injected integer keys: the k largest keys survive, lowest index wins ties
        k = randomk_keep_count(n, K)
        order = torch.sort(keys.view(-1).to(torch.int64), descending=True, stable=True).indices
        mask = torch.zeros(n, dtype=torch.bool, device=g.device)
        mask[order[:k]] = True
    return g * mask.to(g.dtype)


def _thresholdv(g: torch.Tensor, V: float) -> torch.Tensor:
    out = g.clone()
    out[g.abs() < V] = 0
    return out


def _adaptive(g: torch.Tensor) -> torch.Tensor:
    out = g.clone()
    if g.numel() == 0:
        return out
    gmax = g.abs().max()
    out[(g * 2).abs() < gmax] = 0
    return out


def _terngrad(g: torch.Tensor, rand: Optional[torch.Tensor], generator) -> torch.Tensor:
    if g.numel() == 0:
        return g.clone()
    mag = g.abs()
    s = mag.max()
    prob = mag / s  # NaN when s == 0; NaN comparisons are False -> zeros
    if rand is None:
        rand = torch.rand(g.shape, dtype=g.dtype, device=g.device, generator=generator)
    b = (rand.to(prob.dtype) < prob).to(g.dtype)
    return g.sign() * s * b


def _dithering(g: torch.Tensor, q: int, rand: Optional[torch.Tensor], generator,
               zero_guard: bool) -> torch.Tensor:
    if g.numel() == 0:
        return g.clone()
    norm = torch.norm(g.float()).to(g.dtype)
    if zero_guard and float(norm) == 0.0:
        return torch.zeros_like(g)
    if rand is None:
        rand = torch.rand(g.shape, dtype=g.dtype, device=g.device, generator=generator)
    level = torch.floor(g.abs() / norm * q + rand.to(g.dtype))
    out = g.sign() * norm * (level / q)
    return torch.where(torch.isinf(out), torch.zeros_like(out), out)


def compress_dense(g: torch.Tensor, method: Optional[str], K=None, V=None, qstates=None, *,
                   rand: Optional[torch.Tensor] = None, generator=None,
                   zero_guard: bool = True) -> torch.Tensor:
    """Compress the flat view of ``g`` and return it as a dense tensor.

    ``rand`` injects randomness: uniforms in [0,1) for TernGrad / dithering, integer
    keys for Random-k (the k largest keys survive).
    """
    shape = g.shape
    flat = g.reshape(-1)
    m = effective_method(method, K, V, qstates)
    if m == TOPK:
        out = _topk(flat, float(K))
    elif m == RANDOMK:
        out = _randomk(flat, float(K), rand, generator)
    elif m == THRESHOLDV:
        out = _thresholdv(flat, float(V))
    elif m == ADAPTIVE:
        out = _adaptive(flat)
    elif m == TERNGRAD:
        out = _terngrad(flat, None if rand is None else rand.reshape(-1), generator)
    elif m == DITHERING:
        out = _dithering(flat, int(qstates), None if rand is None else rand.reshape(-1),
                         generator, zero_guard)
    else:
        out = flat.clone()
    return out.view(shape)


def compress_with_feedback(g: torch.Tensor, residual: Optional[torch.Tensor], method, K=None,
                           V=None, qstates=None, **kw) -> torch.Tensor:
    """Error-feedback step: ``acc = g + e; c = C(acc); e <- acc - c``; returns ``c``.

    The reference has this only for Random-k
    (IMAGENET/training/sparsified_ddp.py:409-413); here it composes with every
    compressor.  ``residual`` is updated in place (fp32 recommended).
    """
    if residual is None:
        return compress_dense(g, method, K, V, qstates, **kw)
    acc = g.reshape(-1).to(residual.dtype) + residual.view(-1)
    c = compress_dense(acc, method, K, V, qstates, **kw)
    residual.view(-1).copy_(acc - c)
    return c.to(g.dtype).view(g.shape)
