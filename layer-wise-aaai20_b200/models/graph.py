"""Nested-dict network DSL and its executor.

A network is a nested dict whose leaves are modules or ``(module, [input names])``
pairs; ``build_graph`` flattens it to ``name -> (module, inputs)`` where a leaf
without explicit inputs consumes the previous leaf (the first consumes ``input``).
``Network`` runs the nodes in insertion order and returns the dict of all
activations.  Behaviour follows the reference (CIFAR10/core.py:121-141,
CIFAR10/torch_backend.py:69-118); the implementation is independent.
"""
from __future__ import annotations

from collections import namedtuple
from typing import Dict, List, Tuple

import torch
import torch.nn as nn

from ..utils.misc import path_iter, union  # noqa: F401  (re-exported)

SEP = "_"
RelativePath = namedtuple("RelativePath", ("parts",))


def rel_path(*parts):
    return RelativePath(parts)


def _resolve(ref, prefix: Tuple[str, ...]) -> str:
    if isinstance(ref, RelativePath):
        return SEP.join(tuple(prefix) + tuple(ref.parts))
    if isinstance(ref, str):
        return ref
    return SEP.join(ref)


def build_graph(net: Dict) -> Dict[str, Tuple[nn.Module, List[str]]]:
    flat = list(path_iter(net))
    graph = {}
    previous = "input"
    for path, leaf in flat:
        prefix, name = path[:-1], SEP.join(path)
        if isinstance(leaf, tuple):
            module, inputs = leaf
            inputs = [_resolve(i, prefix) for i in inputs]
        else:
            module, inputs = leaf, [previous]
        graph[name] = (module, inputs)
        previous = name
    return graph


class Identity(nn.Module):
    def forward(self, x):
        return x


class Mul(nn.Module):
    def __init__(self, weight):
        super().__init__()
        self.weight = weight

    def forward(self, x):
        if x.is_cuda:
            from ..ops import misc as _misc
            from ..ops.nn import get_backend
            if get_backend() != "torch":
                return _misc.scale(x, self.weight)
        return x * self.weight


class Flatten(nn.Module):
    def forward(self, x):
        return x.reshape(x.size(0), x.size(1))


class Add(nn.Module):
    def forward(self, x, y):
        return x + y


class Concat(nn.Module):
    def forward(self, *xs):
        return torch.cat(xs, 1)


class Correct(nn.Module):
    def forward(self, classifier, target):
        if classifier.is_cuda:
            from ..ops import misc as _misc
            from ..ops.nn import get_backend
            if get_backend() != "torch":
                return _misc.correct(classifier.detach(), target)
        return classifier.max(dim=1)[1] == target


class Network(nn.Module):
    """Executes a graph; ``forward(dict)`` returns every intermediate by name."""

    def __init__(self, net: Dict):
        super().__init__()
        self.graph = build_graph(net)
        for name, (module, _) in self.graph.items():
            setattr(self, name, module)

    def forward(self, inputs: Dict):
        cache = dict(inputs)
        for name, (_, srcs) in self.graph.items():
            cache[name] = getattr(self, name)(*[cache[s] for s in srcs])
        self.cache = cache
        return cache


def trainable_params(model: nn.Module):
    return [p for p in model.parameters() if p.requires_grad]


# ----- graph export (reference core.py:346-411, pydot/graphviz) -------------------------------
def to_dot(graph: Dict[str, Tuple[nn.Module, List[str]]], name: str = "net") -> str:
    """Plain-text Graphviz ``dot`` source of a built graph (no graphviz/pydot needed)."""
    lines = [f"digraph {name} {{", "  rankdir=LR;", '  node [shape=box, style="rounded,filled", fillcolor="#eeeeee"];']
    nodes = set(graph)
    for node, (module, inputs) in graph.items():
        label = f"{node}\\n{type(module).__name__}"
        lines.append(f'  "{node}" [label="{label}"];')
        for src in inputs:
            if src not in nodes:
                lines.append(f'  "{src}" [shape=ellipse, fillcolor="#ffffff"];')
            lines.append(f'  "{src}" -> "{node}";')
    lines.append("}")
    return "\n".join(lines)


def remove_by_type(net: Dict, node_type) -> Dict:
    """Drop every node of ``node_type`` from a *built graph*, re-wiring consumers to its input."""
    graph = build_graph(net) if not all(isinstance(v, tuple) and isinstance(v[1], list)
                                        for v in net.values()) else dict(net)
    alias = {}
    for name, (module, inputs) in graph.items():
        if isinstance(module, node_type):
            alias[name] = inputs[0]

    def root(n):
        while n in alias:
            n = alias[n]
        return n

    return {n: (m, [root(i) for i in ins]) for n, (m, ins) in graph.items()
            if not isinstance(m, node_type)}
