#!/usr/bin/env python
"""Write the Graphviz source of a model-zoo graph network (the reference renders SVG through pydot,
CIFAR10/core.py:355-403; here it is plain text -- pipe it into `dot -Tsvg` where graphviz exists)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200 import models  # noqa: E402

if __name__ == "__main__":
    net = models.union(models.resnet9(), models.losses())
    print(models.to_dot(models.build_graph(net), "resnet9"))
