"""Small helpers: nested-dict utilities and launcher helpers.

``union`` / ``path_iter`` serve the network DSL (reference CIFAR10/core.py:121-127);
``text_pickle`` / ``text_unpickle`` carry the phases list on a command line
(reference IMAGENET/util.py:48-61); ``is_set`` / ``random_id`` as in
IMAGENET/util.py:12-22.
"""
from __future__ import annotations

import base64
import os
import pickle
import random
import string


def union(*dicts):
    out = {}
    for d in dicts:
        out.update(d)
    return out


def path_iter(nested, prefix=()):
    """Depth-first walk yielding ``(path_tuple, leaf)`` for every non-dict leaf."""
    for name, val in nested.items():
        if isinstance(val, dict):
            yield from path_iter(val, prefix + (name,))
        else:
            yield prefix + (name,), val


def text_pickle(obj) -> str:
    return base64.b64encode(pickle.dumps(obj)).decode("ascii")


def text_unpickle(text: str):
    return pickle.loads(base64.b64decode(text.encode("ascii")))


def is_set(name: str) -> bool:
    """True if environment variable ``name`` is set to something other than 0/empty."""
    val = os.environ.get(name, "0")
    return val not in ("0", "")


def random_id(k: int = 5) -> str:
    return "".join(random.choices(string.ascii_lowercase + string.digits, k=k))
