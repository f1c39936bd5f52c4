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


# ---- launcher / environment helpers (reference IMAGENET/util.py:24-47, :64-82, :127-132) ---------------
def ossystem(cmd: str, shell: bool = True) -> str:
    """Run a shell command and return its combined output (like ``os.system`` with capture)."""
    import subprocess
    p = subprocess.Popen(cmd, shell=shell, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    out, _ = p.communicate()
    return out.decode("utf-8", errors="replace")


def log_environment(sink=None) -> dict:
    """Collect environment facts worth keeping with a run (versions, GPUs, relevant env vars)."""
    import platform
    import sys
    info = {"python": sys.version.split()[0], "platform": platform.platform()}
    try:
        import torch
        info["torch"] = torch.__version__
        info["cuda"] = torch.version.cuda
        if torch.cuda.is_available():
            info["gpus"] = [torch.cuda.get_device_name(i) for i in range(torch.cuda.device_count())]
    except Exception:
        pass
    for k, v in os.environ.items():
        if k.startswith(("NCCL_", "CUDA_", "WORLD_SIZE", "RANK", "LOCAL_RANK", "MASTER_")):
            info[f"env.{k}"] = v
    if sink is not None:
        for k, v in info.items():
            sink(f"{k}: {v}")
    return info


def format_env(**env) -> str:
    """``K1=v1 K2=v2`` prefix for a shell command."""
    return " ".join(f"{k}={v}" for k, v in env.items())


def format_env_export(**env) -> str:
    return " ".join(f"export {k}={v};" for k, v in env.items())


def run_parallel(fn, items, workers: int = 8):
    """Apply ``fn`` to every item on a thread pool; returns results in order."""
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=max(1, workers)) as ex:
        return list(ex.map(fn, items))
