"""ctypes binding of the native sm_100a library.

The library is plain C ABI (see ``include/lwc_b200.h``): every entry point takes raw
device pointers plus a ``cudaStream_t`` and returns a ``cudaError_t``-style int
(0 = success, negative = library-level argument error).  It is built in-tree by
``build()`` with ``nvcc -gencode arch=compute_100a,code=sm_100a`` so that the ``.so``
travels with the repository snapshot; nothing is JIT-compiled at run time.
"""
from __future__ import annotations

import ctypes
import glob
import os
import shutil
import subprocess
import sys
from typing import List, Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_HERE)
_REPO = os.path.dirname(_PKG)
CSRC = os.path.join(_PKG, "csrc")
INCLUDE = os.path.join(_REPO, "include")
LIB_PATH = os.path.join(_HERE, "liblwc_b200.so")

_lib: Optional[ctypes.CDLL] = None
_load_error: Optional[str] = None

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O3",
    "--expt-relaxed-constexpr", "-Xptxas", "-v",
]


class _Counters:
    """Host-side count of kernels this library launched (bench.py reports it as gpu_launches)."""

    def __init__(self):
        self.kernels = 0

    def reset(self):
        self.kernels = 0

    def add(self, n: int):
        self.kernels += n


counters = _Counters()


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found")
    return cand


def sources() -> List[str]:
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cpp")))


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) \
        + glob.glob(os.path.join(INCLUDE, "*.h"))
    return any(os.path.getmtime(s) > t for s in deps)


def build(force: bool = False, verbose: bool = False, jobs: int = 8) -> str:
    """Compile every ``csrc/*.cu|*.cpp`` for sm_100a and link ``liblwc_b200.so`` in-tree."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = _nvcc()
    objdir = os.path.join(_HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    from concurrent.futures import ThreadPoolExecutor

    def compile_one(src):
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
            glob.glob(os.path.join(INCLUDE, "*.h"))
        newest = max([os.path.getmtime(src)] + [os.path.getmtime(h) for h in hdrs])
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > newest:
            return obj, ""
        cmd = [nvcc, *NVCC_FLAGS, "-I", INCLUDE, "-I", CSRC, "-x", "cu", "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=jobs) as ex:
        results = list(ex.map(compile_one, sources()))
    objs = [o for o, _ in results]
    log = "\n".join(l for _, l in results if l)
    with open(os.path.join(objdir, "ptxas.log"), "a") as f:
        f.write(log)
    if verbose and log:
        print(log, file=sys.stderr)
    tmp = LIB_PATH + ".tmp"
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp, *objs,
           "-lcudart", "-lcuda"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


def _load() -> Optional[ctypes.CDLL]:
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        _load_error = f"{LIB_PATH} not built (run python -c 'import __graft_entry__ as g; g.build()')"
        return None
    try:
        import torch  # noqa: F401  (loads libcudart the library links against)
        _lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        from . import abi
        abi.declare(_lib)
    except OSError as e:  # e.g. no CUDA driver on a CPU-only box
        _load_error = str(e)
        _lib = None
    return _lib


def available() -> bool:
    """True when the library is loaded and a CUDA device is usable."""
    try:
        import torch
        if not torch.cuda.is_available():
            return False
    except Exception:
        return False
    return _load() is not None


def lib() -> ctypes.CDLL:
    l = _load()
    if l is None:
        raise RuntimeError(f"native library unavailable: {_load_error}")
    return l


def require():
    """Fail loudly when a CUDA device is present but the extension is missing."""
    import torch
    if torch.cuda.is_available() and _load() is None:
        raise RuntimeError("CUDA device present but liblwc_b200.so is not loadable: "
                           f"{_load_error}. Build it with __graft_entry__.build().")
    return _lib
