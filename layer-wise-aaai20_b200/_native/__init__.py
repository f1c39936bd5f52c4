"""Loader for the in-tree native library ``liblwc_b200.so`` (built by ``build()``)."""
from .lib import available, lib, require, build, counters, LIB_PATH  # noqa: F401
