"""Loader for the in-tree native library ``liblwc_b200.so`` (built by ``build.py``)."""
from .lib import available, lib, require, build, LIB_PATH  # noqa: F401
