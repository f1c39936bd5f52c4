"""Operators: reference-math compressors, fused CUDA compress/exchange kernels, SGD, layers."""
from . import compress_ref  # noqa: F401
from .compress_ref import (compress_dense, compress_with_feedback, canonical_method,  # noqa: F401
                           effective_method, METHODS)
