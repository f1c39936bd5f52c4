"""Convolution dispatch (native implicit-GEMM kernels where available)."""


def maybe_conv2d(module, x):
    """Return the native result, or ``None`` to fall back to the stock op."""
    return None
