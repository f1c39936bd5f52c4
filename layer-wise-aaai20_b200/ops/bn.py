"""Batch-norm dispatch (fused native kernels where available)."""


def maybe_batch_norm(module, x):
    return None
