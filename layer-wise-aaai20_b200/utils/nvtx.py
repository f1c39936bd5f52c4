"""NVTX ranges for Nsight timelines (no-ops without CUDA).  The reference has no profiler hooks at all
(SURVEY 5.1); phases of the step (fwd / bwd / exchange / opt) are annotated with these."""
from __future__ import annotations

import contextlib

import torch


@contextlib.contextmanager
def nvtx_range(name: str):
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()
