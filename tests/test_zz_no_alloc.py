"""The allocation-free Top-k entry point must agree with the ordinary one.

``lwc_topk_select_f32_noalloc`` exists for callers that have no way to obtain device scratch:
code living under a memory pool that owns all of HBM, plug-ins that receive only pointers and
streams, or sections that cannot tolerate ``cudaMalloc``'s device-wide synchronisation.  Its
scratch is a fixed array inside the library, shared by every call, so this test makes two calls
in a row on different inputs (the second must not be confused by what the first left behind)
and compares each against ``lwc_topk_select_f32`` fed with caller-allocated workspace.  Both are
reached through the C ABI exactly as an external caller would.  The file sorts last on purpose:
if it fails under ``pytest -x`` every other test has already run.
"""
import ctypes as C

import pytest
import torch

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


def _select(lib, x, keep, cap, use_noalloc):
    n = x.numel()
    idx = torch.full((cap,), -1, dtype=torch.int32, device="cuda")
    val = torch.zeros(cap, dtype=torch.float32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if use_noalloc:
        rc = lib.lwc_topk_select_f32_noalloc(x.data_ptr(), n, keep, cap, idx.data_ptr(), val.data_ptr(),
                                             cnt.data_ptr(), st)
    else:
        ws = torch.empty(lib.lwc_topk_workspace_bytes(n), dtype=torch.uint8, device="cuda")
        rc = lib.lwc_topk_select_f32(x.data_ptr(), n, keep, cap, idx.data_ptr(), val.data_ptr(), cnt.data_ptr(),
                                     ws.data_ptr(), st)
    assert rc == 0, rc
    torch.cuda.synchronize()
    k = int(cnt.item())
    return k, idx[:k].clone(), val[:k].clone()


def test_noalloc_topk_matches_workspace_variant_twice_in_a_row():
    from layerwise_b200 import _native
    lib = _native.lib()
    g = torch.Generator(device="cuda").manual_seed(123)
    # two consecutive calls: different length, different keep count, different data
    cases = [(300001, 3001), (70000, 35)]
    for n, keep in cases:
        x = torch.randn(n, device="cuda", generator=g)
        cap = keep + 64
        k_ref, idx_ref, val_ref = _select(lib, x, keep, cap, use_noalloc=False)
        k_new, idx_new, val_new = _select(lib, x, keep, cap, use_noalloc=True)
        assert k_new == k_ref == keep
        assert torch.equal(idx_new, idx_ref)
        assert torch.allclose(val_new, val_ref, atol=1e-6, rtol=1e-6)
        # and both agree with a plain fp32 PyTorch top-k of |x|
        want = x.abs().topk(keep).indices.sort().values
        assert torch.equal(idx_new.long(), want)


def test_noalloc_topk_allocates_nothing_even_at_its_maximum_size():
    """The point of the entry point is that it obtains no device memory: free memory as the driver reports it, and
    torch's allocator statistics, must be the same before and after a call -- including a call at the documented
    maximum n = 2**24, which also shows that the selection state of such a tensor fits the built-in scratch."""
    from layerwise_b200 import _native
    lib = _native.lib()
    n, keep = 1 << 24, 4096
    cap = keep + 64
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(n, device="cuda", generator=g)
    idx = torch.full((cap,), -1, dtype=torch.int32, device="cuda")
    val = torch.zeros(cap, dtype=torch.float32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # first call outside the measured window: it looks the scratch address up (once per device)
    assert lib.lwc_topk_select_f32_noalloc(x.data_ptr(), n, keep, cap, idx.data_ptr(), val.data_ptr(), cnt.data_ptr(), st) == 0
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    alloc0, reserved0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    assert lib.lwc_topk_select_f32_noalloc(x.data_ptr(), n, keep, cap, idx.data_ptr(), val.data_ptr(), cnt.data_ptr(), st) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert torch.cuda.memory_allocated() == alloc0 and torch.cuda.memory_reserved() == reserved0
    k = int(cnt.item())
    assert k == keep
    want = x.abs().topk(keep).indices.sort().values
    assert torch.equal(idx[:k].long(), want)


def test_noalloc_topk_rejects_oversize_without_touching_outputs():
    from layerwise_b200 import _native
    lib = _native.lib()
    too_big = (1 << 24) + 1
    x = torch.zeros(8, device="cuda")           # never dereferenced: the size check comes first
    idx = torch.zeros(8, dtype=torch.int32, device="cuda")
    val = torch.zeros(8, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = lib.lwc_topk_select_f32_noalloc(x.data_ptr(), too_big, 10, 8, idx.data_ptr(), val.data_ptr(), cnt.data_ptr(),
                                         C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -3                              # LWC_ETOOBIG
    assert b"maximum" in lib.lwc_error_string(rc)
