"""Fused CUDA exchange kernels vs the fp32 PyTorch oracle (single GPU, W virtual ranks)."""
import math

import pytest
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200.ops import compress_ref as cr
from layerwise_b200.parallel.exchange import CompressionConfig

pytestmark = pytest.mark.gpu

SIZES = [64, 1000, 8192, 8193, 50000, 300001]


def _world(W, sizes, cfg, entire, dtype=torch.float32, ef=False, seed=0):
    from layerwise_b200.parallel.fused import ExchangePlan, bind_local_world
    g = torch.Generator(device="cuda").manual_seed(seed)
    grads = [[torch.randn(n, device="cuda", generator=g).to(dtype) for n in sizes] for _ in range(W)]
    outs = [[torch.empty(n, device="cuda", dtype=dtype) for n in sizes] for _ in range(W)]
    resids = [[torch.zeros(n, device="cuda") for n in sizes] if ef else None for _ in range(W)]
    plans = [ExchangePlan(grads[r], outs[r], resids[r], cfg, W, r, entire) for r in range(W)]
    lw = bind_local_world(plans)
    return grads, outs, resids, plans, lw


def _run(plans, step):
    for p in plans:
        p.send(step)
    for p in plans:
        p.recv(step)
    torch.cuda.synchronize()


def _oracle(grads, cfg, entire, rand=None, resids=None):
    W = len(grads)
    T = len(grads[0])
    if entire:
        flat = []
        for r in range(W):
            f = torch.cat([t.float() for t in grads[r]])
            rr = None if resids is None else resids[r]
            rd = None if rand is None else torch.cat([x for x in rand[r]])
            flat.append(cr.compress_with_feedback(f, rr, cfg.method_name, cfg.ratio, cfg.threshold, cfg.qstates, rand=rd))
        s = sum(flat) / W
        out, off = [], 0
        for t in grads[0]:
            out.append(s[off:off + t.numel()])
            off += t.numel()
        return out
    out = []
    for i in range(T):
        acc = 0
        for r in range(W):
            rr = None if resids is None else resids[r][i]
            rd = None if rand is None else rand[r][i]
            acc = acc + cr.compress_with_feedback(grads[r][i].float(), rr, cfg.method_name, cfg.ratio,
                                                  cfg.threshold, cfg.qstates, rand=rd)
        out.append(acc / W)
    return out


@pytest.mark.parametrize("W", [1, 2, 4])
@pytest.mark.parametrize("entire", [False, True])
@pytest.mark.parametrize("method,kw", [("Topk", dict(ratio=0.01)), ("Topk", dict(ratio=0.3)),
                                       ("Thresholdv", dict(threshold=1.5)),
                                       ("AdaptiveThreshold", {}), ("none", {})])
def test_deterministic_methods_match_oracle(W, entire, method, kw):
    cfg = CompressionConfig(compress="entiremodel" if entire else "layerwise", method=method,
                            capacity_factor=1.5, dense_capacity=1.0, **kw)
    grads, outs, _, plans, lw = _world(W, SIZES, cfg, entire)
    _run(plans, 0)
    exp = _oracle(grads, cfg, entire)
    for r in range(W):
        for e, o in zip(exp, outs[r]):
            assert torch.allclose(e, o, atol=1e-6, rtol=1e-6)
        for o0, o in zip(outs[0], outs[r]):
            assert torch.equal(o0, o)        # bitwise-identical replicas
    st = plans[0].stats()
    assert st["dropped"] == 0 and st["timeouts"] == 0


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_topk_low_precision_grads_with_ties(dtype):
    cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.05, capacity_factor=4.0)
    grads, outs, _, plans, lw = _world(2, [4096, 70000], cfg, False, dtype=dtype)
    _run(plans, 0)
    exp = _oracle(grads, cfg, False)
    for e, o in zip(exp, outs[0]):
        assert torch.allclose(e, o.float(), atol=2e-2, rtol=1e-2)
    assert plans[0].stats()["dropped"] == 0


def test_topk_zero_gradient_sends_nothing():
    cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.1)
    grads, outs, _, plans, lw = _world(2, [5000], cfg, False)
    for r in range(2):
        grads[r][0].zero_()
        grads[r][0][:7] = torch.arange(1.0, 8.0, device="cuda")
    _run(plans, 0)
    assert torch.allclose(outs[0][0], grads[0][0])     # threshold 0: everything non-zero survives
    assert plans[0].stats()["sent"] == 7


def test_error_feedback_multi_step_matches_oracle():
    cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.02, error_feedback=True,
                            capacity_factor=2.0)
    W, sizes = 2, [1000, 20000]
    grads, outs, resids, plans, lw = _world(W, sizes, cfg, False, ef=True)
    ref_res = [[torch.zeros(n, device="cuda") for n in sizes] for _ in range(W)]
    g = torch.Generator(device="cuda").manual_seed(5)
    for step in range(4):
        for r in range(W):
            for t in grads[r]:
                t.copy_(torch.randn(t.shape, device="cuda", generator=g))
        exp = _oracle(grads, cfg, False, resids=ref_res)
        _run(plans, step)
        for e, o in zip(exp, outs[0]):
            assert torch.allclose(e, o, atol=1e-6)
        for r in range(W):
            for a, b in zip(ref_res[r], resids[r]):
                assert torch.allclose(a, b, atol=1e-6)


@pytest.mark.parametrize("entire", [False, True])
def test_randomk_injected_keys_exact(entire):
    cfg = CompressionConfig(compress="entiremodel" if entire else "layerwise", method="Randomk", ratio=0.03)
    W, sizes = 2, [1000, 30000]
    grads, outs, _, plans, lw = _world(W, sizes, cfg, entire)
    g = torch.Generator(device="cuda").manual_seed(9)
    keys = [[torch.randint(0, 2**31 - 1, (n,), device="cuda", generator=g, dtype=torch.int64).to(torch.int32)
             for n in sizes] for _ in range(W)]
    for r in range(W):
        plans[r].set_random(keys[r])
    _run(plans, 0)
    exp = _oracle(grads, cfg, entire, rand=[[k.to(torch.int64) for k in keys[r]] for r in range(W)])
    for e, o in zip(exp, outs[0]):
        assert torch.allclose(e, o, atol=1e-6)


def test_randomk_counter_rng_exact_count_and_shared_mask():
    cfg = CompressionConfig(compress="layerwise", method="Randomk", ratio=0.01)
    W, sizes = 2, [1000, 123457]
    grads, outs, _, plans, lw = _world(W, sizes, cfg, False)
    for r in range(W):
        for t in grads[r]:
            t.abs_().add_(1.0)          # no zeros
    _run(plans, 0)
    for n, o in zip(sizes, outs[0]):
        assert int((o != 0).sum()) == cr.randomk_keep_count(n, 0.01)    # same mask on both ranks
    first = [o.clone() for o in outs[0]]
    _run(plans, 1)
    assert not torch.equal((first[1] != 0), (outs[0][1] != 0))          # new mask every step


@pytest.mark.parametrize("method,q", [("TernGrad", 0), ("RandomDithering", 255), ("RandomDithering", 15)])
@pytest.mark.parametrize("entire", [False, True])
def test_quantisers_injected_uniforms(method, q, entire):
    cfg = CompressionConfig(compress="entiremodel" if entire else "layerwise", method=method, qstates=q or 255,
                            error_feedback=True)
    W, sizes = 2, [777, 40000]
    grads, outs, resids, plans, lw = _world(W, sizes, cfg, entire, ef=True)
    g = torch.Generator(device="cuda").manual_seed(11)
    us = [[torch.rand(n, device="cuda", generator=g) for n in sizes] for _ in range(W)]
    for r in range(W):
        plans[r].set_random(us[r])
    _run(plans, 0)
    if entire:
        ref_res = [torch.zeros(sum(sizes), device="cuda") for _ in range(W)]
    else:
        ref_res = [[torch.zeros(n, device="cuda") for n in sizes] for _ in range(W)]
    exp = _oracle(grads, cfg, entire, rand=us, resids=ref_res)
    for e, o in zip(exp, outs[0]):
        # the norm is reduced in a different order than torch.norm: allow one-level flips to be rare
        bad = (~torch.isclose(e, o, atol=1e-5, rtol=1e-4)).float().mean()
        assert bad < 2e-3
    for o0, o1 in zip(outs[0], outs[1]):
        assert torch.equal(o0, o1)


def test_quantisers_counter_rng_unbiased():
    cfg = CompressionConfig(compress="layerwise", method="RandomDithering", qstates=7)
    grads, outs, _, plans, lw = _world(1, [20000], cfg, False)
    acc = torch.zeros(20000, device="cuda")
    reps = 200
    for s in range(reps):
        _run(plans, s)
        acc += outs[0][0]
    err = (acc / reps - grads[0][0]).abs().mean()
    assert err < 0.5 * grads[0][0].norm() / 7 / math.sqrt(reps) * 3


def test_capacity_clamp_keeps_residual():
    cfg = CompressionConfig(compress="layerwise", method="Thresholdv", threshold=0.1, error_feedback=True,
                            dense_capacity=0.05)
    grads, outs, resids, plans, lw = _world(1, [40000], cfg, False, ef=True)
    g0 = grads[0][0].clone()
    _run(plans, 0)
    st = plans[0].stats()
    assert st["dropped"] > 0 and st["sent"] == plans[0].cap[0]
    assert torch.allclose(outs[0][0] + resids[0][0], g0, atol=1e-6)     # nothing lost


def test_standalone_topk_select():
    import ctypes as C
    from layerwise_b200 import _native
    lib = _native.lib()
    n, keep = 200000, 2001
    x = torch.randn(n, device="cuda")
    ws = torch.empty(lib.lwc_topk_workspace_bytes(n), dtype=torch.uint8, device="cuda")
    idx = torch.empty(keep + 64, dtype=torch.int32, device="cuda")
    val = torch.empty(keep + 64, dtype=torch.float32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.lwc_topk_select_f32(x.data_ptr(), n, keep, keep + 64, idx.data_ptr(), val.data_ptr(), cnt.data_ptr(),
                                 ws.data_ptr(), C.c_void_p(st))
    assert rc == 0
    torch.cuda.synchronize()
    k = int(cnt.item())
    assert k == keep
    ref = x.abs().topk(keep).indices.sort().values
    assert torch.equal(idx[:k].long(), ref)
    assert torch.equal(val[:k], x[ref])


def test_missing_peer_times_out_instead_of_hanging():
    """A receiver whose peer never sends must leave its kernel after the configured timeout and count it."""
    from layerwise_b200 import _native
    lib = _native.lib()
    cfg = CompressionConfig(compress="layerwise", method="Topk", ratio=0.1)
    grads, outs, _, plans, lw = _world(2, [4096], cfg, False)
    lib.lwc_set_flag_timeout_ms(50)
    try:
        plans[0].send(0)
        plans[0].recv(0)            # rank 1 never sent step 0
        torch.cuda.synchronize()
        assert plans[0].stats()["timeouts"] >= 1
        # fail closed: the time-out poisoned the step, so the fused optimizer must not touch weights or momentum
        import ctypes as C
        from layerwise_b200.ops.sgd import FusedSGD
        v = C.c_uint(0)
        assert lib.lwc_poison_read(C.byref(v)) == 0 and v.value == 1
        w = [torch.randn(5000, device="cuda", requires_grad=True), torch.randn(70, 3, device="cuda", requires_grad=True)]
        opt = FusedSGD(w, lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-4)
        lib.lwc_poison_clear()
        for t in w:
            t.grad = torch.randn_like(t)
        opt.step()                   # a normal first step (creates the momentum buffers)
        before = [t.detach().clone() for t in w]
        mom = [opt.state[t]["momentum_buffer"].clone() for t in w]
        plans[0].send(1)
        plans[0].recv(1)            # times out again -> poisoned
        opt.step()
        torch.cuda.synchronize()
        for t, b0, m0 in zip(w, before, mom):
            assert torch.equal(t.detach(), b0) and torch.equal(opt.state[t]["momentum_buffer"], m0)
        lib.lwc_poison_clear()
        opt.step()
        assert not torch.equal(w[0].detach(), before[0])       # cleared: updates resume
    finally:
        lib.lwc_poison_clear()
        lib.lwc_set_flag_timeout_ms(10000)


def _tiny_net():
    torch.manual_seed(0)
    # GEMM-only network: its backward is bitwise reproducible run to run (cuDNN wgrad kernels are not), so the
    # three exchange modes can be compared exactly
    net = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(3 * 16 * 16, 256), torch.nn.ReLU(),
                              torch.nn.Linear(256, 64), torch.nn.ReLU(), torch.nn.Linear(64, 10)).cuda()
    return net


@pytest.mark.parametrize("method,kw", [("Topk", dict(ratio=0.05)), ("none", {})])
def test_hook_overlap_and_grad_steal_match_plain_exchange(method, kw):
    """Hook-driven (side stream, stolen gradient tensors) and after-backward exchanges give identical averages."""
    from layerwise_b200.parallel.exchange import GradientExchanger
    x = torch.randn(8, 3, 16, 16, device="cuda")
    y = torch.randint(0, 10, (8,), device="cuda")
    results = []
    for overlap, steal in ((False, False), (True, True), (False, True)):
        net = _tiny_net()
        cfg = CompressionConfig(compress="layerwise" if method != "none" else "none", method=method, error_feedback=True, **kw)
        ex = GradientExchanger(net, cfg, engine="fused", overlap=overlap, steal=steal, dense="slots")
        outs = []
        for _ in range(3):
            if hasattr(ex.engine, "zero_grad"):
                ex.engine.zero_grad()
            torch.nn.functional.cross_entropy(net(x), y).backward()
            ex.exchange()
            outs.append([p.grad.detach().clone() for p in net.parameters()])
        results.append(outs)
        ex.engine.close()
    for other in results[1:]:
        for sa, sb in zip(results[0], other):
            for a, b in zip(sa, sb):
                assert torch.equal(a, b)


@pytest.mark.parametrize("method,kw", [("Topk", dict(ratio=0.01)), ("Thresholdv", dict(threshold=1e-30)),
                                       ("AdaptiveThreshold", {})])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_no_error_feedback_never_drops_at_default_capacity(method, kw, dtype):
    """Reference semantics without error feedback: every tie at the k-th value and every |g| >= V is sent
    (CIFAR10/core.py:178-193).  All-equal magnitudes and a tiny V select (almost) the whole tensor; with the default
    CompressionConfig the slots must hold it -- nothing may be cut."""
    cfg = CompressionConfig(compress="layerwise", method=method, **kw)          # default capacities, EF off
    grads, outs, _, plans, lw = _world(2, [4096, 50000], cfg, False, dtype=dtype)
    for r in range(2):
        grads[r][0].copy_(torch.where(torch.rand(4096, device="cuda") < 0.5, -0.5, 0.5).to(dtype))     # all |g| equal
    _run(plans, 0)
    exp = _oracle(grads, cfg, False)
    for e, o in zip(exp, outs[0]):
        assert torch.allclose(e, o.float(), atol=2e-2 if dtype == torch.bfloat16 else 1e-6, rtol=1e-2)
    assert plans[0].stats()["dropped"] == 0 and plans[1].stats()["dropped"] == 0


@pytest.mark.parametrize("method,kw", [("Topk", dict(ratio=0.01)), ("Randomk", dict(ratio=0.05)), ("TernGrad", {}),
                                       ("RandomDithering", dict(qstates=255)), ("none", {})])
def test_eight_virtual_ranks(method, kw):
    """W = 8 (the box size): every rank's receiver sees eight sources; replicas must agree bit for bit."""
    cfg = CompressionConfig(compress="layerwise", method=method, **kw)
    W = 8
    grads, outs, _, plans, lw = _world(W, [64, 8193, 100000], cfg, False)
    _run(plans, 0)
    for r in range(1, W):
        for o0, o in zip(outs[0], outs[r]):
            assert torch.equal(o0, o)
    if method in ("Topk", "none"):
        exp = _oracle(grads, cfg, False)
        for e, o in zip(exp, outs[0]):
            assert torch.allclose(e, o, atol=1e-6, rtol=1e-6)
    st = plans[0].stats()
    assert st["timeouts"] == 0 and st["dropped"] == 0


@pytest.mark.parametrize("W", [1, 2, 4])
@pytest.mark.parametrize("entire", [False, True])
@pytest.mark.parametrize("ef", [False, True])
def test_shared_seed_randomk_ships_values_only(W, entire, ef):
    """Shared-seed Random-k: 4 bytes per kept element on the wire (no indices), zeros included, the receiver rebuilds
    the positions from the hash (IMAGENET/training/sparsified_ddp.py:410-412).  Expected result: mean over ranks of
    g_r * mask with the mask every rank derives for (seed, step, tensor)."""
    cfg = CompressionConfig(compress="entiremodel" if entire else "layerwise", method="Randomk", ratio=0.02,
                            error_feedback=ef, shared_seed=True)
    sizes = [64, 8193, 100001]
    grads, outs, resids, plans, lw = _world(W, sizes, cfg, entire, ef=ef)
    assert all(p.values_only for p in plans)
    for r in range(W):
        grads[r][1][::3] = 0.0                      # exact zeros must travel too (positions are value-independent)
    g0 = [[t.clone() for t in grads[r]] for r in range(W)]
    for step in range(2):
        # the mask of this step: a one-rank plan with all-ones gradients returns exactly the mask
        ones = [torch.ones(n, device="cuda") for n in sizes]
        mo = [torch.empty(n, device="cuda") for n in sizes]
        from layerwise_b200.parallel.fused import ExchangePlan, bind_local_world
        cfg1 = CompressionConfig(compress=cfg.compress, method="Randomk", ratio=0.02, shared_seed=True)
        mp = ExchangePlan(ones, mo, None, cfg1, 1, 0, entire)
        keep = bind_local_world([mp])
        mp.send(step)
        mp.recv(step)
        _run(plans, step)
        masks = [m != 0 for m in mo]
        kept = sum(int(m.sum()) for m in masks)
        for i, n in enumerate(sizes):
            if not entire:
                assert int(masks[i].sum()) == cr.randomk_keep_count(n, 0.02)
            if ef and step == 1:
                acc = [g0[r][i] + g0[r][i] * (~prev_masks[i]) for r in range(W)]     # residual of step 0 + new gradient
            else:
                acc = [g0[r][i] for r in range(W)]
            exp = sum(a * masks[i] for a in acc) / W
            for r in range(W):
                assert torch.allclose(outs[r][i], exp, atol=1e-6, rtol=1e-6)
                assert torch.equal(outs[r][i], outs[0][i])
            if ef:
                for r in range(W):
                    assert torch.allclose(resids[r][i], acc[r] * (~masks[i]), atol=1e-6)
        prev_masks = masks
        st = plans[0].stats()
        assert st["dropped"] == 0 and st["bytes"] == st["sent"] * 4
        del keep


def test_per_rank_seed_randomk_draws_different_masks_and_ships_indices():
    cfg = CompressionConfig(compress="layerwise", method="Randomk", ratio=0.05, shared_seed=False)
    grads, outs, _, plans, lw = _world(2, [50000], cfg, False)
    assert not any(p.values_only for p in plans)
    for r in range(2):
        grads[r][0].fill_(1.0)
    _run(plans, 0)
    nz = int((outs[0][0] != 0).sum())
    k = cr.randomk_keep_count(50000, 0.05)
    assert k < nz <= 2 * k                      # two independent masks: union larger than one mask
    assert torch.equal(outs[0][0], outs[1][0])
    assert plans[0].stats()["bytes"] == plans[0].stats()["sent"] * 8


def test_peer_broadcast_virtual_ranks():
    """lwc_peer_broadcast with W = 4 virtual ranks on one GPU: the root's tensors (mixed dtypes, odd sizes) arrive in every
    other rank's tensors, twice in a row (both staging parities)."""
    from layerwise_b200.parallel.peer import LocalWorld, PeerBroadcaster

    class _Heap:
        def __init__(self, ptrs, rank):
            self.ptrs, self.world, self.rank = ptrs, len(ptrs), rank

        def close(self):
            pass

    W = 4
    shapes = [((64,), torch.float32), ((3, 5, 7), torch.float32), ((1,), torch.int64), ((1000, 33), torch.bfloat16), ((17,), torch.float16)]
    ranks = [[(torch.randn(s, device="cuda") * (r + 1)).to(dt) if dt.is_floating_point else torch.full(s, r + 7, device="cuda", dtype=dt)
              for s, dt in shapes] for r in range(W)]
    nbytes = sum((t.numel() * t.element_size() + 15) // 16 * 16 for t in ranks[0])
    cap = (nbytes + 255) // 256 * 256
    lw = LocalWorld(W, 2 * cap + 512)
    bcs = [PeerBroadcaster(ranks[r], heap=_Heap(lw.ptrs, r), root=1) for r in range(W)]
    for rep in range(2):
        want = [t.clone() for t in ranks[1]]
        bcs[1].broadcast()                       # the root first: nothing waits on an unlaunched kernel
        for r in (0, 2, 3):
            bcs[r].broadcast()
        torch.cuda.synchronize()
        for r in range(W):
            for a, b in zip(ranks[r], want):
                assert torch.equal(a, b)
        for t in ranks[1]:
            if t.dtype.is_floating_point:
                t.mul_(-0.5)
            else:
                t.add_(1)
    lw.close()
