"""Fused CUDA exchange engine (B200): plans, buckets, streams, hooks.

``ExchangePlan`` wraps one native plan (``csrc/exchange.cu``): a bucket of gradient tensors
whose compress -> pack -> peer-store -> flag -> receive -> scatter-add/dequantise pipeline
runs as a fixed, short launch sequence with no host synchronisation.

``FusedEngine`` owns the gradient arena (every ``p.grad`` is a view into one flat buffer
per dtype, so "flattening" the model gradient costs nothing -- the reference concatenates
and re-slices every step, CIFAR10/core.py:229-236,:292-301), the fp32 error-feedback
arena, the peer heap with receive slots, and the bucket schedule.  With ``overlap=True`` it
registers post-accumulate-grad hooks and launches each bucket's send on a side stream the
moment its last gradient is produced, in reverse parameter order -- the role of the
reference's hook-driven DDP wrappers (IMAGENET/training/sparsified_ddp.py:383-450) for
every compressor, not only Random-k.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from .. import _native
from .._native import abi
from ..ops import compress_ref as cref
from . import comm
from .peer import DeviceBuffer, LocalWorld, PeerHeap

FLAG_BYTES = abi.LWC_MAX_WORLD * 8


def _ptr_array(ptrs: Sequence[int]):
    arr = (C.c_void_p * len(ptrs))()
    for i, p in enumerate(ptrs):
        arr[i] = p
    return arr


def keep_and_cap(method: str, n: int, cfg) -> (int, int):
    """Per-group keep count and payload capacity (entries).

    With error feedback a payload that outgrows its slot is clamped and the excess stays in the residual -- nothing
    is lost, so the slots are sized for the expected count plus headroom.  WITHOUT error feedback (the reference's
    menu semantics: every ``|g| >= V`` and every tie at the k-th value is sent, CIFAR10/core.py:178-193) a clamped
    entry would simply vanish, so the slot holds the whole tensor: ties and loose thresholds can never drop a
    gradient (the price is slot memory, n * 8 bytes per tensor and rank)."""
    lossless = not getattr(cfg, "error_feedback", False)
    if method == cref.TOPK:
        keep = cref.topk_keep_count(n, cfg.ratio)
        cap = n if lossless else min(n, int(math.ceil(keep * cfg.capacity_factor)) + 64)
    elif method == cref.RANDOMK:
        keep = cref.randomk_keep_count(n, cfg.ratio)
        cap = keep
    elif method in (cref.THRESHOLDV, cref.ADAPTIVE):
        keep = 0
        cap = n if lossless else min(n, max(int(math.ceil(n * cfg.dense_capacity)), 1024))
    else:
        keep, cap = 0, 0
    return max(keep, 0), max(cap, 1)


class ExchangePlan:
    """One native plan.  ``grads/outs/resids`` are flat tensors with stable addresses."""

    def __init__(self, grads: List[torch.Tensor], outs: List[torch.Tensor],
                 resids: Optional[List[Optional[torch.Tensor]]], cfg, world: int, rank: int,
                 entire: bool, uids: Optional[List[int]] = None):
        self.lib = _native.lib()
        self.cfg, self.world, self.rank = cfg, world, rank
        self.method = cfg.method_name
        T = len(grads)
        self.keepalive = (grads, outs, resids)
        numel = [int(g.numel()) for g in grads]
        if entire:
            group_of = [0] * T
            gsize = [sum(numel)]
        else:
            group_of = list(range(T))
            gsize = numel
        kc = [keep_and_cap(self.method, n, cfg) for n in gsize]
        self.keep = [k for k, _ in kc]
        self.cap = [c for _, c in kc]
        d = abi.LwcPlanDesc()
        d.num_tensors, d.num_groups = T, len(gsize)
        d.method = abi.METHOD_CODES[self.method]
        d.grad_dtype = abi.dtype_code(grads[0].dtype)
        d.out_dtype = abi.dtype_code(outs[0].dtype)
        d.world, d.rank = world, rank
        d.qstates = int(cfg.qstates or 0) if self.method == cref.DITHERING else 0
        d.threshold = float(cfg.threshold or 0.0)
        shared = bool(getattr(cfg, "shared_seed", True))
        # per-rank masks (the reference's CIFAR Random-k draws an independent randperm on every rank, core.py:186):
        # mix the rank into the key; shared masks (sparsified DDP) keep one seed for all ranks
        seed = int(cfg.seed) if shared else int(cfg.seed) + 0x9E3779B97F4A7C15 * (rank + 1)
        d.seed = seed & 0xFFFFFFFFFFFFFFFF
        self._arrs = dict(
            grad=_ptr_array([g.data_ptr() for g in grads]),
            out=_ptr_array([o.data_ptr() for o in outs]),
            resid=_ptr_array([0 if (resids is None or r is None) else r.data_ptr()
                              for r in (resids or [None] * T)]),
            numel=(C.c_int64 * T)(*numel),
            group_of=(C.c_int32 * T)(*group_of),
            uid=(C.c_int32 * T)(*(uids or list(range(T)))),
            keep=(C.c_int64 * len(gsize))(*self.keep),
            cap=(C.c_int64 * len(gsize))(*self.cap),
        )
        a = self._arrs
        d.grad_dev, d.out_dev, d.resid_dev = a["grad"], a["out"], a["resid"]
        d.numel, d.group_of, d.uid, d.keep, d.cap = a["numel"], a["group_of"], a["uid"], a["keep"], a["cap"]
        self.handle = C.c_void_p()
        abi.check(self.lib.lwc_plan_create(C.byref(self.handle), C.byref(d)), "lwc_plan_create")
        # shared-seed Random-k: values-only wire format (indices are implied by the hash, sparsified_ddp.py:410-412)
        self.values_only = self.method == cref.RANDOMK and shared
        if self.values_only:
            abi.check(self.lib.lwc_plan_set_values_only(self.handle, 1), "lwc_plan_set_values_only")
        self.slot_bytes = int(self.lib.lwc_plan_slot_bytes(self.handle))
        self.region_bytes = int(self.lib.lwc_plan_region_bytes(self.handle))
        self.numel = sum(numel)
        # kernels per send (the memset is not one of ours) and per receive
        self.kernels_per_send = {cref.TOPK: 5, cref.RANDOMK: 5, cref.THRESHOLDV: 2 + int(resids is not None),
                                 cref.ADAPTIVE: 3, cref.TERNGRAD: 2, cref.DITHERING: 2}.get(self.method, 1)

    def bind(self, region_ptrs: Sequence[int], flag_ptrs: Sequence[int]):
        abi.check(self.lib.lwc_plan_bind(self.handle, _ptr_array(region_ptrs), _ptr_array(flag_ptrs)),
                  "lwc_plan_bind")

    def bind_multicast(self, mc_region_ptr: int):
        """NVLS variant: payload stores go once to the multicast mapping of the region (0 switches back)."""
        abi.check(self.lib.lwc_plan_bind_multicast(self.handle, C.c_void_p(mc_region_ptr or None)), "lwc_plan_bind_multicast")

    def set_random(self, rand: Optional[List[Optional[torch.Tensor]]]):
        self._rand_keep = rand
        if rand is None:
            abi.check(self.lib.lwc_plan_set_random(self.handle, None), "lwc_plan_set_random")
        else:
            arr = _ptr_array([0 if r is None else r.data_ptr() for r in rand])
            abi.check(self.lib.lwc_plan_set_random(self.handle, arr), "lwc_plan_set_random")

    def update_grads(self, ptrs: Sequence[int], stream=None):
        """Re-point the input gradients (stream-ordered table upload)."""
        st = (stream or torch.cuda.current_stream()).cuda_stream
        abi.check(self.lib.lwc_plan_update_grads(self.handle, _ptr_array(ptrs), C.c_void_p(st)), "lwc_plan_update_grads")

    def use_device_step(self, enable: bool, start_step: int):
        """Keep the step counter on the device (kernels read it, the receiver bumps it) so that a captured
        send/recv sequence can be replayed by a CUDA graph.  The caller synchronises the device first."""
        abi.check(self.lib.lwc_plan_use_device_step(self.handle, int(bool(enable)), C.c_uint64(start_step)),
                  "lwc_plan_use_device_step")

    def send(self, step: int, stream=None):
        st = (stream or torch.cuda.current_stream()).cuda_stream
        abi.check(self.lib.lwc_plan_send(self.handle, C.c_uint64(step), C.c_void_p(st)), "lwc_plan_send")
        _native.counters.add(self.kernels_per_send)

    def recv(self, step: int, stream=None):
        st = (stream or torch.cuda.current_stream()).cuda_stream
        abi.check(self.lib.lwc_plan_recv(self.handle, C.c_uint64(step), C.c_void_p(st)), "lwc_plan_recv")
        _native.counters.add(1)

    def stats(self) -> Dict[str, int]:
        out = (C.c_uint64 * 4)()
        st = torch.cuda.current_stream().cuda_stream
        abi.check(self.lib.lwc_plan_stats(self.handle, out, C.c_void_p(st)), "lwc_plan_stats")
        return {"sent": int(out[0]), "dropped": int(out[1]), "timeouts": int(out[2]), "bytes": int(out[3])}

    def close(self):
        if self.handle:
            self.lib.lwc_plan_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def bind_local_world(plans: List[ExchangePlan]) -> LocalWorld:
    """Wire W plans (virtual ranks of one process) to each other; returns the backing heaps."""
    W = len(plans)
    region = max(p.region_bytes for p in plans)
    region = (region + 255) // 256 * 256
    lw = LocalWorld(W, region + FLAG_BYTES)
    regions = [ptr for ptr in lw.ptrs]
    flags = [ptr + region for ptr in lw.ptrs]
    for p in plans:
        p.bind(regions, flags)
    return lw


# --------------------------------------------------------------------------------------------------------
class _Bucket:
    def __init__(self, idxs: List[int]):
        self.idxs = idxs
        self.plan: Optional[ExchangePlan] = None
        self.pending = 0
        self.launched = False


class FusedEngine:
    name = "fused"

    def __init__(self, params: List[torch.nn.Parameter], cfg, group=None, overlap: bool = False,
                 out_dtype: Optional[torch.dtype] = None, num_buckets: Optional[int] = None, steal: bool = True,
                 nvls: bool = False):
        _native.require()
        self.params, self.cfg, self.group = params, cfg, group
        # steal=True: ``zero_grad`` sets ``p.grad = None`` so autograd hands over its freshly written
        # gradient tensors (no accumulate/copy pass); the plans read them in place and write the averaged
        # gradient into the arena, to which ``p.grad`` is re-pointed after the exchange.
        self.steal = steal
        self.world = comm.world_size(group)
        self.rank = comm.rank(group)
        if self.world > abi.LWC_MAX_WORLD:
            raise ValueError(f"fused engine supports up to {abi.LWC_MAX_WORLD} ranks per node")
        self.device = params[0].device
        self.step = 0
        self.graph_mode = False
        self.overlap = overlap
        dts = {p.dtype for p in params}
        self.entire = cfg.compress == "entiremodel"
        # entire-model selection runs over one flat gradient; with mixed precision (bf16 weights, fp32 batch-norm
        # parameters) there is one flat gradient -- hence one selection group -- per precision class
        # ---- arenas: grads (per dtype) and fp32 error-feedback residuals ----------------------------
        self._grad_arena: Dict[torch.dtype, torch.Tensor] = {}
        self.grad_views: List[torch.Tensor] = [None] * len(params)
        for dt in dts:
            idx = [i for i, p in enumerate(params) if p.dtype == dt]
            offs, total = [], 0
            for i in idx:
                offs.append(total)
                total += (params[i].numel() + 63) // 64 * 64      # 256-byte aligned views (fp32)
            arena = torch.zeros(total, dtype=dt, device=self.device)
            self._grad_arena[dt] = arena
            for i, o in zip(idx, offs):
                self.grad_views[i] = arena[o:o + params[i].numel()]
        use_ef = cfg.error_feedback and cfg.method_name != cref.NONE
        self.resid_views: Optional[List[torch.Tensor]] = None
        if use_ef:
            total = sum((p.numel() + 63) // 64 * 64 for p in params)
            self._resid_arena = torch.zeros(total, dtype=torch.float32, device=self.device)
            self.resid_views, o = [], 0
            for p in params:
                self.resid_views.append(self._resid_arena[o:o + p.numel()])
                o += (p.numel() + 63) // 64 * 64
        self._adopt_grads()
        # ---- buckets -----------------------------------------------------------------------------------------
        order = list(range(len(params)))[::-1] if overlap else list(range(len(params)))
        if self.entire:
            buckets = [[i for i, p in enumerate(params) if p.dtype == dt] for dt in sorted(dts, key=str)]
        else:
            buckets = self._make_buckets(order, num_buckets)
        self.buckets = [_Bucket(b) for b in buckets]
        for b in self.buckets:
            b.idxs = sorted(b.idxs)
            grads = [self.grad_views[i] for i in b.idxs]
            resids = None if self.resid_views is None else [self.resid_views[i] for i in b.idxs]
            b.plan = ExchangePlan(grads, grads, resids, cfg, self.world, self.rank, self.entire,
                                  uids=list(b.idxs))
        # ---- peer heap: regions of every bucket + flag words ---------------------------------------------
        offs, total = [], 0
        for b in self.buckets:
            offs.append(total)
            total += (b.plan.region_bytes + 255) // 256 * 256
        flag_off = total
        total += FLAG_BYTES * len(self.buckets)
        # nvls=True: allocate the heap through the symmetric-memory allocator and, if the host exposes NVSwitch
        # multicast, send every payload with ONE multimem store instead of one unicast store per peer
        self.heap = PeerHeap(total, group=group, device=self.device, multicast=nvls)
        self.nvls = bool(nvls and self.heap.mc_ptr)
        for bi, b in enumerate(self.buckets):
            b.plan.bind([ptr + offs[bi] for ptr in self.heap.ptrs],
                        [ptr + flag_off + bi * FLAG_BYTES for ptr in self.heap.ptrs])
            if self.nvls:
                b.plan.bind_multicast(self.heap.mc_ptr + offs[bi])
        self.comm_stream = torch.cuda.Stream(device=self.device) if overlap else None
        self._hooks = []
        self._param_bucket = {}
        self._shaped_views = [self._shaped(i) for i in range(len(params))]
        self._arena_ptr = [v.data_ptr() for v in self.grad_views]
        self._cur_ptr: List[Optional[int]] = list(self._arena_ptr)
        for b in self.buckets:
            b.ptrs = [self._arena_ptr[i] for i in b.idxs]
        if overlap:
            for bi, b in enumerate(self.buckets):
                for i in b.idxs:
                    self._param_bucket[i] = bi
            for i, p in enumerate(params):
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
            self._reset_pending()

    # ---- helpers ---------------------------------------------------------------------------------------------
    def _make_buckets(self, order: List[int], num_buckets: Optional[int]) -> List[List[int]]:
        by_dtype: Dict[torch.dtype, List[int]] = {}
        for i in order:
            by_dtype.setdefault(self.params[i].dtype, []).append(i)
        out = []
        for dt, idx in by_dtype.items():
            if not self.overlap and num_buckets is None:
                out.append(idx)
                continue
            total_bytes = sum(self.params[i].numel() * self.params[i].element_size() for i in idx)
            cap = self.cfg.bucket_cap_mb * (1 << 20) if num_buckets is None else total_bytes / num_buckets
            cur, size = [], 0
            for i in idx:
                nbytes = self.params[i].numel() * self.params[i].element_size()
                if cur and size + nbytes > cap:
                    out.append(cur)
                    cur, size = [], 0
                cur.append(i)
                size += nbytes
            if cur:
                out.append(cur)
        return out

    def _shaped(self, i: int) -> torch.Tensor:
        """Arena view with the parameter's own shape *and strides* (e.g. channels_last weights),
        so that flat element k of the gradient matches flat element k of the parameter."""
        p, v = self.params[i], self.grad_views[i]
        dense = p.is_contiguous() or (p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last))
        if not dense:
            raise ValueError("fused engine needs dense parameters (contiguous or channels_last)")
        return torch.as_strided(v, p.size(), p.stride())

    def _adopt_grads(self):
        """Make every ``p.grad`` a view of the arena (copying an existing gradient once)."""
        for i, (p, v) in enumerate(zip(self.params, self.grad_views)):
            g = p.grad
            if g is not None and g.data_ptr() == v.data_ptr():
                continue
            shaped = self._shaped(i)
            if g is not None:
                shaped.copy_(g)
            p.grad = shaped

    def zero_grad(self):
        """steal mode: drop the gradients (``None``) so the next backward writes fresh tensors that are
        consumed in place; otherwise one memset per dtype arena instead of one launch per parameter."""
        if self.steal:
            for i, p in enumerate(self.params):
                p.grad = None
                self._cur_ptr[i] = None
            return
        for arena in self._grad_arena.values():
            arena.zero_()

    def _note_grad(self, i: int, param):
        """Record where this step's gradient of parameter i lives (or move it into the arena)."""
        g = param.grad
        if g is None:
            return
        if g.data_ptr() == self._arena_ptr[i]:
            self._cur_ptr[i] = self._arena_ptr[i]
        elif self.steal and g.dtype == param.dtype and g.stride() == param.stride() and g.data_ptr() % 16 == 0:
            self._cur_ptr[i] = g.data_ptr()
        else:
            self._shaped_views[i].copy_(g)
            param.grad = self._shaped_views[i]
            self._cur_ptr[i] = self._arena_ptr[i]

    def _send(self, b, stream=None):
        ptrs = []
        for i in b.idxs:
            ptr = self._cur_ptr[i]
            if ptr is None:                      # parameter produced no gradient this step
                self.grad_views[i].zero_()
                ptr = self._arena_ptr[i]
            ptrs.append(ptr)
        if ptrs != b.ptrs:
            b.plan.update_grads(ptrs, stream)
            b.ptrs = ptrs
        b.plan.send(self.step, stream)

    def enable_graph_mode(self):
        """Make send/recv replayable by a CUDA graph: from here on the step counter of every plan lives on the
        device (``self.step`` stays a host mirror, advanced by :meth:`note_replay`).  Collective across ranks in
        the sense that every rank must switch at the same step.  One live graph per engine: a later capture
        reuses each plan's pinned pointer-table staging."""
        if self.graph_mode:
            return
        torch.cuda.synchronize(self.device)
        for b in self.buckets:
            b.plan.use_device_step(True, self.step)
        self.graph_mode = True

    def note_replay(self, n: int = 1):
        """A captured step was replayed ``n`` times: advance the host mirror of the step counter."""
        self.step += n
        for b in self.buckets:
            b.ptrs = None            # the replay re-uploaded ITS pointer table: the next eager step must upload again

    def _repoint(self):
        """After the receivers are queued: every ``p.grad`` is the arena view holding the average."""
        for p, v in zip(self.params, self._shaped_views):
            p.grad = v

    # ---- after-backward mode -----------------------------------------------------------------------------
    @torch.no_grad()
    def exchange(self):
        if self.overlap:
            return self.finish()
        for i, p in enumerate(self.params):
            self._cur_ptr[i] = None
            self._note_grad(i, p)
        for b in self.buckets:
            self._send(b)
        for b in self.buckets:
            b.plan.recv(self.step)
        self._repoint()
        self.step += 1

    # ---- hook-driven mode ----------------------------------------------------------------------------------
    def _reset_pending(self):
        for b in self.buckets:
            b.pending = len(b.idxs)
            b.launched = False

    def _make_hook(self, i: int):
        def hook(param):
            b = self.buckets[self._param_bucket[i]]
            self._note_grad(i, param)
            b.pending -= 1
            if b.pending == 0 and not b.launched:
                b.launched = True
                self.comm_stream.wait_stream(torch.cuda.current_stream())
                self._send(b, self.comm_stream)
        return hook

    @torch.no_grad()
    def finish(self):
        """Join: launch any bucket whose hooks did not all fire, run the receivers."""
        if not self.overlap:
            return
        cur = torch.cuda.current_stream()
        self.comm_stream.wait_stream(cur)
        for b in self.buckets:
            if not b.launched:
                for i in b.idxs:                 # gradients that arrived without completing the bucket
                    if self._cur_ptr[i] is None:
                        self._note_grad(i, self.params[i])
                self._send(b, self.comm_stream)
        for b in self.buckets:
            b.plan.recv(self.step, self.comm_stream)
        cur.wait_stream(self.comm_stream)
        self._repoint()
        self.step += 1
        self._reset_pending()

    @torch.no_grad()
    def benchmark_exchange(self, iters: int = 10) -> float:
        """Device time (ms) of one full compress -> exchange -> decompress of this engine's buckets, run back
        to back on the current stream over the arena contents (no backward to hide behind).  Advances the step
        counter and the error-feedback state like real exchanges do."""
        for b in self.buckets:
            ptrs = [self._arena_ptr[i] for i in b.idxs]
            if ptrs != b.ptrs:
                b.plan.update_grads(ptrs)
                b.ptrs = ptrs

        def once():
            for b in self.buckets:
                b.plan.send(self.step)
            for b in self.buckets:
                b.plan.recv(self.step)
            self.step += 1
        for _ in range(2):
            once()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            once()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    # ---- state -------------------------------------------------------------------------------------------------
    def state_dict(self):
        res = None if self.resid_views is None else [r.detach().cpu().clone() for r in self.resid_views]
        return {"residual": res, "step": self.step}

    def load_state_dict(self, sd):
        # the step counter also keys the peer flags: never move it backwards on a live engine
        self.step = max(self.step, int(sd.get("step", 0)))
        if self.graph_mode:
            torch.cuda.synchronize(self.device)
            for b in self.buckets:
                b.plan.use_device_step(True, self.step)
        res = sd.get("residual")
        if res is not None and self.resid_views is not None:
            if isinstance(res, torch.Tensor):      # entire-model flat residual from the torch engine
                off = 0
                for v in self.resid_views:
                    v.copy_(res[off:off + v.numel()])
                    off += v.numel()
            else:
                for v, r in zip(self.resid_views, res):
                    v.copy_(r)

    def stats(self):
        tot = {"sent": 0, "dropped": 0, "timeouts": 0, "bytes": 0}
        for b in self.buckets:
            for k, v in b.plan.stats().items():
                tot[k] += v
        return tot

    def close(self):
        for h in self._hooks:
            h.remove()
        for b in self.buckets:
            b.plan.close()
        self.heap.close()


# --------------------------------------------------------------------------------------------------------
class DenseEngine:
    """Uncompressed gradient averaging (``--compress none`` and the dense DDP wrapper) as a two-shot
    all-reduce over peer memory (``lwc_dense_allreduce``): the gradient arenas themselves live in the
    peer-mapped heap, rank r reduces slice r of every arena and stores the average back everywhere.
    Buckets are contiguous arena ranges in reverse parameter order; with ``overlap=True`` each is
    launched from the hook of its last gradient on a side stream.
    Reference: per-tensor ``dist.all_reduce`` + ``/W`` (CIFAR10/core.py:316-319) and the bucketed hook-driven
    all-reduce of the vendored DDP (IMAGENET/training/ddp.py:394-475)."""

    name = "fused-dense"
    FLAGS = 2 * abi.LWC_MAX_WORLD * 8       # ready row + done row
    AUX = FLAGS + 64                        # + ticket / timeout counters

    def __init__(self, params: List[torch.nn.Parameter], cfg, group=None, overlap: bool = False, **_unused):
        _native.require()
        self.lib = _native.lib()
        self.params, self.cfg, self.group = params, cfg, group
        self.world, self.rank = comm.world_size(group), comm.rank(group)
        if self.world > abi.LWC_MAX_WORLD:
            raise ValueError(f"fused engine supports up to {abi.LWC_MAX_WORLD} ranks per node")
        self.device = params[0].device
        self.step = 0
        self.overlap = overlap
        # ---- arena layout inside the heap ------------------------------------------------------------------
        by_dtype: Dict[torch.dtype, List[int]] = {}
        for i, p in enumerate(params):
            by_dtype.setdefault(p.dtype, []).append(i)
        self._layout = {}
        heap_bytes = 0
        offs_in_arena = {}
        for dt, idx in by_dtype.items():
            esz = torch.empty((), dtype=dt).element_size()
            total = 0
            for i in idx:
                offs_in_arena[i] = total
                total += (params[i].numel() + 63) // 64 * 64
            self._layout[dt] = (heap_bytes, total, esz, idx)
            heap_bytes += (total * esz + 255) // 256 * 256
        # ---- buckets: contiguous ranges per dtype, reverse order --------------------------------------------
        cap = cfg.bucket_cap_mb * (1 << 20) if overlap else float("inf")
        self.buckets = []
        for dt, (base, total, esz, idx) in self._layout.items():
            cur: List[int] = []
            size = 0
            for i in reversed(idx):
                nbytes = params[i].numel() * esz
                if cur and size + nbytes > cap:
                    self.buckets.append({"dtype": dt, "idxs": cur})
                    cur, size = [], 0
                cur.append(i)
                size += nbytes
            if cur:
                self.buckets.append({"dtype": dt, "idxs": cur})
        aux_base = heap_bytes
        heap_bytes += self.AUX * len(self.buckets)
        self.heap = PeerHeap(heap_bytes, group=group, device=self.device)
        self._grad_arena = {}
        self.grad_views: List[torch.Tensor] = [None] * len(params)
        for dt, (base, total, esz, idx) in self._layout.items():
            arena = self.heap.local.view(dt, total, base)
            self._grad_arena[dt] = arena
            for i in idx:
                self.grad_views[i] = arena[offs_in_arena[i]:offs_in_arena[i] + params[i].numel()]
        for bi, b in enumerate(self.buckets):
            base, total, esz, _ = self._layout[b["dtype"]]
            lo = min(offs_in_arena[i] for i in b["idxs"])
            hi = max(offs_in_arena[i] + (params[i].numel() + 63) // 64 * 64 for i in b["idxs"])
            b.update(off=lo, n=hi - lo, pending=0, launched=False,
                     arenas=_ptr_array([ptr + base for ptr in self.heap.ptrs]),
                     flags=_ptr_array([ptr + aux_base + bi * self.AUX for ptr in self.heap.ptrs]),
                     counters=self.heap.local.ptr + aux_base + bi * self.AUX + self.FLAGS)
        self._adopt_grads()
        self.comm_stream = torch.cuda.Stream(device=self.device) if overlap else None
        self._hooks = []
        if overlap:
            self._param_bucket = {i: bi for bi, b in enumerate(self.buckets) for i in b["idxs"]}
            for i, p in enumerate(params):
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
            self._reset_pending()

    _shaped = FusedEngine._shaped
    _adopt_grads = FusedEngine._adopt_grads

    def zero_grad(self):
        for arena in self._grad_arena.values():
            arena.zero_()

    def _launch(self, b, stream=None):
        if self.world == 1:
            return
        st = (stream or torch.cuda.current_stream()).cuda_stream
        abi.check(self.lib.lwc_dense_allreduce(b["arenas"], b["flags"], C.c_void_p(b["counters"]), b["off"], b["n"],
                                               abi.dtype_code(b["dtype"]), self.world, self.rank, C.c_uint64(self.step),
                                               C.c_void_p(st)), "lwc_dense_allreduce")
        _native.counters.add(2)

    @torch.no_grad()
    def exchange(self):
        if self.overlap:
            return self.finish()
        self._adopt_grads()
        for b in self.buckets:
            self._launch(b)
        self.step += 1

    def _reset_pending(self):
        for b in self.buckets:
            b["pending"] = len(b["idxs"])
            b["launched"] = False

    def _make_hook(self, i: int):
        def hook(param):
            b = self.buckets[self._param_bucket[i]]
            if param.grad is not None and param.grad.data_ptr() != self.grad_views[i].data_ptr():
                shaped = self._shaped(i)
                shaped.copy_(param.grad)
                param.grad = shaped
            b["pending"] -= 1
            if b["pending"] == 0 and not b["launched"]:
                b["launched"] = True
                self.comm_stream.wait_stream(torch.cuda.current_stream())
                self._launch(b, self.comm_stream)
        return hook

    @torch.no_grad()
    def finish(self):
        if not self.overlap:
            return
        cur = torch.cuda.current_stream()
        self.comm_stream.wait_stream(cur)
        for b in self.buckets:
            if not b["launched"]:
                self._launch(b, self.comm_stream)
        cur.wait_stream(self.comm_stream)
        self.step += 1
        self._reset_pending()

    def state_dict(self):
        return {"residual": None, "step": self.step}

    def load_state_dict(self, sd):
        self.step = max(self.step, int(sd.get("step", 0)))

    def stats(self):
        timeouts = 0
        torch.cuda.synchronize()
        aux = self.heap.local.view(torch.int64, (self.AUX * len(self.buckets)) // 8,
                                   self.heap.nbytes - self.AUX * len(self.buckets))
        per = self.AUX // 8
        for bi in range(len(self.buckets)):
            timeouts += int(aux[bi * per + self.FLAGS // 8 + 1].item())
        nbytes = sum(b["n"] * self._layout[b["dtype"]][2] for b in self.buckets)
        return {"sent": 0, "dropped": 0, "timeouts": timeouts, "bytes": int(2 * nbytes * (self.world - 1) / max(self.world, 1))}

    def close(self):
        for h in self._hooks:
            h.remove()
        self.heap.close()
