"""ImageNet ResNet-50 trainer (progressive resizing recipe) with compressed gradient exchange.

Capabilities follow the reference's per-rank trainer (IMAGENET/training/train_imagenet_nv.py):
``main`` :149-253 (process group, model, fp16 + fp32 masters, BN weight-decay split, SGD
with Nesterov momentum, resume, phases -> DataManager + Scheduler, epoch loop, checkpoints),
``train`` :388-478 (hot loop), ``validate`` / ``distributed_predict`` :481-542,
``DataManager`` :545-598, ``save_checkpoint`` :663-669, ``accuracy``/``correct`` :672-689.

B200 design: the step is ``H2D(uint8) -> normalise -> forward -> loss -> backward ->
fused compress/exchange/decompress (hooked, overlapped) -> fused SGD on fp32 masters that
also writes the bf16 model copy``.  Metrics stay on the device and are reduced once per
``print_freq`` steps (the reference all-reduces and ``.cpu()``s four floats every step,
:447-450).  Defects D3-D5, D12-D14 of SURVEY 2.6 are fixed: ``--phases`` is honoured,
``entiremodel`` selects the entire-model exchange, the wrapper and the explicit exchange
are exclusive, master weights are checkpointed.
"""
from __future__ import annotations

import copy
import os
import time
from datetime import datetime
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from ..models import resnet as resnet_zoo
from ..ops import nn as lnn
from ..ops.sgd import FusedSGD
from ..parallel import comm
from ..parallel.exchange import CompressionConfig, GradientExchanger
from ..utils import mixed
from ..utils.checkpoint import load_checkpoint, save_checkpoint
from ..utils.logging import FileLogger, TensorboardLogger
from ..utils.timing import maybe_phase
from ..utils.meter import AverageMeter, NetworkMeter, TimeMeter
from ..utils.schedule import PhaseScheduler

IMAGENET_MEAN = (0.485 * 255, 0.456 * 255, 0.406 * 255)
IMAGENET_STD = (0.229 * 255, 0.224 * 255, 0.225 * 255)


def default_phases(lr: float = 1.0) -> List[dict]:
    """The single-machine schedule hard-coded in the reference (train_imagenet_nv.py:204-219)."""
    s224, s288 = 224 / 512, 128 / 512
    return [
        {"ep": 0, "sz": 128, "bs": 512, "trndir": "-sz/160"},
        {"ep": (0, 5), "lr": (lr, lr * 2)},
        {"ep": 5, "lr": lr},
        {"ep": 14, "sz": 224, "bs": 224, "lr": lr * s224},
        {"ep": 16, "lr": lr / 10 * s224},
        {"ep": 27, "lr": lr / 100 * s224},
        {"ep": 32, "sz": 288, "bs": 128, "min_scale": 0.5, "rect_val": True, "lr": lr / 100 * s288},
        {"ep": (33, 35), "lr": lr / 1000 * s288},
    ]


def correct(output: torch.Tensor, target: torch.Tensor, topk=(1,)):
    """Number of samples whose label is within the top-k predictions, one tensor per k."""
    if output.is_cuda and set(topk) <= {1, 5} and output.shape[1] >= 5:
        from ..ops import misc as _misc
        from ..ops.nn import get_backend
        if get_backend() != "torch":
            hits = _misc.topk_hits(output.detach().contiguous(), target)       # one kernel, device counters
            return [hits[0:1] if k == 1 else hits[1:2] for k in topk]
    maxk = max(topk)
    pred = output.topk(maxk, 1, True, True).indices.t()
    hit = pred.eq(target.view(1, -1).expand_as(pred))
    return [hit[:k].reshape(-1).float().sum(0, keepdim=True) for k in topk]


def accuracy(output, target, topk=(1,)):
    n = target.size(0)
    return [c.mul_(100.0 / n) for c in correct(output, target, topk)]


class InputNormalizer(nn.Module):
    """uint8 NCHW batch -> normalised compute dtype (reference dataloader.py:88-93)."""

    def __init__(self, dtype=torch.float32, channels_last=True, s2d=False):
        super().__init__()
        self.dtype, self.channels_last, self.s2d = dtype, channels_last, s2d
        self.register_buffer("mean", torch.tensor(IMAGENET_MEAN).view(1, 3, 1, 1), persistent=False)
        self.register_buffer("std", torch.tensor(IMAGENET_STD).view(1, 3, 1, 1), persistent=False)

    def forward(self, x):
        if self.s2d and x.is_cuda and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0:
            # one fused kernel: normalise + channels-last + 2x2 space-to-depth for the ResNet stem
            from ..ops.input import normalize_u8
            return normalize_u8(x.contiguous(), self.dtype, s2d=True, keep_plain=self.training)
        x = x.float().sub_(self.mean).div_(self.std).to(self.dtype)
        return x.contiguous(memory_format=torch.channels_last) if self.channels_last else x


class Trainer:
    """One rank's training state: model, precision policy, exchanger, optimiser, scheduler."""

    def __init__(self, arch: str = "resnet50", num_classes: int = 1000, dtype: str = "bf16",
                 cfg: Optional[CompressionConfig] = None, momentum: float = 0.9,
                 weight_decay: float = 1e-4, init_bn0: bool = True, no_bn_wd: bool = True,
                 loss_scale: float = 1.0, lr: float = 0.0, device=None, overlap: bool = True,
                 engine: str = "auto", channels_last: bool = True, model: Optional[nn.Module] = None,
                 wrapper: Optional[str] = None, randk: float = 1.0, seed: int = 2147483647,
                 cuda_graph: bool = False):
        self.device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
        self.phase_timer = None           # a utils.timing.PhaseTimer while profile_phases() runs
        if self.device.type == "cuda":
            torch.backends.cudnn.benchmark = True     # as the reference does (train_imagenet_nv.py:95): autotune conv algos
            if os.environ.get("LWC_CUDNN_BENCH_LIMIT") is not None:      # 0 = try every algorithm (default: the first 10)
                torch.backends.cudnn.benchmark_limit = int(os.environ["LWC_CUDNN_BENCH_LIMIT"])
        self.dtype = {"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}[dtype]
        self.loss_scale = loss_scale if self.dtype == torch.float16 else 1.0
        torch.manual_seed(seed)          # identical initial replicas (reference :117-118)
        net = model if model is not None else getattr(resnet_zoo, arch)(num_classes=num_classes, bn0=init_bn0)
        net = net.to(self.device)
        if channels_last and self.device.type == "cuda":
            net = net.to(memory_format=torch.channels_last)
        if self.dtype != torch.float32:
            net = mixed.bn_to_float(net.to(self.dtype))
        self.net = net
        on_gpu = self.device.type == "cuda"
        s2d = on_gpu and channels_last and isinstance(net, resnet_zoo.ResNet)
        self.normalize = InputNormalizer(self.dtype, channels_last and on_gpu, s2d=s2d).to(self.device)
        self.criterion = lnn.CrossEntropyLoss().to(self.device)
        self.criterion.trust_targets = True       # labels come from the dataset / sampler: always class indices
        self.cfg = cfg or CompressionConfig()
        self.model_params = [p for p in net.parameters() if p.requires_grad]
        if comm.is_initialized() and comm.world_size() > 1:
            comm.broadcast_coalesced([t for t in net.state_dict().values() if torch.is_tensor(t)])
        # ---- fp32 masters ----------------------------------------------------------------------------------
        if self.dtype != torch.float32:
            from ..ops.sgd import allow_lowp_grad
            self.master_params = [allow_lowp_grad(p.detach().clone().float().requires_grad_(True))
                                  if p.dtype != torch.float32 else p for p in self.model_params]
        else:
            self.master_params = self.model_params
        # ---- gradient exchange -----------------------------------------------------------------------------
        self.wrapper = None
        if wrapper in ("ddp", "sparsified"):
            from ..parallel import ddp as _ddp
            if wrapper == "ddp":
                self.wrapper = _ddp.DistributedDataParallel(net)
            else:
                self.wrapper = _ddp.RandomKSparsifiedDDP(net, randk=randk, seed=seed)
            self.exchanger = self.wrapper.exchanger
        else:
            self.exchanger = GradientExchanger(self.model_params, self.cfg, engine=engine,
                                               overlap=overlap and self.device.type == "cuda")
        # ---- optimiser: masters hold .grad views of the model grads (same arena, no copy) -------
        groups = (mixed.bnwd_optim_params(net, self.model_params, self.master_params) if no_bn_wd
                  else [{"params": self.master_params}])
        self.optimizer = FusedSGD(groups, lr=lr, momentum=momentum, weight_decay=weight_decay,
                                  nesterov=momentum > 0, grad_scale=1.0 / self.loss_scale,
                                  model_params=self._model_in_group_order(groups))
        self.steps = 0
        self._graph = None
        # cuda_graph=True: after 3 eager steps with one batch shape, capture the step for that shape and replay it
        self.cuda_graph = bool(cuda_graph) and self.device.type == "cuda" and self.wrapper is None \
            and hasattr(self.exchanger.engine, "enable_graph_mode")
        self._shape_key, self._shape_run = None, 0
        self._side = None
        self._graphable = self.device.type == "cuda" and self.wrapper is None \
            and hasattr(self.exchanger.engine, "enable_graph_mode")

    def _model_in_group_order(self, groups):
        by_master = {id(m): p for m, p in zip(self.master_params, self.model_params)}
        return [by_master[id(m)] for g in groups for m in g["params"]]

    # ---- one optimisation step ---------------------------------------------------------------------------
    def forward_backward(self, x: torch.Tensor, target: torch.Tensor):
        with maybe_phase(self.phase_timer, "forward"):
            out = (self.wrapper if self.wrapper is not None else self.net)(x)
            loss = self.criterion(out, target)
        # the backward hooks launch compress + send of every finished bucket on the comm stream: that work overlaps
        # this phase and is not part of "exchange (exposed)" below
        with maybe_phase(self.phase_timer, "backward (+hooked compress/send)"):
            (loss * self.loss_scale if self.loss_scale != 1.0 else loss).backward()
        return out, loss

    # ---- whole-step CUDA graph (opt-in) ---------------------------------------------------------------------
    def capture_graph(self, images: torch.Tensor, target: torch.Tensor, warmup: int = 3):
        """Capture ``normalise -> forward -> backward -> compress/exchange -> SGD`` for this batch shape into ONE
        CUDA graph; later ``train_step`` calls with the same shape copy the batch into static buffers and replay
        it (one launch instead of ~1.5k).  Needs the fused engine: its step counter and the learning rate move to
        device memory so the replay sees them change.  ``warmup`` eager steps run first (cuDNN autotuning,
        momentum buffers, lazily created plans); they are real optimisation steps on the given batch (pass 0 when
        eager steps of this shape have already run -- capturing itself executes nothing).  In a
        multi-rank job every rank must capture at the same step.  Returned logits / loss are static tensors that
        the next replay overwrites."""
        eng = self.exchanger.engine
        if self.device.type != "cuda" or self.wrapper is not None or not hasattr(eng, "enable_graph_mode"):
            raise RuntimeError("capture_graph needs a CUDA device and the fused exchange engine")
        from .. import _native
        self._graph = None
        self._g_images = images.to(self.device, copy=True)
        self._g_target = target.to(self.device, copy=True)
        eng.enable_graph_mode()
        self.optimizer.enable_device_lr(self.device)
        self.net.train()
        self.normalize.train()
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device)
        side = self._side
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._step_body(self._g_images, self._g_target)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        n0, step0 = _native.counters.kernels, eng.step
        with torch.cuda.graph(graph, capture_error_mode="thread_local"):
            self._g_out = self._step_body(self._g_images, self._g_target)
        self._graph_launches = _native.counters.kernels - n0
        _native.counters.add(-self._graph_launches)      # nothing ran during capture
        eng.step = step0
        self.steps -= 1
        self._graph = graph
        return graph

    def _replay(self, images, target):
        self._g_images.copy_(images, non_blocking=True)
        self._g_target.copy_(target, non_blocking=True)
        self.optimizer.sync_device_lr()
        self._graph.replay()
        from .. import _native
        _native.counters.add(self._graph_launches)
        self.exchanger.engine.note_replay()
        self.steps += 1
        return self._g_out

    def _step_body(self, images, target):
        pt = self.phase_timer
        with maybe_phase(pt, "normalize"):
            x = self.normalize(images) if images.dtype == torch.uint8 else images
        out, loss = self.forward_backward(x, target)
        with maybe_phase(pt, "exchange (exposed: join + receive)"):
            self.exchanger.exchange()
        with maybe_phase(pt, "optimizer"):
            self._attach_master_grads()
            self.optimizer.step()
            self.zero_grad()
        self.steps += 1
        return out, loss

    def profile_phases(self, images, target, steps: int = 5) -> str:
        """Run ``steps`` eager training steps with CUDA-event timing of every phase (also NVTX ranges) and return the
        per-phase table, max over ranks (the reference has only wall-clock meters, meter.py:49-60)."""
        from ..utils.timing import PhaseTimer
        graph, self._graph = self._graph, None
        cg, self.cuda_graph = self.cuda_graph, False
        self.phase_timer = PhaseTimer()
        try:
            for _ in range(steps):
                self.train_step(images, target)
            return self.phase_timer.table(steps)
        finally:
            self.phase_timer = None
            self._graph, self.cuda_graph = graph, cg

    def train_step(self, images: torch.Tensor, target: torch.Tensor):
        """``images``: uint8 NCHW (host pinned or device) or already-normalised float; returns (logits, loss)."""
        self.net.train()
        self.normalize.train()
        if (self._graph is not None and images.shape == self._g_images.shape
                and images.dtype == self._g_images.dtype):
            return self._replay(images, target)
        if self.cuda_graph:
            key = (tuple(images.shape), images.dtype)
            self._shape_run = self._shape_run + 1 if key == self._shape_key else 1
            self._shape_key = key
            if self._shape_run > 3:
                self.capture_graph(images, target, warmup=0)
                return self._replay(images, target)
        if self._graphable:
            # Eager steps of a trainer that may be captured later run on a stream of its own: autograd state created on
            # the legacy default stream (gradient accumulators kept alive by a loss the caller still holds) makes a
            # later capture fail with "legacy stream depend on a capturing blocking stream".
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.device)
            cur = torch.cuda.current_stream()
            self._side.wait_stream(cur)
            with torch.cuda.stream(self._side):
                if images.device != self.device:
                    images = images.to(self.device, non_blocking=True)
                    target = target.to(self.device, non_blocking=True)
                res = self._step_body(images, target)
            cur.wait_stream(self._side)
            return res
        if images.device != self.device:
            images = images.to(self.device, non_blocking=True)
            target = target.to(self.device, non_blocking=True)
        x = self.normalize(images) if images.dtype == torch.uint8 else images
        out, loss = self.forward_backward(x, target)
        if self.wrapper is not None:
            self.wrapper.finish_reduction()      # wrapper XOR explicit exchange
        else:
            self.exchanger.exchange()
        self._attach_master_grads()
        self.optimizer.step()
        self.zero_grad()
        self.steps += 1
        return out, loss

    def _attach_master_grads(self):
        for m, p in zip(self.master_params, self.model_params):
            if m is not p:
                m.grad = p.grad

    def zero_grad(self):
        eng = self.exchanger.engine
        if hasattr(eng, "zero_grad"):
            eng.zero_grad()
        else:
            for p in self.model_params:
                if p.grad is not None:
                    p.grad.zero_()

    @torch.no_grad()
    def predict(self, images, target):
        self.net.eval()
        self.normalize.eval()
        if images.device != self.device:
            images = images.to(self.device, non_blocking=True)
            target = target.to(self.device, non_blocking=True)
        x = self.normalize(images) if images.dtype == torch.uint8 else images
        out = self.net(x)
        return out, self.criterion(out, target)

    # ---- checkpoints (reference format + optional extras) ---------------------------------------
    def save(self, epoch, best_top5, filename, is_best=False, logdir=None):
        masters = [m for m, p in zip(self.master_params, self.model_params) if m is not p]
        return save_checkpoint(epoch, self.net, best_top5, self.optimizer, is_best=is_best, filename=filename,
                               logdir=logdir, exchanger=self.exchanger, masters=masters or None)

    def save_rank_state(self, filename):
        """Call on EVERY rank next to ``save`` (which only rank 0 calls): writes this rank's error-feedback residual
        into a sidecar of the checkpoint (``utils.checkpoint.save_rank_residual``)."""
        from ..utils.checkpoint import save_rank_residual
        if comm.world_size() > 1:
            return save_rank_residual(filename, self.exchanger, comm.rank())
        return None

    def resume(self, path):
        masters = [m for m, p in zip(self.master_params, self.model_params) if m is not p]
        epoch, best = load_checkpoint(path, self.net, self.optimizer, map_location=self.device,
                                      exchanger=self.exchanger, masters=masters or None,
                                      rank=comm.rank(), world=comm.world_size())
        if not masters:
            return epoch, best
        ck = torch.load(path, map_location="cpu", weights_only=False)
        if ck.get("masters") is None:        # reference-written file: rebuild masters from the model
            for m, p in zip(self.master_params, self.model_params):
                if m is not p:
                    m.data.copy_(p.data)
        return epoch, best


# ---- distributed evaluation ----------------------------------------------------------------------------------
@torch.no_grad()
def predict_metrics(trainer: Trainer, images, target) -> torch.Tensor:
    """Device tensor [samples, batches, loss, top-1 hits, top-5 hits] of this rank's batch (zeros when the rank's last
    batch is empty); nothing is read back to the host."""
    bs = images.size(0)
    m = torch.zeros(5, device=trainer.device)
    if bs:
        out, loss = trainer.predict(images, target)
        c1, c5 = correct(out.float(), target.to(trainer.device), topk=(1, 5))
        m[0] = float(bs)
        m[1] = 1.0
        m[2] = loss.float().reshape(())
        m[3] = c1.reshape(())
        m[4] = c5.reshape(())
    return m


@torch.no_grad()
def distributed_predict(trainer: Trainer, images, target):
    """Evaluation that tolerates ranks with an empty last batch (reference :523-542); one reduction + host read per
    call -- ``validate`` uses ``predict_metrics`` and reads back once per ``print_freq`` batches instead."""
    m = comm.sum_tensor(predict_metrics(trainer, images, target))
    total, valid, loss_sum, c1, c5 = m.tolist()
    return c1 * 100.0 / max(total, 1), c5 * 100.0 / max(total, 1), loss_sum / max(valid, 1), total


def validate(trainer: Trainer, val_loader, epoch: int, log=None, tb=None, short_epoch=False, print_freq=50):
    losses, top1, top5 = AverageMeter(), AverageMeter(), AverageMeter()
    t0 = time.time()
    n = len(val_loader)
    # per-batch metrics stay on the device ([n_batches_in_window, 5]); ONE all-reduce + host read per print_freq
    # batches instead of one per batch (reference :536-537 synchronises every eval batch)
    window = []

    def flush():
        if not window:
            return
        rows = comm.sum_tensor(torch.stack(window)).tolist()
        window.clear()
        for total, valid, loss_sum, c1, c5 in rows:
            losses.update(loss_sum / max(valid, 1), total)
            top1.update(c1 * 100.0 / max(total, 1), total)
            top5.update(c5 * 100.0 / max(total, 1), total)

    for i, (images, target) in enumerate(val_loader):
        if short_epoch and i > 10:
            break
        window.append(predict_metrics(trainer, images, target))
        if (i + 1) % print_freq == 0 or i + 1 == n:
            flush()
        if log is not None and ((i + 1) % print_freq == 0 or i + 1 == n):
            log.verbose(f"Test:  [{epoch}][{i + 1}/{n}]\tLoss {losses.val:.4f} ({losses.avg:.4f})\t"
                        f"Acc@1 {top1.val:.3f} ({top1.avg:.3f})\tAcc@5 {top5.val:.3f} ({top5.avg:.3f})")
    flush()
    if tb is not None:
        tb.log_eval(top1.avg, top5.avg, time.time() - t0)
        tb.log("epoch", epoch)
    return top1.avg, top5.avg


def train_epoch(trainer: Trainer, loader, scheduler: PhaseScheduler, epoch: int, log=None, tb=None,
                short_epoch=False, print_freq=50):
    timer, net_meter = TimeMeter(), NetworkMeter()
    losses, top1, top5 = AverageMeter(), AverageMeter(), AverageMeter()
    n = len(loader)
    # device-side accumulators: [samples, loss*bs, correct1, correct5]
    acc = torch.zeros(4, device=trainer.device)
    for i, (images, target) in enumerate(loader):
        if short_epoch and i > 10:
            break
        timer.batch_start()
        scheduler.update_lr(epoch, i + 1, n)
        out, loss = trainer.train_step(images, target)
        timer.batch_end()
        with torch.no_grad():
            tgt = target.to(trainer.device, non_blocking=True)
            c1, c5 = correct(out.detach().float(), tgt, topk=(1, 5))
            bs = float(tgt.size(0))
            acc += torch.stack([torch.tensor(bs, device=trainer.device), loss.detach().float() * bs,
                                c1.reshape(()), c5.reshape(())])
        if (i + 1) % print_freq == 0 or i + 1 == n or (short_epoch and i == 10):
            tot, lsum, s1, s5 = comm.sum_tensor(acc).tolist()       # the only host sync of the loop
            acc.zero_()
            health = trainer.exchanger.check_health()               # raises when a peer's flags stopped arriving
            if tb is not None and health:
                tb.log("net/exchange_dropped", health.get("dropped", 0))
            tot = max(tot, 1.0)
            losses.update(lsum / tot, tot)
            top1.update(100.0 * s1 / tot, tot)
            top5.update(100.0 * s5 / tot, tot)
            if tb is not None:
                tb.log_memory()
                tb.log_trn_times(timer.batch_time.val, timer.data_time.val, images.size(0))
                tb.log_trn_loss(losses.val, top1.val, top5.val)
                rg, tg = net_meter.update_bandwidth()
                tb.log("sizes/batch_total", tot)
                tb.log("net/recv_gbit", rg)
                tb.log("net/transmit_gbit", tg)
            if log is not None:
                log.verbose(f"Epoch: [{epoch}][{i + 1}/{n}]\tTime {timer.batch_time.val:.3f} ({timer.batch_time.avg:.3f})\t"
                            f"Loss {losses.val:.4f} ({losses.avg:.4f})\tAcc@1 {top1.val:.3f} ({top1.avg:.3f})\t"
                            f"Acc@5 {top5.val:.3f} ({top5.avg:.3f})\tData {timer.data_time.val:.3f} ({timer.data_time.avg:.3f})")
        if tb is not None:
            tb.update_step_count(images.size(0) * comm.world_size())
    return losses.avg, top1.avg, top5.avg
