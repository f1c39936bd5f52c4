"""Console / file / scalar loggers.

* ``TableLogger`` -- fixed-width stdout table (reference CIFAR10/core.py:31-37).
* ``TSVLogger`` -- DAWNBench ``epoch\\thours\\ttop1Accuracy`` lines (CIFAR10/dawn.py:89-96).
* ``StatsLogger`` -- per-batch ``loss``/``correct`` accumulation (core.py:161-173),
  kept on device; no host sync until ``mean`` is asked for.
* ``FileLogger`` -- ``verbose.log`` / ``event.log`` / ``debug.log`` + console
  (IMAGENET/training/logger.py:74-121).
* ``ScalarLogger`` / ``TensorboardLogger`` -- scalar streams with x-axis = examples seen,
  exported to ``scalars.json`` (logger.py:13-68).  tensorboardX / wandb are optional:
  when missing the scalars are still collected and exported.
"""
from __future__ import annotations

import json
import logging as _pylog
import os
import sys
from typing import Dict, List

import numpy as np
import torch


class TableLogger:
    def __init__(self, stream=None):
        self.stream = stream or sys.stdout
        self.keys = None

    def append(self, output: Dict):
        if self.keys is None:
            self.keys = list(output.keys())
            print(*(f"{k:>12s}" for k in self.keys), file=self.stream)
        row = []
        for k in self.keys:
            v = output[k]
            if isinstance(v, (float, np.floating)):
                row.append(f"{float(v):12.4f}")
            else:
                row.append(f"{v!s:>12}")
        print(*row, file=self.stream)


class TSVLogger:
    def __init__(self):
        self.log = ["epoch\thours\ttop1Accuracy"]

    def append(self, output: Dict):
        epoch = output["epoch"]
        hours = output["total time"] / 3600.0
        acc = output["test acc"] * 100.0
        self.log.append(f"{epoch}\t{hours:.8f}\t{acc:.2f}")

    def __str__(self):
        return "\n".join(self.log)


class StatsLogger:
    def __init__(self, keys):
        self._stats = {k: [] for k in keys}

    def append(self, output: Dict):
        for k, bucket in self._stats.items():
            bucket.append(output[k].detach())

    def stats(self, key) -> torch.Tensor:
        return torch.cat([t.reshape(-1) for t in self._stats[key]])

    def mean(self, key) -> float:
        s = self.stats(key)
        if s.numel() == 0:
            return float("nan")
        return float(s.double().mean().item())


class NoOp:
    """Swallows every attribute access / call (logger.py:124-127)."""

    def __getattr__(self, *_):
        def no_op(*args, **kwargs):
            return None
        return no_op


class FileLogger:
    def __init__(self, output_dir: str = "", is_master: bool = False, is_rank0: bool = False):
        self.output_dir = output_dir
        if is_master and output_dir:
            os.makedirs(output_dir, exist_ok=True)
        self.logger = self._build(output_dir, is_rank0) if is_rank0 else NoOp()

    @staticmethod
    def _build(output_dir: str, log_to_file: bool):
        logger = _pylog.getLogger(f"layerwise_b200.{output_dir or 'console'}.{id(output_dir)}")
        logger.setLevel(_pylog.DEBUG)
        logger.propagate = False
        fmt = _pylog.Formatter("%(message)s")
        tfmt = _pylog.Formatter("%(asctime)s - %(message)s")
        if output_dir and log_to_file:
            os.makedirs(output_dir, exist_ok=True)
            for name, level, f in (("verbose.log", _pylog.DEBUG, fmt), ("event.log", _pylog.WARNING, tfmt),
                                   ("debug.log", _pylog.DEBUG, tfmt)):
                h = _pylog.FileHandler(os.path.join(output_dir, name))
                h.setLevel(level)
                h.setFormatter(f)
                logger.addHandler(h)
        ch = _pylog.StreamHandler(sys.stdout)
        ch.setLevel(_pylog.INFO)
        ch.setFormatter(fmt)
        logger.addHandler(ch)
        return logger

    def console(self, *args):
        self.logger.debug(" ".join(str(a) for a in args))

    def event(self, *args):
        self.logger.warning(" ".join(str(a) for a in args))

    def verbose(self, *args):
        self.logger.info(" ".join(str(a) for a in args))


class ScalarLogger:
    """Collects ``tag -> [(examples_seen, value)]``; the sink-independent core."""

    def __init__(self, output_dir: str = "", is_master: bool = False):
        self.output_dir = output_dir
        self.is_master = is_master
        self.current_step = 0
        self.scalars: Dict[str, List] = {}

    def log(self, tag: str, val):
        if not self.is_master:
            return
        self.scalars.setdefault(tag, []).append((self.current_step, float(val)))

    def update_step_count(self, batch_total):
        self.current_step += int(batch_total)

    def close(self):
        if self.is_master and self.output_dir:
            os.makedirs(self.output_dir, exist_ok=True)
            with open(os.path.join(self.output_dir, "scalars.json"), "w") as f:
                json.dump(self.scalars, f)


class TensorboardLogger(ScalarLogger):
    """Scalar logger with the reference's tag families and optional TB / wandb mirrors."""

    def __init__(self, output_dir: str = "", is_master: bool = False, world_size: int = 1):
        super().__init__(output_dir, is_master)
        self.world_size = world_size
        self.writer = None
        self.wandb = None
        if is_master and output_dir:
            try:
                from tensorboardX import SummaryWriter  # optional
                self.writer = SummaryWriter(output_dir)
            except Exception:
                try:
                    from torch.utils.tensorboard import SummaryWriter  # optional
                    self.writer = SummaryWriter(output_dir)
                except Exception:
                    self.writer = None

    def attach_wandb(self, wandb_module):
        self.wandb = wandb_module

    def log(self, tag, val):
        super().log(tag, val)
        if not self.is_master:
            return
        if self.writer is not None:
            self.writer.add_scalar(tag, val, self.current_step)
        if self.wandb is not None:
            self.wandb.log({tag: val}, step=int(self.current_step))

    def close(self):
        super().close()
        if self.writer is not None:
            self.writer.close()

    def log_memory(self):
        if not (self.is_master and torch.cuda.is_available()):
            return
        gb = 1e-9
        self.log("memory/allocated_gb", torch.cuda.memory_allocated() * gb)
        self.log("memory/max_allocated_gb", torch.cuda.max_memory_allocated() * gb)
        self.log("memory/cached_gb", torch.cuda.memory_reserved() * gb)
        self.log("memory/max_cached_gb", torch.cuda.max_memory_reserved() * gb)

    def log_trn_times(self, batch_time, data_time, batch_size):
        if not self.is_master:
            return
        self.log("times/step", 1000 * batch_time)
        self.log("times/data", 1000 * data_time)
        per_gpu = batch_size / max(batch_time, 1e-12)
        self.log("times/1gpu_images_per_sec", per_gpu)
        self.log("times/allgpu_images_per_sec", self.world_size * per_gpu)

    def log_trn_loss(self, loss, top1, top5):
        self.log("losses/xent", loss)
        self.log("losses/train_1", top1)
        self.log("losses/train_5", top5)

    def log_eval(self, top1, top5, seconds):
        self.log("losses/test_1", top1)
        self.log("losses/test_5", top5)
        self.log("times/eval_sec", seconds)

    def log_size(self, bs=None, sz=None):
        if bs:
            self.log("sizes/batch", bs)
        if sz:
            self.log("sizes/image", sz)
