"""Checkpoint save / resume in the reference's on-disk format.

The reference writes ``{'epoch': epoch+1, 'state_dict', 'best_top5', 'optimizer'}``
with ``torch.save`` (IMAGENET/training/train_imagenet_nv.py:663-669) and reads the
same four keys on ``--resume`` (:193-198).  Files written here load in the reference
and vice versa.  Extra, optional keys the reference loader ignores:
``ef_residual`` (error-feedback state), ``masters`` (fp32 master weights, fixes D14),
``rng`` and ``exchange_step``.
"""
from __future__ import annotations

import os
import shutil
from typing import Optional

import torch


def residual_sidecar(filename: str, rank: int) -> str:
    """Per-rank file next to a checkpoint that holds THAT rank's error-feedback residual."""
    return f"{filename}.ef_rank{rank}"


def save_rank_residual(filename: str, exchanger, rank: int) -> Optional[str]:
    """Error-feedback residuals differ from rank to rank (each rank defers its own unsent gradient mass), while the
    checkpoint proper is written by rank 0 only (reference train_imagenet_nv.py:247).  Every rank therefore writes its
    residual into a small sidecar file; resume reads the sidecar of the resuming rank.  Returns the path written."""
    if exchanger is None or not hasattr(exchanger, "state_dict"):
        return None
    ex = exchanger.state_dict()
    if ex.get("residual") is None:
        return None
    path = residual_sidecar(filename, rank)
    torch.save({"residual": ex["residual"], "step": ex.get("step", 0), "rank": rank}, path)
    return path


def save_checkpoint(epoch, model, best_top5, optimizer, is_best=False,
                    filename="checkpoint.pth.tar", logdir: Optional[str] = None,
                    exchanger=None, masters=None, save_rng: bool = True):
    state = {
        "epoch": epoch + 1,
        "state_dict": model.state_dict(),
        "best_top5": best_top5,
        "optimizer": optimizer.state_dict(),
    }
    if exchanger is not None and hasattr(exchanger, "state_dict"):
        ex = exchanger.state_dict()
        state["ef_residual"] = ex.get("residual")
        state["exchange_step"] = ex.get("step", 0)
    if masters is not None:
        state["masters"] = [m.detach().clone() for m in masters]
    if save_rng:
        state["rng"] = {"cpu": torch.get_rng_state()}
        if torch.cuda.is_available():
            state["rng"]["cuda"] = torch.cuda.get_rng_state()
    torch.save(state, filename)
    if is_best and logdir:
        os.makedirs(logdir, exist_ok=True)
        shutil.copyfile(filename, os.path.join(logdir, os.path.basename(filename)))
    return state


def load_checkpoint(path, model, optimizer=None, map_location=None, exchanger=None,
                    masters=None, restore_rng: bool = False, strict: bool = True, rank: int = 0, world: int = 1):
    """Restore what :func:`save_checkpoint` wrote; returns ``(start_epoch, best_top5)``.

    Error-feedback residual: the resuming rank's own sidecar (:func:`save_rank_residual`) when it exists; the residual
    inside the checkpoint only in a single-rank job (it is rank 0's); otherwise the residual restarts from zero -- the
    deferred gradient mass of the other ranks was never saved, and loading rank 0's copy everywhere would apply it
    ``world`` times."""
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    sd = ckpt["state_dict"]
    own = model.state_dict()
    # tolerate the ``module.`` prefix of wrapped models in either direction
    if own and sd and (next(iter(own)).startswith("module.") != next(iter(sd)).startswith("module.")):
        if next(iter(sd)).startswith("module."):
            sd = {k[len("module."):]: v for k, v in sd.items()}
        else:
            sd = {"module." + k: v for k, v in sd.items()}
    model.load_state_dict(sd, strict=strict)
    if optimizer is not None and "optimizer" in ckpt:
        optimizer.load_state_dict(ckpt["optimizer"])
    if exchanger is not None:
        side = residual_sidecar(path, rank)
        if os.path.exists(side):
            sd = torch.load(side, map_location="cpu", weights_only=False)
            exchanger.load_state_dict({"residual": sd.get("residual"), "step": max(sd.get("step", 0), ckpt.get("exchange_step", 0))})
        elif ckpt.get("ef_residual") is not None and world == 1:
            exchanger.load_state_dict({"residual": ckpt["ef_residual"], "step": ckpt.get("exchange_step", 0)})
        elif "exchange_step" in ckpt:
            exchanger.load_state_dict({"residual": None, "step": ckpt.get("exchange_step", 0)})
    if masters is not None and ckpt.get("masters") is not None:
        for dst, src in zip(masters, ckpt["masters"]):
            dst.data.copy_(src)
    if restore_rng and "rng" in ckpt:
        torch.set_rng_state(ckpt["rng"]["cpu"].cpu())
        if torch.cuda.is_available() and "cuda" in ckpt["rng"]:
            torch.cuda.set_rng_state(ckpt["rng"]["cuda"].cpu())
    return ckpt.get("epoch", 0), ckpt.get("best_top5", 0.0)
