"""Checkpoint writer / loader.

Default file format is exactly the reference's (multigpu.py:109-113): ``torch.save`` of the *unwrapped* module's
``state_dict()`` (an ``OrderedDict[str, Tensor]`` with torch-compatible key names and ``_metadata``) to
``checkpoint.pt`` in the working directory, overwritten each time - so files load interchangeably with the
reference's ``VGG().load_state_dict(...)``.  ``save_snapshot`` / ``load_snapshot`` add the resume state the
reference never saves (optimizer, scheduler, epoch, RNG) as a superset in a separate file.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Any, Dict, Optional

import torch
from torch import nn

DEFAULT_PATH = "checkpoint.pt"


def unwrap(model: nn.Module) -> nn.Module:
    """Strip a data-parallel wrapper (``.module``), as multigpu.py:110 does."""
    return model.module if hasattr(model, "module") and isinstance(getattr(model, "module"), nn.Module) else model


def model_state_dict(model: nn.Module) -> "OrderedDict[str, torch.Tensor]":
    return unwrap(model).state_dict()


def save_checkpoint(model: nn.Module, path: str = DEFAULT_PATH) -> str:
    """Write the model state_dict exactly like the reference (tensors keep their device tag)."""
    ckp = model_state_dict(model)
    torch.save(ckp, path)
    return path


def load_checkpoint(model: nn.Module, path: str = DEFAULT_PATH, strict: bool = True, map_location: Any = None):
    """Load a ``checkpoint.pt`` written by this framework or by the reference scripts."""
    target = unwrap(model)
    if map_location is None:
        try:
            map_location = next(target.parameters()).device
        except StopIteration:
            map_location = "cpu"
    sd = torch.load(path, map_location=map_location, weights_only=True)
    return target.load_state_dict(sd, strict=strict)


def save_snapshot(path: str, model: nn.Module, optimizer: Optional[torch.optim.Optimizer] = None, scheduler=None,
                  epoch: int = 0, extra: Optional[Dict[str, Any]] = None) -> str:
    """Full training snapshot (superset of the reference format) for resume."""
    snap = {
        "MODEL_STATE": model_state_dict(model),
        "OPTIMIZER_STATE": optimizer.state_dict() if optimizer is not None else None,
        "SCHEDULER_STATE": _scheduler_state(scheduler),
        "EPOCHS_RUN": epoch,
        "RNG_CPU": torch.get_rng_state(),
        "RNG_CUDA": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None,
        "EXTRA": extra or {},
    }
    tmp = path + ".tmp"
    torch.save(snap, tmp)
    os.replace(tmp, path)
    return path


def _scheduler_state(scheduler):
    if scheduler is None:
        return None
    sd = dict(scheduler.state_dict())
    # LambdaLR keeps un-picklable lambdas under 'lr_lambdas' as None already; make sure of it
    if "lr_lambdas" in sd:
        sd["lr_lambdas"] = [None if callable(f) and getattr(f, "__name__", "") == "<lambda>" else f for f in sd["lr_lambdas"]]
    return sd


def load_snapshot(path: str, model: nn.Module, optimizer: Optional[torch.optim.Optimizer] = None, scheduler=None,
                  map_location: Any = "cpu", restore_rng: bool = False) -> int:
    """Restore a snapshot; returns the epoch to resume from."""
    snap = torch.load(path, map_location=map_location, weights_only=False)
    unwrap(model).load_state_dict(snap["MODEL_STATE"])
    if optimizer is not None and snap.get("OPTIMIZER_STATE") is not None:
        optimizer.load_state_dict(snap["OPTIMIZER_STATE"])
    if scheduler is not None and snap.get("SCHEDULER_STATE") is not None:
        sd = dict(snap["SCHEDULER_STATE"])
        sd.pop("lr_lambdas", None)
        scheduler.__dict__.update(sd)
    if restore_rng:
        torch.set_rng_state(snap["RNG_CPU"].cpu())
        if snap.get("RNG_CUDA") is not None and torch.cuda.is_available():
            torch.cuda.set_rng_state_all([s.cpu() for s in snap["RNG_CUDA"]])
    return int(snap.get("EPOCHS_RUN", 0))
