"""Process-group bootstrap and per-GPU spawn (reference: ``ddp_setup`` multigpu.py:24-33, ``mp.spawn`` :263,
``destroy_process_group`` :250).

``ddp_setup(rank, world_size)`` keeps the reference signature and its localhost:12355 rendezvous, but selects the
backend from where it runs: on CUDA the control plane is a gloo+nccl torch process group (NCCL is kept only as
baseline / oracle) and gradient traffic goes through the peer-to-peer NVLink backend
(``parallel.p2p.P2PBackend``); on a CPU-only host everything runs on gloo (BASELINE.json config #1).
"""
from __future__ import annotations

import datetime
import os
import sys
import traceback
from typing import Callable, Optional, Sequence

import torch
import torch.distributed as dist

from . import config

_STATE = {"backend": None, "comm": None, "device": None}


def resolve_backend(requested: Optional[str] = None) -> str:
    req = (requested or config.get().backend or "auto").lower()
    if req != "auto":
        return req
    return "p2p" if torch.cuda.is_available() else "gloo"


def ddp_setup(rank: int, world_size: int, backend: Optional[str] = None, device_index: Optional[int] = None,
              timeout_s: float = 600.0) -> str:
    """
    Args:
        rank: Unique identifier of each process
        world_size: Total number of processes
    Returns the gradient-communication backend in use ("p2p" | "nccl" | "gloo").
    """
    cfg = config.get()
    os.environ.setdefault("MASTER_ADDR", cfg.master_addr)
    os.environ.setdefault("MASTER_PORT", str(cfg.master_port))
    be = resolve_backend(backend)
    use_cuda = be in ("p2p", "nccl")
    if use_cuda:
        if not torch.cuda.is_available():
            raise RuntimeError(f"backend {be!r} needs CUDA")
        dev = rank if device_index is None else device_index
        torch.cuda.set_device(dev)
        _STATE["device"] = torch.device("cuda", dev)
        pg_backend = "cpu:gloo,cuda:nccl"
    else:
        _STATE["device"] = torch.device("cpu")
        pg_backend = "gloo"
    if not dist.is_initialized():
        dist.init_process_group(backend=pg_backend, rank=rank, world_size=world_size,
                                timeout=datetime.timedelta(seconds=timeout_s))
    _STATE["backend"] = be
    _install_rank_excepthook(rank)
    return be


def setup_from_env(backend: Optional[str] = None) -> str:
    """Bootstrap under torchrun-style env (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    return ddp_setup(rank, world, backend=backend, device_index=local)


def backend() -> Optional[str]:
    return _STATE["backend"]


def is_distributed() -> bool:
    return dist.is_available() and dist.is_initialized()


def world_size() -> int:
    return dist.get_world_size() if is_distributed() else 1


def rank() -> int:
    return dist.get_rank() if is_distributed() else 0


def get_comm():
    """The process-wide gradient communication backend object (created on first use)."""
    if _STATE["comm"] is None:
        from .parallel import comm as _comm
        _STATE["comm"] = _comm.create(_STATE["backend"] or resolve_backend(None))
    return _STATE["comm"]


def teardown() -> None:
    """Unmap peers / free communication buffers and destroy the process group (multigpu.py:250)."""
    c = _STATE.get("comm")
    if c is not None:
        try:
            c.close()
        finally:
            _STATE["comm"] = None
    if is_distributed():
        dist.destroy_process_group()
    _STATE["backend"] = None


destroy_process_group = teardown


def _install_rank_excepthook(rank: int) -> None:
    old = sys.excepthook

    def hook(tp, val, tb):
        msg = "".join(traceback.format_exception(tp, val, tb))
        sys.stderr.write("".join(f"[rank{rank}] {ln}\n" for ln in msg.splitlines()))
    if getattr(old, "_ddpx", False):
        return
    hook._ddpx = True  # type: ignore[attr-defined]
    sys.excepthook = hook


def spawn(fn: Callable, args: Sequence = (), nprocs: Optional[int] = None, join: bool = True, start_method: str = "spawn"):
    """One process per device, ``fn(i, *args)``; fail-fast like ``torch.multiprocessing.spawn`` (a raising child
    terminates its siblings and the parent re-raises with the child's traceback, spawn.py:158-211)."""
    import torch.multiprocessing as mp
    if nprocs is None:
        nprocs = torch.cuda.device_count() if torch.cuda.is_available() else 1
    return mp.spawn(fn, args=tuple(args), nprocs=nprocs, join=join, start_method=start_method)
