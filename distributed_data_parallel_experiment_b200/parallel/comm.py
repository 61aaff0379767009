"""Gradient-communication backends.

* ``TorchComm`` - torch.distributed collectives (gloo on CPU; NCCL on CUDA as baseline / oracle only).
* ``P2PComm``   - the product path on B200: peer-mapped symmetric buffers + hand-written sm_100a kernels over
                  NVLink 5 / NVSwitch (parallel/p2p.py).
* ``NullComm``  - world size 1.

All expose the same small surface used by the reducer and the DataParallel wrapper.
"""
from __future__ import annotations

from typing import Any, List, Optional

import torch
import torch.distributed as dist


class _Done:
    def wait(self):
        return None


class CommBase:
    name = "base"
    supports_fused_update = False

    def __init__(self):
        self.world_size = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0

    # ---- memory -------------------------------------------------------------------------------
    def alloc(self, numel: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
        """Flat buffer for buckets / flat parameters (symmetric memory on the P2P backend)."""
        return torch.zeros(numel, dtype=dtype, device=device)

    # ---- collectives --------------------------------------------------------------------------
    def broadcast_(self, flat: torch.Tensor, src: int = 0):
        raise NotImplementedError

    def all_reduce_sum_(self, flat: torch.Tensor, async_op: bool = True):
        raise NotImplementedError

    def broadcast_object(self, obj: Any, src: int = 0) -> Any:
        if self.world_size == 1:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=src)
        return box[0]

    def all_gather_object(self, obj: Any) -> List[Any]:
        if self.world_size == 1:
            return [obj]
        out = [None] * self.world_size
        dist.all_gather_object(out, obj)
        return out

    def barrier(self):
        if self.world_size > 1:
            dist.barrier()

    def close(self):
        pass


class NullComm(CommBase):
    name = "null"

    def broadcast_(self, flat, src=0):
        return _Done()

    def all_reduce_sum_(self, flat, async_op=True):
        return _Done()


class TorchComm(CommBase):
    """gloo (CPU tensors) or NCCL (CUDA tensors) through the default process group."""

    def __init__(self, name: str = "gloo"):
        super().__init__()
        self.name = name

    def broadcast_(self, flat, src=0):
        if self.world_size == 1:
            return _Done()
        return dist.broadcast(flat, src=src, async_op=False) or _Done()

    def all_reduce_sum_(self, flat, async_op=True):
        if self.world_size == 1:
            return _Done()
        w = dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=async_op)
        return w if w is not None else _Done()


def create(name: Optional[str]) -> CommBase:
    name = (name or "auto").lower()
    if not dist.is_initialized() or dist.get_world_size() == 1:
        if name == "p2p" and torch.cuda.is_available():
            from .p2p import P2PComm
            return P2PComm()
        return NullComm()
    if name == "auto":
        name = "p2p" if torch.cuda.is_available() else "gloo"
    if name == "p2p":
        from .p2p import P2PComm
        return P2PComm()
    if name in ("nccl", "gloo"):
        return TorchComm(name)
    raise ValueError(f"unknown communication backend {name!r}")
