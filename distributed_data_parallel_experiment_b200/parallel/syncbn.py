"""Cross-rank BatchNorm statistics (SyncBatchNorm).

The reference considered ``nn.SyncBatchNorm.convert_sync_batchnorm`` and left it commented out (multigpu.py:127), so
per-rank statistics are the default here too; this is the optional extra.  ``StatSync`` sums small
``[2, C]`` fp32 statistic blocks over all ranks: forward ``(sum x, sum x^2)``, backward ``(sum dy, sum dy*xhat)``.

* P2P backend: one persistent symmetric buffer per (C, slot), allocated collectively at first use (every rank runs the
  same model in the same order), reduced by the same fused all-reduce kernel as the gradient buckets - a 4 KB message is
  one latency-bound launch;
* NCCL / gloo: ``torch.distributed.all_reduce``.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.distributed as dist


class StatSync:
    def __init__(self, comm=None):
        if comm is None:
            from .. import launch
            comm = launch.get_comm()
        self.comm = comm
        self.world_size = comm.world_size
        self._bufs: Dict[Tuple[int, str], torch.Tensor] = {}

    def reduce(self, stats: torch.Tensor, key: str = "") -> torch.Tensor:
        """Sum of ``stats`` (fp32, numel a multiple of 8) over all ranks, as a new tensor.  ``key`` names the caller
        (layer + direction): every key owns one persistent symmetric buffer on the P2P backend."""
        if self.world_size == 1:
            return stats
        flat = stats.reshape(-1)
        if getattr(self.comm, "name", "") == "p2p":
            n = flat.numel()
            buf = self._bufs.get((n, key))
            if buf is None:
                buf = self._bufs[(n, key)] = self.comm.alloc(n, torch.float32, stats.device)
            buf.copy_(flat)
            self.comm.all_reduce_sum_(buf, async_op=False)
            return buf.clone().view_as(stats)
        out = flat.clone()
        dist.all_reduce(out, op=dist.ReduceOp.SUM)
        return out.view_as(stats)


_default: Dict[int, StatSync] = {}


def default_sync() -> StatSync:
    """One StatSync per communication backend instance (buffers are cached in it)."""
    from .. import launch
    comm = launch.get_comm()
    s = _default.get(id(comm))
    if s is None:
        _default.clear()
        s = _default[id(comm)] = StatSync(comm)
    return s
