"""Debug aids: cross-rank replica checksums and NVTX ranges.

* ``DDPX_DEBUG_CHECKSUM=1`` - after every optimizer step the trainer sums every parameter in fp64, gathers the sums from
  all ranks and raises if any rank disagrees with rank 0 (a desynchronised replica is the one bug class data parallelism
  adds; torch offers ``TORCH_DISTRIBUTED_DEBUG`` for its reducer, reducer.hpp:547-549).  On the P2P backend replicas are
  bit-identical by construction, so the tolerance is exact there; over NCCL / gloo all-reduce is also deterministic
  across ranks, so exact equality is demanded everywhere.
* ``DDPX_NVTX=1`` - ``nvtx_range("forward")`` etc. become real NVTX ranges (visible to Nsight Systems / Compute);
  otherwise they cost one attribute lookup.
"""
from __future__ import annotations

import contextlib
import os
from typing import Iterable, List

import torch


def checksum_enabled() -> bool:
    return os.environ.get("DDPX_DEBUG_CHECKSUM", "0") not in ("", "0")


def nvtx_enabled() -> bool:
    return os.environ.get("DDPX_NVTX", "0") not in ("", "0") and torch.cuda.is_available()


@contextlib.contextmanager
def nvtx_range(name: str):
    if nvtx_enabled():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


def replica_checksums(params: Iterable[torch.Tensor]) -> List[float]:
    """[sum of all parameters, sum of squares] in fp64 - cheap, order-independent enough to catch a diverged replica."""
    s = torch.zeros((), dtype=torch.float64)
    q = torch.zeros((), dtype=torch.float64)
    for p in params:
        d = p.detach().double()
        s = s.to(d.device) + d.sum()
        q = q.to(d.device) + (d * d).sum()
    return [float(s), float(q)]


def assert_replicas_in_sync(params: Iterable[torch.Tensor], comm, step: int) -> List[float]:
    """Gather every rank's checksums; raise on the first mismatch.  Returns this rank's checksums."""
    mine = replica_checksums(params)
    if comm is None or comm.world_size == 1:
        return mine
    everyone = comm.all_gather_object(mine)
    for r, other in enumerate(everyone):
        if other != everyone[0]:
            raise RuntimeError(f"replica divergence at step {step}: rank {r} checksums {other} != rank 0 {everyone[0]}")
    return mine


def replicas_in_sync(params: Iterable[torch.Tensor], tol: float = 0.0) -> bool:
    """True when every rank's fp64 parameter checksums agree (collective over torch.distributed's default group; a
    single process is trivially in sync).  The P2P backend reduces every element exactly once in one order on one rank,
    so its replicas are bit-identical and ``tol`` stays 0; NCCL-style reductions may need a relative tolerance."""
    import torch.distributed as dist
    mine = replica_checksums(params)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return True
    everyone = [None] * dist.get_world_size()
    dist.all_gather_object(everyone, mine)
    ref = everyone[0]
    for other in everyone:
        for a, b in zip(other, ref):
            if a != b and not (abs(a - b) <= tol * max(abs(a), abs(b), 1e-300)):
                return False
    return True
