"""``DataParallel``: replicated-parameter data parallelism with a bucketed, overlapped gradient reducer.

Stands where ``torch.nn.parallel.DistributedDataParallel`` stands in the reference (``DDP(model, device_ids=[gpu_id])``,
multigpu.py:89) and keeps its observable semantics with all-default arguments:

* construction: parameter shapes are verified across ranks, then parameters and buffers are overwritten with rank 0's
  (X2, X3);
* every forward in training first overwrites the module buffers (BatchNorm running stats) with rank 0's (X4);
* backward averages gradients over ranks, bucket by bucket in reverse-definition order, overlapped with the rest of
  backward (X5); buckets are rebuilt once in the observed ready order with the [1 MiB, bucket_cap_mb] caps (X6);
* ``no_sync()`` accumulates locally and reduces on the next synchronised backward;
* ``.module`` is the wrapped model.

It is not a wrapper around torch DDP: buckets live in the communication backend's (peer-mapped) memory and on the P2P
backend one hand-written kernel per bucket does all-reduce + 1/world + optimizer update (``optimizer=`` argument).
"""
from __future__ import annotations

import contextlib
from typing import Optional

import torch
from torch import nn

from .. import config
from .reducer import Reducer


class DataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0,
                 broadcast_buffers: bool = True, bucket_cap_mb: Optional[float] = None,
                 find_unused_parameters: bool = False, gradient_as_bucket_view: bool = True,
                 comm=None, optimizer: Optional[torch.optim.Optimizer] = None, init_sync: bool = True):
        super().__init__()
        if find_unused_parameters:
            raise NotImplementedError("find_unused_parameters=True is not supported (the reference uses the default False)")
        self.module = module
        self.device_ids = device_ids
        cfg = config.get()
        self.broadcast_buffers = broadcast_buffers and cfg.broadcast_buffers
        cap_mb = cfg.bucket_cap_mb if bucket_cap_mb is None else bucket_cap_mb
        self.bucket_bytes_cap = int(cap_mb * 1024 * 1024)
        # torch: the 1 MiB first-bucket cap applies only with the default bucket_cap_mb (distributed.py:1255-1280)
        self._first_bucket_bytes = int(cfg.first_bucket_mb * 1024 * 1024) if bucket_cap_mb is None else self.bucket_bytes_cap
        if comm is None:
            from .. import launch
            comm = launch.get_comm()
        self.comm = comm
        self.world_size = comm.world_size
        self.rank = comm.rank
        self.require_backward_grad_sync = True
        self.require_forward_param_sync = True
        self._names, self._params = [], []
        for name, p in module.named_parameters():
            if p.requires_grad:
                self._names.append(name)
                self._params.append(p)
        if not self._params:
            raise RuntimeError("DataParallel: module has no parameter that requires a gradient")
        self._buffers_list = [b for _, b in module.named_buffers()]
        self._verify_params()
        self.fused = None
        if optimizer is not None and getattr(comm, "supports_fused_update", False) and cfg.fuse_optimizer:
            from .p2p import FusedUpdate
            self.fused = FusedUpdate.try_create(comm, self._params, optimizer)
        if init_sync:
            self._sync_module_states()
        self.reducer = Reducer(self._params, comm, self.bucket_bytes_cap,
                               first_bucket_bytes=self._first_bucket_bytes, fused=self.fused)
        # X4 moved off the critical path: rank 0's buffers (BatchNorm running statistics) are final once ITS forward has
        # run, so the broadcast that torch DDP issues at the head of the next forward is enqueued on the communication
        # stream when the first gradient bucket launches and overlaps the rest of backward (it used to open every step
        # with a flag round trip on the compute stream: 25 us + rank skew).  DDPX_EARLY_BUFFER_SYNC=0 restores the old place.
        self._buffers_synced_early = False
        self._early_event = None
        if (self.broadcast_buffers and self._buffers_list and self.world_size > 1 and cfg.early_buffer_sync
                and getattr(comm, "stream", None) is not None and hasattr(comm, "broadcast_tensors_")):
            self.reducer.on_first_launch = self._early_buffer_sync

    # ------------------------------------------------------------------------------------------ construction
    def _verify_params(self):
        """All ranks must hold the same number of parameters with the same shapes (reducer.hpp:602-605)."""
        if self.world_size == 1:
            return
        sig = [(n, tuple(p.shape), str(p.dtype)) for n, p in zip(self._names, self._params)]
        sigs = self.comm.all_gather_object(sig)
        for r, s in enumerate(sigs):
            if s != sigs[0]:
                raise RuntimeError(f"DataParallel: parameter set of rank {r} does not match rank 0 "
                                   f"({len(s)} vs {len(sigs[0])} tensors or differing shapes)")

    def _flat_broadcast(self, tensors):
        """Coalesced broadcast from rank 0: one flat buffer per dtype (broadcast_coalesced, c10d/comm.hpp:13-17)."""
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault(t.dtype, []).append(t)
        for dtype, ts in by_dtype.items():
            if hasattr(self.comm, "broadcast_tensors_"):
                self.comm.broadcast_tensors_(ts, src=0)
                continue
            flat = torch.cat([t.detach().reshape(-1) for t in ts]) if len(ts) > 1 else ts[0].detach().reshape(-1).clone()
            self.comm.broadcast_(flat, src=0)
            if self.rank != 0:
                off = 0
                for t in ts:
                    n = t.numel()
                    t.detach().copy_(flat[off:off + n].view_as(t))
                    off += n

    def _sync_module_states(self):
        if self.world_size == 1:
            return
        with torch.no_grad():
            self._flat_broadcast(list(self.module.parameters()))
            if self._buffers_list:
                self._flat_broadcast(self._buffers_list)

    def _sync_buffers(self):
        if self.world_size == 1 or not self._buffers_list:
            return
        with torch.no_grad():
            self._flat_broadcast(self._buffers_list)

    def _early_buffer_sync(self):
        """Called by the reducer right after the first bucket of a synchronised backward was enqueued."""
        comm = self.comm
        with torch.no_grad(), torch.cuda.stream(comm.stream):
            self._flat_broadcast(self._buffers_list)
            ev = torch.cuda.Event()
            ev.record(comm.stream)
        self._early_event = ev
        self._buffers_synced_early = True

    # ------------------------------------------------------------------------------------------ forward
    def forward(self, *inputs, **kwargs):
        grad_on = torch.is_grad_enabled()
        if grad_on and self.require_backward_grad_sync:
            self.reducer.rebuild_buckets()
        if self._buffers_synced_early:
            # this forward's buffer broadcast already ran during the previous backward (see __init__)
            self._buffers_synced_early = False
            torch.cuda.current_stream().wait_event(self._early_event)
        elif self.require_forward_param_sync and self.broadcast_buffers and self._buffers_list:
            self._sync_buffers()
        out = self.module(*inputs, **kwargs)
        if grad_on and self.require_backward_grad_sync:
            self.require_forward_param_sync = True
            self.reducer.prepare_for_backward()
        else:
            self.require_forward_param_sync = False
        return out

    @contextlib.contextmanager
    def no_sync(self):
        """Skip gradient synchronisation inside the context; gradients accumulate in ``param.grad`` and are reduced by
        the first backward after leaving it (torch distributed.py:1474-1500)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    # ------------------------------------------------------------------------------------------ misc
    def train(self, mode: bool = True):
        super().train(mode)
        return self

    def get_logging_data(self) -> dict:
        """Subset of c10d::Logger's DDPLoggingData (logger.hpp:42-127)."""
        r = self.reducer
        return {"world_size": self.world_size, "rank": self.rank, "backend": self.comm.name,
                "bucket_cap_bytes": self.bucket_bytes_cap, "num_buckets": len(r.buckets), "bucket_sizes": r.bucket_sizes(),
                "rebuilt_bucket_sizes": r.stats["rebuilt_bucket_sizes"], "has_rebuilt_buckets": r.has_rebuilt,
                "iteration": r.iteration, "fused_optimizer": self.fused is not None}

    def measure_exposed_comm(self, on: bool = True):
        """Start / stop recording the exposed-communication time of every synchronised backward."""
        self.reducer.measure_exposed = on
        if on:
            self.reducer.stats["exposed_events"] = []

    def exposed_comm_ms(self):
        """Per-iteration milliseconds the compute stream waited for gradient buckets after backward's last kernel."""
        evs = self.reducer.stats.get("exposed_events", [])
        if not evs:
            return []
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in evs]

    def close(self):
        self.reducer.close()
