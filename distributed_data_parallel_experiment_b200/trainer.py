"""``Trainer`` - the framework's public training loop, same surface as the reference's (multigpu.py:74-119,
singlegpu.py:85-128):

    Trainer(model, train_data, optimizer, gpu_id, save_every[, scheduler]).train(max_epochs)

Differences, all supersets: ``scheduler`` is optional (the upstream tutorial's 5-argument form works), the loss is
pluggable (cross-entropy default, MSE for the toy regressor), the data-parallel wrapper is chosen from the process
group (none when single process), and ``train_step`` is the end-to-end public call (pinned host batch in, host loss
out).  Print lines, save cadence (``epoch % save_every == 0``, rank 0 only, ``checkpoint.pt``) and the per-batch
``scheduler.step()`` are the reference's.
"""
from __future__ import annotations

from typing import Iterable, Callable, Optional, Union

import torch
import torch.distributed as dist

from . import checkpoint, config, ops, optim as _optim
from .parallel import DataParallel
from .utils import debug as _debug, watchdog as _watchdog


def _resolve_loss(loss_fn: Union[str, Callable, None]) -> Callable:
    if loss_fn is None or loss_fn == "cross_entropy":
        return ops.cross_entropy
    if loss_fn == "mse":
        return ops.mse_loss
    if callable(loss_fn):
        return loss_fn
    raise ValueError(f"unknown loss {loss_fn!r}")


class Trainer:
    def __init__(
        self,
        model: torch.nn.Module,
        train_data,
        optimizer: torch.optim.Optimizer,
        gpu_id: int,
        save_every: int,
        scheduler=None,
        loss_fn: Union[str, Callable, None] = None,
        distributed: Optional[bool] = None,
        accum_steps: int = 1,
        checkpoint_path: Optional[str] = None,
    ) -> None:
        self.gpu_id = gpu_id
        self.device = torch.device("cuda", gpu_id) if torch.cuda.is_available() and gpu_id is not None and gpu_id >= 0 \
            and _model_wants_cuda(model) else torch.device("cpu")
        self.model = model.to(self.device)
        self.train_data = train_data
        self.optimizer = optimizer
        self.save_every = save_every
        self.scheduler = scheduler
        self.loss_fn = _resolve_loss(loss_fn)
        self.accum_steps = max(1, int(accum_steps))
        self.checkpoint_path = checkpoint_path or config.get().checkpoint_path
        self._micro = 0
        self.last_loss = None
        if distributed is None:
            distributed = dist.is_available() and dist.is_initialized()
        self.distributed = distributed
        if self.device.type == "cuda":
            _optim.fuse(self.optimizer)            # optimizer.step() -> one native multi-tensor kernel
        if distributed:
            # multigpu.py:89 - wrap; the optimizer was built on the raw module's parameters and keeps working
            self.model = DataParallel(self.model, device_ids=[gpu_id] if self.device.type == "cuda" else None,
                                      optimizer=self.optimizer)
        # failure detection / debug aids - all off unless their environment variable is set
        self.global_step = 0
        self._rank = dist.get_rank() if (dist.is_available() and dist.is_initialized()) else 0
        self._fault = _watchdog.fault_plan_from_env()
        self._watchdog = _watchdog.from_env(self._rank, self._describe_state)
        self._checksum = _debug.checksum_enabled()
        static = {"world_size": dist.get_world_size() if self.distributed else 1}
        if isinstance(self.model, DataParallel):
            info = self.model.get_logging_data()
            static.update(backend=info["backend"], fused_optimizer=info["fused_optimizer"])
        from . import metrics as _metrics
        self._metrics = _metrics.StepMetrics(config.get().metrics_jsonl, self._rank, self.device, static,
                                             comm_probe=self._comm_probe if isinstance(self.model, DataParallel) else None)
        if self._metrics.enabled and isinstance(self.model, DataParallel) and self.device.type == "cuda":
            self.model.measure_exposed_comm(True)

    def _comm_probe(self) -> dict:
        """Communication fields of a metrics line: per-bucket kernel time / algorithm / bus bandwidth from the native
        reducer's device timers, and the exposed communication of the last synchronised backward."""
        red = self.model.reducer
        out = {}
        buckets = red.bucket_times() if hasattr(red, "bucket_times") else []
        if buckets:
            out["buckets"] = [{k: (round(v, 4) if isinstance(v, float) else v) for k, v in b.items()} for b in buckets]
            out["comm_ms"] = round(sum(b["ms"] for b in buckets if b["ms"] is not None), 4)
        evs = red.stats.get("exposed_events") or []
        if evs:
            a, b = evs[-1]
            b.synchronize()
            out["exposed_comm_ms"] = round(a.elapsed_time(b), 4)
            del evs[:-4]
        return out

    def _describe_state(self) -> str:
        parts = [f"global step {self.global_step}"]
        red = getattr(self.model, "reducer", None)
        if red is not None:
            parts.append(f"reducer iteration {red.iteration}, buckets launched {red.stats['launched']}, "
                         f"parameters without gradient {red._unmarked() if red.expect_finalize else []}")
        comm = getattr(self.model, "comm", None)
        if comm is not None and hasattr(comm, "error_word"):
            parts.append(f"device error word {comm.error_word():#x}")
        return "; ".join(parts)

    # ------------------------------------------------------------------------------------------ one step
    def _run_batch(self, source, targets):
        """zero_grad -> forward -> loss -> backward -> optimizer.step -> scheduler.step (multigpu.py:92-98), with
        optional gradient accumulation over ``accum_steps`` micro-batches under ``no_sync``."""
        if self._fault is not None:
            self._fault.maybe_fire(self._rank, self.global_step)
        if self._watchdog is not None:
            self._watchdog.beat(f"step {self.global_step} begins")
        last = (self._micro + 1) % self.accum_steps == 0
        if self._micro % self.accum_steps == 0:
            self._metrics.begin()
            self.optimizer.zero_grad()
        if self.distributed and not last:
            with self.model.no_sync():
                loss = self._forward_backward(source, targets)
        else:
            loss = self._forward_backward(source, targets)
        self._micro += 1
        if last:
            with _debug.nvtx_range("optimizer"):
                self.optimizer.step()
            if self.scheduler is not None:
                self.scheduler.step()
            if self._checksum:
                _debug.assert_replicas_in_sync(self.model.parameters(), getattr(self.model, "comm", None), self.global_step)
        if last:
            self._metrics.end(self.global_step, int(source.shape[0]) * self.accum_steps)
        self.global_step += 1
        self.last_loss = loss
        return loss

    def _forward_backward(self, source, targets):
        with _debug.nvtx_range("forward"):
            output = self.model(source)
            loss = self.loss_fn(output, targets)
        with _debug.nvtx_range("backward"):
            if self.accum_steps > 1:
                (loss / self.accum_steps).backward()
            else:
                loss.backward()
        return loss.detach()

    def train_step(self, source: torch.Tensor, targets: torch.Tensor) -> float:
        """End-to-end step: copy the (pinned) host batch to the device, run the step, read the loss back."""
        source = source.to(self.device, non_blocking=True)
        targets = targets.to(self.device, non_blocking=True)
        loss = self._run_batch(source, targets)
        return float(loss.item())

    def train_stream(self, batches: Iterable):
        """Pipelined end-to-end loop over host batches ``(source, targets)``; yields one host-side loss per batch, in order.

        While step *i* computes, batch *i+1* is uploaded on a copy stream (pinned memory makes that a true DMA) and the
        loss of step *i-1* is read back, so neither the PCIe copy nor the read-back sits between two steps' kernels.
        Every batch is still copied and every loss still read - only the order of the host's waits changes.
        CPU runs degrade to the plain loop."""
        if self.device.type != "cuda":
            for source, targets in batches:
                yield self.train_step(source, targets)
            return
        compute = torch.cuda.current_stream(self.device)
        copy = getattr(self, "_copy_stream", None)
        if copy is None:
            copy = self._copy_stream = torch.cuda.Stream(self.device)

        def upload(batch):
            ev = torch.cuda.Event()
            with torch.cuda.stream(copy):       # buffers are allocated on the copy stream; record_stream() below keeps
                src = batch[0].to(self.device, non_blocking=True)   # the allocator from reusing them under the step
                tgt = batch[1].to(self.device, non_blocking=True)
                ev.record(copy)
            return src, tgt, ev

        it = iter(batches)
        try:
            nxt = upload(next(it))
        except StopIteration:
            return
        # losses land in pinned host slots by async copies; the host waits on the slot's own event, i.e. for step i-1
        # only, never for the step it has just enqueued
        slots = getattr(self, "_loss_slots", None)
        if slots is None:
            slots = self._loss_slots = [torch.empty((), dtype=torch.float32).pin_memory() for _ in range(2)]
        prev = None                                   # (slot, event) of the previous step
        i = 0
        while nxt is not None:
            src, tgt, ready = nxt
            compute.wait_event(ready)
            src.record_stream(compute); tgt.record_stream(compute)
            loss = self._run_batch(src, tgt)
            slot = slots[i & 1]
            slot.copy_(loss.detach().float().reshape(()), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(compute)
            try:
                nxt = upload(next(it))
            except StopIteration:
                nxt = None
            if prev is not None:
                prev[1].synchronize()
                yield float(prev[0])
            prev = (slot, ev)
            i += 1
        if prev is not None:
            prev[1].synchronize()
            yield float(prev[0])

    # ------------------------------------------------------------------------------------------ epochs
    def _run_epoch(self, epoch):
        b_sz = _batch_size_of(self.train_data)
        print(f"[GPU{self.gpu_id}] Epoch {epoch} | Batchsize: {b_sz} | Steps: {len(self.train_data)}")
        sampler = getattr(self.train_data, "sampler", None)
        if self.distributed and hasattr(sampler, "set_epoch"):
            sampler.set_epoch(epoch)                     # multigpu.py:103
        if self.device.type == "cuda":
            # same steps as the reference's loop (multigpu.py:104-107), with the upload of batch i+1 overlapping step i
            # (the reference copies blockingly between steps - SURVEY.md K16)
            for _ in self.train_stream(self.train_data):
                pass
            return
        for source, targets in self.train_data:
            source = source.to(self.device, non_blocking=True)
            targets = targets.to(self.device, non_blocking=True)
            self._run_batch(source, targets)

    def _save_checkpoint(self, epoch):
        PATH = self.checkpoint_path
        checkpoint.save_checkpoint(self.model, PATH)     # unwrapped module's state_dict, torch.save (multigpu.py:110-112)
        print(f"Epoch {epoch} | Training checkpoint saved at {PATH}")

    def train(self, max_epochs: int):
        for epoch in range(max_epochs):
            if self._watchdog is not None:
                self._watchdog.resume()
            self._run_epoch(epoch)
            if self._watchdog is not None:
                self._watchdog.pause()          # checkpoint writes and whatever the caller does next are not hangs
            self._metrics.flush(block=True)
            # multigpu.py:118 gates on gpu_id == 0 (rank 0); singlegpu.py:127 saves unconditionally
            if (not self.distributed or self.gpu_id == 0) and epoch % self.save_every == 0:
                self._save_checkpoint(epoch)


def _model_wants_cuda(model) -> bool:
    """CPU/gloo runs (BASELINE.json config #1) pass a rank as ``gpu_id`` on a host that may also have GPUs; keep the model
    where the process group lives."""
    if dist.is_available() and dist.is_initialized():
        from . import launch
        be = launch.backend()
        if be == "gloo":
            return False
    return True


def _batch_size_of(loader):
    """The reference prints ``len(next(iter(train_data))[0])`` - which loads and discards a whole batch
    (multigpu.py:101); read the configured size instead and only fall back to a probe when it is unknown."""
    bs = getattr(loader, "batch_size", None)
    if bs is not None:
        return bs
    try:
        return len(next(iter(loader))[0])
    except Exception:
        return 0
