"""Model-size helper (reference multigpu.py:208-221), device-timed step metrics and JSONL logging.

The reference only prints wall-clock time around ``trainer.train()`` (multigpu.py:230-235); here step times
are taken with CUDA events on the launching stream and the exposed-communication time is the gap between the
end of backward compute and the completion of the last gradient bucket.
"""
from __future__ import annotations

import json
import time
from typing import Optional

import torch

# bit-based units, exactly as the reference defines them (multigpu.py:218-221)
Byte = 8
KiB = 1024 * Byte
MiB = 1024 * KiB
GiB = 1024 * MiB


def get_model_size(model: torch.nn.Module, data_width: int = 32) -> int:
    """Model size in bits: number of parameter elements x ``data_width`` (buffers excluded)."""
    num_elements = 0
    for param in model.parameters():
        num_elements += param.numel()
    return num_elements * data_width


class StepTimer:
    """CUDA-event (or perf_counter on CPU) timer for training steps."""

    def __init__(self, device: Optional[torch.device] = None):
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self._pairs = []
        self._t0 = None

    def start(self):
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._t0 = e
        else:
            self._t0 = time.perf_counter()

    def stop(self):
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._pairs.append((self._t0, e))
        else:
            self._pairs.append((self._t0, time.perf_counter()))

    def times_ms(self):
        if self.cuda:
            torch.cuda.synchronize()
            return [a.elapsed_time(b) for a, b in self._pairs]
        return [(b - a) * 1e3 for a, b in self._pairs]

    def reset(self):
        self._pairs = []


class JsonlLogger:
    """Structured step metrics, one JSON object per line (rank, step, device-ms, samples/s, comm-ms, ...)."""

    def __init__(self, path: str):
        self.path = path
        self._fh = open(path, "a") if path else None

    def log(self, **kv):
        if self._fh is None:
            return
        kv.setdefault("ts", time.time())
        self._fh.write(json.dumps(kv) + "\n")
        self._fh.flush()

    def close(self):
        if self._fh is not None:
            self._fh.close()
            self._fh = None


class StepMetrics:
    """Per-step structured metrics for the Trainer, written as JSON lines to ``DDPX_METRICS_JSONL``
    (``{rank}`` in the path is replaced by the rank).  Device time comes from CUDA events that are read back lazily -
    only once they have completed - so logging never puts a synchronisation into the training loop.

    One line per optimizer step: rank, step, device_ms, samples, samples_per_s, loss (when already on the host), world
    size, and - on a data-parallel model - backend, number of buckets, whether the optimizer is fused, and the
    communication fields of ``comm_probe`` (c10d::Logger's runtime stats, logger.hpp:42-127): comm_ms (sum of the
    buckets' kernel times), exposed_comm_ms (how long the compute stream waited for the last buckets) and per bucket
    its bytes, kernel time, algorithm (oneshot / twoshot / nvls) and bus GB/s.  The probe reads the most recent
    iteration's timers, so it is attached to the newest completed step of a flush."""

    def __init__(self, path: str, rank: int = 0, device: Optional[torch.device] = None, static: Optional[dict] = None,
                 comm_probe=None):
        self.comm_probe = comm_probe
        self.rank = rank
        self.log = JsonlLogger(path.replace("{rank}", str(rank)) if path else "")
        self.cuda = torch.cuda.is_available() and device is not None and torch.device(device).type == "cuda"
        self.static = dict(static or {})
        self._pending = []          # (step, samples, start, end)
        self._t0 = None

    @property
    def enabled(self) -> bool:
        return self.log._fh is not None

    def begin(self):
        if not self.enabled:
            return
        if self.cuda:
            self._t0 = torch.cuda.Event(enable_timing=True)
            self._t0.record()
        else:
            self._t0 = time.perf_counter()

    def end(self, step: int, samples: int):
        if not self.enabled or self._t0 is None:
            return
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
        else:
            e = time.perf_counter()
        self._pending.append((step, samples, self._t0, e))
        self._t0 = None
        self.flush(block=False)

    def flush(self, block: bool = True):
        keep, done = [], []
        for step, samples, a, b in self._pending:
            if self.cuda:
                if not block and not b.query():
                    keep.append((step, samples, a, b))
                    continue
                b.synchronize()
                ms = a.elapsed_time(b)
            else:
                ms = (b - a) * 1e3
            done.append((step, samples, ms))
        for i, (step, samples, ms) in enumerate(done):
            extra = {}
            if self.comm_probe is not None and i == len(done) - 1 and not keep:
                try:
                    extra = self.comm_probe() or {}
                except Exception as e:      # diagnostics must not take the training loop down
                    extra = {"comm_probe_error": f"{type(e).__name__}: {e}"}
            self.log.log(rank=self.rank, step=step, device_ms=round(ms, 4), samples=samples,
                         samples_per_s=round(samples / (ms / 1e3), 1) if ms > 0 else None, **self.static, **extra)
        self._pending = keep

    def close(self):
        if self.enabled:
            self.flush(block=True)
        self.log.close()
