"""Failure detection for one rank: a heartbeat watchdog and env-driven fault injection.

The reference has neither; what it inherits is ``mp.spawn``'s fail-fast (kept by ``launch.spawn``) and
ProcessGroupNCCL's watchdog / heartbeat threads (ProcessGroupNCCL.hpp:66-68, 100-106).  The P2P backend has no NCCL
underneath, so the equivalent lives here:

* device side - every flag wait in csrc/runtime/p2p.cu is bounded and leaves an error word (``P2PComm.error_word()``);
* host side - ``StepWatchdog``: a daemon thread that watches a progress counter (bumped once per training step and once
  per gradient bucket launched).  If the counter stands still for ``timeout_s`` it prints what the rank was doing
  (step, last bucket, parameters that never produced a gradient, device error word) and terminates the process, so that
  ``spawn`` tears the job down instead of leaving N-1 ranks spinning on a flag.  ``DDPX_WATCHDOG_S=<seconds>`` arms it
  in ``Trainer``; 0 / unset = off.
* fault injection for tests - ``DDPX_FAULT="kill:rank=1:step=3"`` (exit code 41 at that step),
  ``"raise:rank=0:step=2"`` (Python exception) or ``"hang:rank=1:step=2"`` (sleep forever: exercises the watchdog).
"""
from __future__ import annotations

import os
import sys
import threading
import time
from typing import Callable, Optional

FAULT_EXIT_CODE = 41
WATCHDOG_EXIT_CODE = 124


class StepWatchdog:
    def __init__(self, timeout_s: float, rank: int = 0, describe: Optional[Callable[[], str]] = None,
                 on_timeout: Optional[Callable[[str], None]] = None, poll_s: Optional[float] = None):
        self.timeout_s = float(timeout_s)
        self.rank = rank
        self.describe = describe
        self.on_timeout = on_timeout
        self.poll_s = poll_s if poll_s is not None else max(0.01, min(1.0, self.timeout_s / 4))
        self._count = 0
        self._what = "init"
        self._last_change = time.monotonic()
        self._stop = threading.Event()
        self._paused = False
        self._thread = None
        self.fired = False

    # ---- producer side (training thread)
    def beat(self, what: str = "") -> None:
        self._count += 1
        if what:
            self._what = what
        self._last_change = time.monotonic()

    def pause(self) -> None:
        """Long legitimate host-side waits (evaluation, checkpoint writes) are not hangs."""
        self._paused = True

    def resume(self) -> None:
        self._last_change = time.monotonic()
        self._paused = False

    # ---- watcher
    def start(self) -> "StepWatchdog":
        if self.timeout_s <= 0 or self._thread is not None:
            return self
        self._thread = threading.Thread(target=self._run, name=f"ddpx-watchdog-{self.rank}", daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2 * self.poll_s + 0.1)
            self._thread = None

    def _run(self) -> None:
        while not self._stop.wait(self.poll_s):
            if self._paused:
                continue
            idle = time.monotonic() - self._last_change
            if idle >= self.timeout_s:
                self.fired = True
                msg = (f"[rank{self.rank}] watchdog: no progress for {idle:.1f} s (limit {self.timeout_s:.1f} s); "
                       f"last event #{self._count}: {self._what}")
                if self.describe is not None:
                    try:
                        msg += " | " + self.describe()
                    except Exception as e:           # never let diagnostics mask the timeout itself
                        msg += f" | (describe failed: {e!r})"
                if self.on_timeout is not None:
                    self.on_timeout(msg)
                    return
                print(msg, file=sys.stderr, flush=True)
                os._exit(WATCHDOG_EXIT_CODE)


def from_env(rank: int, describe: Optional[Callable[[], str]] = None) -> Optional[StepWatchdog]:
    try:
        t = float(os.environ.get("DDPX_WATCHDOG_S", "0") or 0)
    except ValueError:
        t = 0.0
    if t <= 0:
        return None
    return StepWatchdog(t, rank=rank, describe=describe).start()


# ------------------------------------------------------------------------------------------------ fault injection
class FaultPlan:
    """Parsed ``DDPX_FAULT``: ``<kind>:rank=<r>:step=<s>`` with kind in kill | raise | hang."""

    def __init__(self, kind: str, rank: int, step: int):
        self.kind, self.rank, self.step = kind, rank, step

    @staticmethod
    def parse(spec: Optional[str]) -> Optional["FaultPlan"]:
        if not spec:
            return None
        parts = spec.split(":")
        kind = parts[0].strip().lower()
        if kind not in ("kill", "raise", "hang"):
            raise ValueError(f"DDPX_FAULT: unknown kind {kind!r} in {spec!r}")
        kv = {}
        for p in parts[1:]:
            k, _, v = p.partition("=")
            kv[k.strip()] = int(v)
        return FaultPlan(kind, kv.get("rank", 0), kv.get("step", 0))

    def maybe_fire(self, rank: int, step: int) -> None:
        if rank != self.rank or step != self.step:
            return
        if self.kind == "kill":
            print(f"[rank{rank}] fault injection: exiting at step {step}", file=sys.stderr, flush=True)
            os._exit(FAULT_EXIT_CODE)
        if self.kind == "raise":
            raise RuntimeError(f"fault injection: rank {rank} raises at step {step}")
        if self.kind == "hang":
            print(f"[rank{rank}] fault injection: hanging at step {step}", file=sys.stderr, flush=True)
            while True:
                time.sleep(3600)


def fault_plan_from_env() -> Optional[FaultPlan]:
    return FaultPlan.parse(os.environ.get("DDPX_FAULT"))
