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


class ProcessRaisedException(RuntimeError):
    """A spawned rank raised; the message carries that rank's traceback."""

    def __init__(self, msg: str, rank: int, pid: int):
        super().__init__(msg)
        self.error_index, self.error_pid = rank, pid


class ProcessExitedException(RuntimeError):
    """A spawned rank died without raising (signal, os._exit, watchdog exit code)."""

    def __init__(self, msg: str, rank: int, pid: int, exit_code: int, signal_name: Optional[str] = None):
        super().__init__(msg)
        self.error_index, self.error_pid, self.exit_code, self.signal_name = rank, pid, exit_code, signal_name


def _child_main(fn, rank, args, error_path):
    """Entry of a spawned rank: die with the parent, prefix uncaught tracebacks with the rank, run ``fn(rank, *args)``,
    hand an exception's traceback to the parent through ``error_path``."""
    try:
        import ctypes
        import signal
        ctypes.CDLL("libc.so.6").prctl(1, signal.SIGINT)         # PR_SET_PDEATHSIG: no orphans when the launcher dies
    except Exception:
        pass
    _install_rank_excepthook(rank)
    try:
        fn(rank, *args)
    except KeyboardInterrupt:
        pass                                                      # the parent is tearing the job down
    except Exception:
        import pickle
        with open(error_path, "wb") as fh:
            pickle.dump(traceback.format_exc(), fh)
        sys.exit(1)


class SpawnContext:
    """Handle on the ranks started by ``spawn(join=False)``; ``join()`` returns True once all of them exited cleanly."""

    def __init__(self, processes, error_paths):
        self.processes, self.error_paths = processes, error_paths

    def pids(self):
        return [p.pid for p in self.processes]

    def _terminate_all(self, grace_s: float = 10.0):
        import time
        for p in self.processes:
            if p.is_alive():
                p.terminate()                                     # SIGTERM first ...
        deadline = time.time() + grace_s
        for p in self.processes:
            p.join(max(0.0, deadline - time.time()))
        for p in self.processes:
            if p.is_alive():
                p.kill()                                          # ... SIGKILL for what ignored it
                p.join()

    def join(self, timeout: Optional[float] = None) -> bool:
        import multiprocessing.connection as mpc
        import pickle
        import signal
        alive = [p for p in self.processes if p.exitcode is None]
        if alive:
            mpc.wait([p.sentinel for p in alive], timeout=timeout)
        failed = None
        for i, p in enumerate(self.processes):
            if p.exitcode is not None and p.exitcode != 0:
                failed = i
                break
        if failed is None:
            return all(p.exitcode == 0 for p in self.processes)
        # fail fast: one rank is gone, the others would wait for it in the next collective forever
        self._terminate_all()
        p = self.processes[failed]
        path = self.error_paths[failed]
        if os.path.exists(path) and os.path.getsize(path) > 0:
            with open(path, "rb") as fh:
                tb = pickle.load(fh)
            raise ProcessRaisedException(f"\n\n-- Process {failed} terminated with the following error:\n{tb}", failed, p.pid)
        code = p.exitcode
        if code < 0:
            try:
                name = signal.Signals(-code).name
            except ValueError:
                name = f"<signal {-code}>"
            raise ProcessExitedException(f"process {failed} terminated with signal {name}", failed, p.pid, code, name)
        raise ProcessExitedException(f"process {failed} terminated with exit code {code}", failed, p.pid, code)


def spawn(fn: Callable, args: Sequence = (), nprocs: Optional[int] = None, join: bool = True, start_method: str = "spawn"):
    """Start one process per device running ``fn(i, *args)`` and (with ``join=True``) wait for them - the launcher of
    ``multigpu.py`` (reference: ``mp.spawn(main, args=(...), nprocs=world_size)``, multigpu.py:263).

    Semantics of ``torch.multiprocessing.spawn`` (spawn.py:79-96, 118-211), implemented here on ``multiprocessing``: the
    child's exception travels to the parent as a pickled traceback and is re-raised as ``ProcessRaisedException``; a child
    that dies without raising gives ``ProcessExitedException`` with its exit code or signal; in both cases the surviving
    ranks are terminated (SIGTERM, then SIGKILL after 10 s) instead of being left waiting in a collective; children die
    with the parent (PR_SET_PDEATHSIG)."""
    import multiprocessing
    import tempfile
    if nprocs is None:
        nprocs = torch.cuda.device_count() if torch.cuda.is_available() else 1
    ctx = multiprocessing.get_context(start_method)
    tmp = tempfile.mkdtemp(prefix="ddpx_spawn_")
    procs, errs = [], []
    for i in range(nprocs):
        err = os.path.join(tmp, f"rank{i}.err")
        p = ctx.Process(target=_child_main, args=(fn, i, tuple(args), err), daemon=False)
        p.start()
        procs.append(p)
        errs.append(err)
    context = SpawnContext(procs, errs)
    if not join:
        return context
    try:
        while not context.join():
            pass
    finally:
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)
    return None
