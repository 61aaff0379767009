"""Environment / flag knobs (reference has none: every value is a literal).

Defaults reproduce the reference; every knob is optional.
"""
from __future__ import annotations

import os
from dataclasses import dataclass


def _env_int(name: str, default: int) -> int:
    v = os.environ.get(name)
    return int(v) if v not in (None, "") else default


def _env_str(name: str, default: str) -> str:
    v = os.environ.get(name)
    return v if v not in (None, "") else default


@dataclass
class RuntimeConfig:
    master_addr: str = "localhost"      # multigpu.py:30
    master_port: int = 12355            # multigpu.py:31
    backend: str = "auto"               # auto | p2p | nccl | gloo
    allreduce_algo: str = "auto"        # auto | oneshot | twoshot | nvls | nccl
    bucket_cap_mb: float = 25.0         # torch DDP default (distributed.py:831-833)
    first_bucket_mb: float = 1.0
    comm_ctas: int = 64                 # CTAs of a communication kernel (64 x 512 threads measured best, 2 and 8 GPUs)
    fuse_optimizer: bool = True
    broadcast_buffers: bool = True      # DataParallel default (torch DDP default); DDPX_BROADCAST_BUFFERS=0 for diagnosis
    early_buffer_sync: bool = True      # buffer broadcast of the next forward runs on the comm stream during backward
    split_tail_bytes: int = 65536       # P2P: the parameters that become ready last (<= this many bytes) get a bucket of their own
    wait_timeout_s: float = 600.0       # device-side flag waits and the control-plane store give up after this long (torch: 10 min)
    checkpoint_path: str = "checkpoint.pt"   # multigpu.py:111
    metrics_jsonl: str = ""

    @staticmethod
    def from_env() -> "RuntimeConfig":
        c = RuntimeConfig()
        c.master_addr = _env_str("DDPX_MASTER_ADDR", c.master_addr)
        c.master_port = _env_int("DDPX_MASTER_PORT", c.master_port)
        c.backend = _env_str("DDPX_BACKEND", c.backend)
        c.allreduce_algo = _env_str("DDPX_ALLREDUCE_ALGO", c.allreduce_algo)
        c.bucket_cap_mb = float(_env_str("DDPX_BUCKET_CAP_MB", str(c.bucket_cap_mb)))
        c.comm_ctas = _env_int("DDPX_COMM_CTAS", c.comm_ctas)
        c.fuse_optimizer = _env_int("DDPX_FUSE_OPTIMIZER", 1) == 1
        c.broadcast_buffers = _env_int("DDPX_BROADCAST_BUFFERS", 1) == 1
        c.early_buffer_sync = _env_int("DDPX_EARLY_BUFFER_SYNC", 1) == 1
        c.split_tail_bytes = _env_int("DDPX_SPLIT_TAIL_BYTES", c.split_tail_bytes)
        c.wait_timeout_s = float(_env_str("DDPX_WAIT_TIMEOUT_S", str(c.wait_timeout_s)))
        c.checkpoint_path = _env_str("DDPX_CHECKPOINT_PATH", c.checkpoint_path)
        c.metrics_jsonl = _env_str("DDPX_METRICS_JSONL", c.metrics_jsonl)
        return c


_cfg: RuntimeConfig | None = None


def get() -> RuntimeConfig:
    global _cfg
    if _cfg is None:
        _cfg = RuntimeConfig.from_env()
    return _cfg


def reset() -> None:
    global _cfg
    _cfg = None
