"""B200-native data-parallel training framework.

Same public surface as UnchartedWhispers/Distributed-Data-Parallel-Experiment (``Trainer``, the
``singlegpu.py`` / ``multigpu.py`` entrypoints, ``ddp_setup``/spawn, sampler sharding with ``set_epoch``,
rank-0 ``checkpoint.pt`` snapshots) on top of hand-written sm_100a kernels and a peer-to-peer NVLink
gradient reducer.  See DESIGN.md for the architecture and SURVEY.md for the reference analysis.
"""
from __future__ import annotations

from . import _lib  # noqa: F401
from . import config, metrics, data, checkpoint, launch  # noqa: F401
from . import ops, nn, models, parallel  # noqa: F401
from . import optim, recipes  # noqa: F401
from .trainer import Trainer  # noqa: F401
from .evaluate import evaluate  # noqa: F401
from .metrics import get_model_size, Byte, KiB, MiB, GiB  # noqa: F401
from .launch import ddp_setup, spawn, teardown  # noqa: F401
from .data import ShardedSampler, prepare_dataloader  # noqa: F401
from .parallel import DataParallel  # noqa: F401

__version__ = "0.1.0"
