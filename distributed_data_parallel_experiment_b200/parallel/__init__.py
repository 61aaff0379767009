"""Data-parallel runtime: bucket assignment, gradient reducer, communication backends, DataParallel wrapper."""
from .buckets import compute_bucket_assignment_by_size, DEFAULT_FIRST_BUCKET_BYTES  # noqa: F401
from .comm import create as create_comm, TorchComm, NullComm  # noqa: F401
from .reducer import Reducer  # noqa: F401
from .ddp import DataParallel  # noqa: F401

DistributedDataParallel = DataParallel
