"""Bucket assignment (thin wrapper over the C++ core, csrc/runtime/reducer_core.cpp).

Same greedy rule as ``torch.distributed._compute_bucket_assignment_by_size`` (c10d/reducer.hpp:592-598), which torch
DDP calls for the reference at distributed.py:1240: for VGG it yields the 9,461,800 / 27,148,288 / 303,360-byte
buckets of SURVEY.md Appendix B.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import torch

from .. import _lib

DEFAULT_FIRST_BUCKET_BYTES = 1024 * 1024          # dist._DEFAULT_FIRST_BUCKET_BYTES
DEFAULT_BUCKET_CAP_BYTES = 25 * 1024 * 1024       # bucket_cap_mb = 25

_DTYPE_IDS = {}


def _key(t: torch.Tensor) -> int:
    k = (t.dtype, str(t.device))
    if k not in _DTYPE_IDS:
        _DTYPE_IDS[k] = len(_DTYPE_IDS)
    return _DTYPE_IDS[k]


def compute_bucket_assignment_by_size(tensors: Sequence[torch.Tensor], bucket_size_limits: Sequence[int],
                                      tensor_indices: Optional[Sequence[int]] = None) -> Tuple[List[List[int]], List[int]]:
    """Returns (bucket_indices, per_bucket_size_limits).  ``tensors`` are visited in the given order; when
    ``tensor_indices`` is given, entry i of ``tensors`` is reported under index ``tensor_indices[i]`` and the bucket order
    is the visiting order, otherwise buckets are sorted by their smallest index."""
    n = len(tensors)
    if n == 0:
        return [], []
    lib = _lib.lib()
    nbytes = (ctypes.c_longlong * n)(*[t.numel() * t.element_size() for t in tensors])
    keys = (ctypes.c_int * n)(*[_key(t) for t in tensors])
    lims = [min(int(x), 2 ** 62) for x in bucket_size_limits]
    limits = (ctypes.c_longlong * len(lims))(*lims)
    out_b = (ctypes.c_int * n)()
    out_l = (ctypes.c_longlong * n)()
    out_n = ctypes.c_int(0)
    rc = lib.ddpx_bucket_assign(n, nbytes, keys, len(lims), limits, 0 if tensor_indices is not None else 1, out_b, out_l,
                                ctypes.byref(out_n))
    _lib.check(rc, "ddpx_bucket_assign")
    buckets: List[List[int]] = [[] for _ in range(out_n.value)]
    for pos in range(n):
        idx = tensor_indices[pos] if tensor_indices is not None else pos
        buckets[out_b[pos]].append(idx)
    return buckets, [int(out_l[b]) for b in range(out_n.value)]
