"""ctypes binding to ``libddpx_kernels.so`` (C ABI in ``include/ddpx_kernels.h``).

The native library is the only compute path on a CUDA device: if it cannot be loaded there, the
ops raise instead of silently falling back to torch kernels. On a CPU-only host (unit tests, gloo
runs) the ops use plain torch math and this module is never asked for the library.
"""
from __future__ import annotations

import ctypes
import os
import threading
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "lib" / "libddpx_kernels.so"

_lock = threading.Lock()
_lib = None
_launches = 0  # number of native kernel-launching calls made through this module

ERR_NAMES = {0: "OK", 1: "ERR_ARG", 2: "ERR_ALIGN", 3: "ERR_SHAPE", 4: "ERR_DRIVER", 5: "ERR_TMAP",
             6: "ERR_LAUNCH", 7: "ERR_LIMIT"}


class NativeLibraryError(RuntimeError):
    pass


def available() -> bool:
    return LIB_PATH.exists()


def lib() -> ctypes.CDLL:
    """Load (once) and return the native library; raises loudly when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not LIB_PATH.exists():
                if os.environ.get("DDPX_AUTOBUILD", "1") == "1":
                    from . import build as _build
                    _build.build(verbose=False)
                if not LIB_PATH.exists():
                    raise NativeLibraryError(
                        f"{LIB_PATH} is missing: run `python -m distributed_data_parallel_experiment_b200.build` "
                        "(the CUDA path never falls back to torch kernels)")
            _lib = ctypes.CDLL(str(LIB_PATH), mode=ctypes.RTLD_GLOBAL)
    return _lib


def count_launch(n: int = 1) -> None:
    global _launches
    _launches += n


def launches() -> int:
    return _launches


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise NativeLibraryError(f"{what} failed: {ERR_NAMES.get(rc, rc)}")


def ptr(t) -> ctypes.c_void_p:
    """Device pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr() -> ctypes.c_void_p:
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
