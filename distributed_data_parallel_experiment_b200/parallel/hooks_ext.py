"""Loader of the C++ autograd-hook extension (``lib/ddpx_autograd.so``, source csrc/torch_ext/autograd_hooks.cpp).

The extension gives the reducer what torch DDP's c10d::Reducer has (reducer.hpp:219-224): post-hooks on the parameters'
gradient accumulators that run in C++ on the autograd thread, so no Python executes per gradient.  It is built in-tree by
``build.build_torch_ext`` (g++ against the installed torch headers); when it is missing it is built on first use, and when
that fails the reducer keeps its Python hooks and says so once.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import threading

from .. import build as _build

_lock = threading.Lock()
_mod = None
_tried = False


def module():
    """The extension module, or None when it cannot be built / imported on this host."""
    global _mod, _tried
    if _mod is not None or _tried:
        return _mod
    with _lock:
        if _mod is not None or _tried:
            return _mod
        _tried = True
        if os.environ.get("DDPX_NATIVE_HOOKS", "1") in ("0", ""):
            return None
        try:
            path = _build.EXT_PATH
            if not path.exists() and os.environ.get("DDPX_AUTOBUILD", "1") == "1":
                _build.build_torch_ext(verbose=False)
            import torch  # noqa: F401  (libtorch must be loaded before the extension)
            spec = importlib.util.spec_from_file_location("ddpx_autograd", str(path))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            _mod = mod
        except Exception as e:  # pragma: no cover - depends on the host toolchain
            print(f"[ddpx] C++ autograd hooks unavailable ({type(e).__name__}: {e}); the reducer uses Python hooks",
                  file=sys.stderr)
            _mod = None
    return _mod
