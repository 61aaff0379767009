"""Python face of the native control-plane store (csrc/runtime/store.cpp).

``NativeStore(host, port, rank, world_size)``: rank 0 starts the server thread inside its own process and everybody
(rank 0 included) connects as a client - the arrangement of torch's ``TCPStore`` that the reference gets from
``init_process_group`` (multigpu.py:30-32).  On top of set / get / add it offers the two collectives the bootstrap
needs: ``barrier`` and ``all_gather`` of small byte strings (CUDA-IPC handles, bucket orders).
"""
from __future__ import annotations

import ctypes
import pickle
from typing import Any, List, Optional

from .. import _lib


def _fns():
    lib = _lib.lib()
    lib.ddpx_store_server_start.restype = ctypes.c_void_p
    lib.ddpx_store_server_start.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    lib.ddpx_store_server_start_on.restype = ctypes.c_void_p
    lib.ddpx_store_server_start_on.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    lib.ddpx_store_server_stop.argtypes = [ctypes.c_void_p]
    lib.ddpx_store_connect.restype = ctypes.c_void_p
    lib.ddpx_store_connect.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int]
    lib.ddpx_store_close.argtypes = [ctypes.c_void_p]
    for name in ("ddpx_store_set", "ddpx_store_get", "ddpx_store_add", "ddpx_store_delete", "ddpx_store_num_keys"):
        getattr(lib, name).restype = ctypes.c_longlong
    lib.ddpx_store_set.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_longlong]
    lib.ddpx_store_get.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int]
    lib.ddpx_store_add.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_longlong]
    lib.ddpx_store_delete.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
    lib.ddpx_store_num_keys.argtypes = [ctypes.c_void_p]
    return lib


class StoreError(RuntimeError):
    pass


def _is_ipv4(h: str) -> bool:
    parts = h.split(".")
    return len(parts) == 4 and all(p.isdigit() and 0 <= int(p) <= 255 for p in parts)


class NativeStore:
    def __init__(self, host: str, port: int, rank: int, world_size: int, timeout_s: float = 60.0,
                 is_server: Optional[bool] = None):
        self._lib = _fns()
        self.rank, self.world_size, self.timeout_ms = rank, world_size, int(timeout_s * 1000)
        self._server = None
        self._gen = {}
        self.port = port
        if is_server if is_server is not None else rank == 0:
            bound = ctypes.c_int(0)
            # listen only on the job's address (the loopback for the single-host jobs of this framework): the store is
            # unauthenticated and values are unpickled by the clients, so it must not be reachable from elsewhere
            bind = host if _is_ipv4(host) else "127.0.0.1"
            self._server = self._lib.ddpx_store_server_start_on(bind.encode(), port, ctypes.byref(bound))
            if not self._server:
                raise StoreError(f"store: cannot listen on port {port}")
            self.port = bound.value
        self._client = self._lib.ddpx_store_connect(host.encode(), self.port, self.timeout_ms)
        if not self._client:
            self.close()
            raise StoreError(f"store: cannot connect to {host}:{self.port} within {timeout_s} s")

    # ---- primitives
    def set(self, key: str, value: bytes) -> None:
        if self._lib.ddpx_store_set(self._client, key.encode(), value, len(value)) < 0:
            raise StoreError(f"store.set({key!r}) failed")

    def get(self, key: str, timeout_s: Optional[float] = None) -> bytes:
        t = self.timeout_ms if timeout_s is None else int(timeout_s * 1000)
        cap = 4096
        while True:
            buf = ctypes.create_string_buffer(cap)
            n = self._lib.ddpx_store_get(self._client, key.encode(), buf, cap, t)
            if n >= 0:
                return buf.raw[:n]
            if n <= -16:
                cap = -16 - n
                continue
            if n == -1:
                raise TimeoutError(f"store.get({key!r}): not set within {t} ms")
            raise StoreError(f"store.get({key!r}): connection lost")

    def add(self, key: str, delta: int = 1) -> int:
        v = self._lib.ddpx_store_add(self._client, key.encode(), delta)
        if v == -2:
            raise StoreError(f"store.add({key!r}): connection lost")
        return int(v)

    def delete(self, key: str) -> bool:
        return self._lib.ddpx_store_delete(self._client, key.encode()) > 0

    def num_keys(self) -> int:
        return int(self._lib.ddpx_store_num_keys(self._client))

    # ---- collectives for the bootstrap
    def _next(self, name: str) -> str:
        g = self._gen.get(name, 0)
        self._gen[name] = g + 1
        return f"{name}/{g}"

    def _retire(self, name: str, keys) -> None:
        """Rank 0 deletes the keys of the generation before the one that just completed: every rank has entered the current
        generation, so nobody reads the previous one any more (a long job must not grow the server's map without bound)."""
        g = self._gen[name] - 2
        if self.rank == 0 and g >= 0:
            for k in keys(f"{name}/{g}"):
                self.delete(k)

    def barrier(self, name: str = "barrier") -> None:
        tag = self._next(name)
        n = self.add(tag + "/arrived", 1)
        if n == self.world_size:
            self.set(tag + "/go", b"1")
        self.get(tag + "/go")
        self._retire(name, lambda t: (t + "/arrived", t + "/go"))

    def all_gather(self, payload: bytes, name: str = "gather") -> List[bytes]:
        tag = self._next(name)
        self.set(f"{tag}/{self.rank}", payload)
        out = [self.get(f"{tag}/{r}") for r in range(self.world_size)]
        self._retire(name, lambda t: [f"{t}/{r}" for r in range(self.world_size)])
        return out

    def all_gather_object(self, obj: Any, name: str = "gather_obj") -> List[Any]:
        return [pickle.loads(b) for b in self.all_gather(pickle.dumps(obj), name)]

    def broadcast_object(self, obj: Any, src: int = 0, name: str = "bcast_obj") -> Any:
        tag = self._next(name)
        if self.rank == src:
            self.set(tag, pickle.dumps(obj))
        return pickle.loads(self.get(tag))

    def close(self, linger_s: float = 5.0) -> None:
        """Clients sign off; the hosting rank keeps the server up until every other rank has signed off (or ``linger_s``
        passed), so a fast rank 0 cannot pull the store from under a peer that is still reading its last key."""
        if getattr(self, "_client", None):
            try:
                if self._server is None:
                    self.add("__signed_off", 1)
                elif self.world_size > 1:
                    import time
                    deadline = time.monotonic() + linger_s
                    while self.add("__signed_off", 0) < self.world_size - 1 and time.monotonic() < deadline:
                        time.sleep(0.005)
            except Exception:
                pass
            self._lib.ddpx_store_close(self._client)
            self._client = None
        if self._server:
            self._lib.ddpx_store_server_stop(self._server)
            self._server = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
