"""Peer-to-peer NVLink communication backend (Python side of csrc/runtime/p2p.cu).

One process per GPU.  ``P2PComm`` owns
* a signal pad and a double-slot staging buffer in *symmetric* memory (same allocation sequence on every rank, every
  peer's copy mapped locally through CUDA IPC handles exchanged over the torch.distributed control plane);
* a high-priority communication stream on which the fused all-reduce(+optimizer) kernels run so they overlap backward;
* two monotonically increasing generation counters (all-reduce, broadcast) that stand in for barriers.

``FusedUpdate`` binds a stock ``torch.optim.SGD`` / ``AdamW`` to the reducer: each bucket's kernel averages the gradients
over peer memory AND applies the optimizer update to the parameters of that bucket, after which ``optimizer.step()`` is
a no-op for that iteration.
"""
from __future__ import annotations

import ctypes
import math
from typing import List, Optional, Sequence

import torch

from .. import _lib, config
from .comm import CommBase

ONESHOT_MAX_BYTES = 256 * 1024
# Algorithm selection ("auto") comes from a MEASURED table (parallel/algo_table.json, written by
# `bench/allreduce_sweep.py --emit-table` from the fused sweep): per world size the largest bucket for which the one-shot
# kernel (every rank reduces the whole bucket from peer memory and updates its replica; one barrier less) beats two-shot,
# and the smallest bucket from which the in-switch reduction (NVLS) beats two-shot.  2 GPUs: one-shot+SGD 24.4 us vs
# two-shot+SGD 24.5 us at 1 MB, 34.7 vs 26.6 us at 4 MB; 8 GPUs: NVLS 585 vs 547 GB/s at 64 MB, 755 vs 621 at 256 MB.


def _load_table():
    import json
    from pathlib import Path
    try:
        raw = json.loads((Path(__file__).with_name("algo_table.json")).read_text())
        return {int(k): v for k, v in raw.items() if not k.startswith("_")}
    except Exception:
        return {}


ALGO_TABLE = _load_table()


def _table_entry(world_size: int) -> dict:
    if world_size in ALGO_TABLE:
        return ALGO_TABLE[world_size]
    smaller = [w for w in ALGO_TABLE if w <= world_size]
    return ALGO_TABLE[max(smaller)] if smaller else {}


def oneshot_max_bytes(world_size: int) -> int:
    return int(_table_entry(world_size).get("oneshot_max_bytes") or ONESHOT_MAX_BYTES)


def nvls_min_bytes(world_size: int):
    """Smallest bucket that goes to multicast-bound memory under ``auto`` (None: never on this world size)."""
    return _table_entry(world_size).get("nvls_min_bytes")


def select_mode(algo: str, nbytes: int, world_size: int, nvls: bool = False) -> int:
    """Kernel flavour for a bucket of `nbytes`: 1 one-shot up to the measured threshold (or when forced), else two-shot (0),
    or the in-switch reduction (2) when the bucket lives in multicast-bound memory."""
    if algo == "oneshot":
        return 1
    if algo == "auto" and nbytes <= oneshot_max_bytes(world_size):
        return 1
    return 2 if nvls else 0
_MAX_SLOTS = 128


class _RawCuda:
    """Expose a raw device allocation to torch through __cuda_array_interface__ (no copy, no ownership)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


class SymBuf:
    __slots__ = ("local", "ptrs", "nbytes", "tensor", "mc_ptr")

    def __init__(self, local, ptrs, nbytes, tensor):
        self.local, self.ptrs, self.nbytes, self.tensor = local, ptrs, nbytes, tensor
        self.mc_ptr = 0          # NVLS: multicast mapping of the same buffer (parallel/nvls.py), 0 = none


class _EventHandle:
    def __init__(self, ev):
        self.ev = ev

    def wait(self):
        torch.cuda.current_stream().wait_event(self.ev)


def _cvoidp_array(vals: Sequence[int]):
    arr = (ctypes.c_void_p * max(1, len(vals)))()
    for i, v in enumerate(vals):
        arr[i] = v
    return arr


class P2PComm(CommBase):
    name = "p2p"
    supports_fused_update = True
    needs_double_buffer = False

    def __init__(self, staging_mb: int = 64):
        super().__init__()
        if not torch.cuda.is_available():
            raise RuntimeError("P2PComm needs CUDA")
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.lib = _lib.lib()
        self.cfg = config.get()
        self._bufs: List[SymBuf] = []
        self._by_ptr = {}
        if self.world_size > 8:
            raise RuntimeError("P2PComm supports up to 8 ranks (one NVSwitch domain)")
        # device-side flag waits give up after config.wait_timeout_s (DDPX_WAIT_TIMEOUT_S)
        _lib.check(self.lib.ddpx_p2p_set_timeout(ctypes.c_double(float(self.cfg.wait_timeout_s))), "ddpx_p2p_set_timeout")
        self.store = self._open_store()
        self.nvls = self._open_nvls()
        self.pad = self._sym_alloc(4096)
        self.slot_bytes = staging_mb * 1024 * 1024
        self.staging = self._sym_alloc(2 * self.slot_bytes)
        # collective sequence numbers live in the native library (the native reducer launches all-reduces too); a new
        # backend object has fresh, zeroed signal pads, so the numbering restarts with it
        self.lib.ddpx_p2p_next_gen.restype = ctypes.c_uint
        self.lib.ddpx_p2p_next_bcast_gen.restype = ctypes.c_uint
        self.lib.ddpx_p2p_reset_gens()
        self._bcast_last = {}
        self.stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.ctas = int(self.cfg.comm_ctas)
        self._pads_arr = _cvoidp_array(self.pad.ptrs)
        self.barrier()

    # ------------------------------------------------------------------------------------------ control plane
    def _open_store(self):
        """Native TCP store (csrc/runtime/store.cpp) hosted by rank 0 for every host-side exchange of this backend:
        CUDA-IPC handles, bucket orders, barriers.  torch's process group is used once, to tell the peers the port.
        ``DDPX_NATIVE_STORE=0`` keeps everything on torch.distributed object collectives."""
        import os
        if self.world_size == 1 or os.environ.get("DDPX_NATIVE_STORE", "1") in ("0", ""):
            return None
        from .store import NativeStore
        host = os.environ.get("MASTER_ADDR", "127.0.0.1")
        if host == "localhost":
            host = "127.0.0.1"
        timeout = max(30.0, float(self.cfg.wait_timeout_s))
        if self.rank == 0:
            st = NativeStore(host, 0, 0, self.world_size, timeout_s=timeout, is_server=True)
            super().broadcast_object(st.port, src=0)
            return st
        port = super().broadcast_object(None, src=0)
        return NativeStore(host, int(port), self.rank, self.world_size, timeout_s=timeout, is_server=False)

    def _open_nvls(self):
        """``allreduce_algo = nvls``: gradient buckets come from the VMM / multicast allocator (parallel/nvls.py) and the
        shard reduction of the all-reduce kernel runs in the NVSwitch.  Needs multicast support on every rank; otherwise
        the backend says so once and stays on the two-shot path."""
        want = self.cfg.allreduce_algo == "nvls" or (self.cfg.allreduce_algo == "auto"
                                                     and nvls_min_bytes(self.world_size) is not None)
        if not want or self.world_size == 1 or self.store is None:
            return None
        from . import nvls
        ok = nvls.supported(self.device.index)
        if not all(self.all_gather_object(bool(ok))):
            if self.rank == 0 and self.cfg.allreduce_algo == "nvls":
                print("[ddpx] NVLS multicast is not available on every rank: using the two-shot all-reduce")
            return None
        torch.zeros(1, device=self.device)                  # the primary context must exist before driver-API calls
        return nvls.NvlsAllocator(self.rank, self.world_size, self.device.index, self)

    def all_gather_object(self, obj):
        if self.store is not None:
            return self.store.all_gather_object(obj)
        return super().all_gather_object(obj)

    def broadcast_object(self, obj, src: int = 0):
        if self.store is not None:
            return self.store.broadcast_object(obj, src=src)
        return super().broadcast_object(obj, src=src)

    def barrier(self):
        if self.store is not None:
            torch.cuda.synchronize(self.device)          # a host barrier after the device work it is meant to fence
            self.store.barrier()
        else:
            super().barrier()

    # ------------------------------------------------------------------------------------------ symmetric memory
    def _sym_alloc(self, nbytes: int) -> SymBuf:
        nbytes = (int(nbytes) + 255) // 256 * 256
        p = ctypes.c_void_p()
        _lib.check(self.lib.ddpx_p2p_malloc(ctypes.byref(p), ctypes.c_longlong(nbytes)), "ddpx_p2p_malloc")
        handle = (ctypes.c_ubyte * 64)()
        ptrs = [0] * self.world_size
        ptrs[self.rank] = p.value
        if self.world_size > 1:
            _lib.check(self.lib.ddpx_p2p_export(p, handle), "ddpx_p2p_export")
            handles = self.all_gather_object(bytes(handle))
            for r, h in enumerate(handles):
                if r == self.rank:
                    continue
                q = ctypes.c_void_p()
                hb = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                _lib.check(self.lib.ddpx_p2p_import(hb, ctypes.byref(q)), f"ddpx_p2p_import(rank {r})")
                ptrs[r] = q.value
        t = torch.as_tensor(_RawCuda(p.value, nbytes), device=self.device)
        sb = SymBuf(p.value, ptrs, nbytes, t)
        self._bufs.append(sb)
        self._by_ptr[p.value] = sb
        return sb

    def alloc(self, numel: int, dtype: torch.dtype, device=None) -> torch.Tensor:
        """Collective: every rank must call with the same sizes in the same order."""
        es = torch.empty((), dtype=dtype).element_size()
        # multicast-bound memory for every buffer under "nvls", and under "auto" for the sizes the measured table gives it
        use_nvls = self.nvls is not None and (self.cfg.allreduce_algo == "nvls"
                                              or numel * es >= (nvls_min_bytes(self.world_size) or (1 << 62)))
        if use_nvls:
            nb = self.nvls.alloc(numel * es)
            sb = SymBuf(nb.local, nb.ptrs, nb.nbytes, nb.tensor)
            sb.mc_ptr = nb.mc_ptr
            self._nvls_bufs = getattr(self, "_nvls_bufs", []) + [sb]
            self._by_ptr[nb.local] = sb
            return sb.tensor[: numel * es].view(dtype)
        sb = self._sym_alloc(numel * es)
        return sb.tensor[: numel * es].view(dtype)

    def free(self, t: torch.Tensor) -> None:
        """Collective: release a symmetric buffer obtained from ``alloc`` (every rank calls it for the same buffer, e.g. the
        whole-model first-iteration bucket once the reducer has rebuilt its buckets).  Peers' mappings of the buffer are
        closed before anybody frees its own copy.  NVLS-backed buffers stay until ``close()``."""
        sb = self._by_ptr.get(t.data_ptr())
        if sb is None or sb.mc_ptr or sb not in self._bufs:
            return
        self.barrier()                                   # nobody still has a kernel reading it (barrier syncs the device)
        for r, q in enumerate(sb.ptrs):
            if r != self.rank and q:
                self.lib.ddpx_p2p_unimport(ctypes.c_void_p(q))
        self.barrier()
        self._bufs.remove(sb)
        del self._by_ptr[sb.local]
        sb.tensor = None
        self.lib.ddpx_p2p_free(ctypes.c_void_p(sb.local))

    def symbuf_of(self, t: torch.Tensor) -> SymBuf:
        sb = self._by_ptr.get(t.data_ptr())
        if sb is None:
            raise RuntimeError("tensor is not the base of a symmetric allocation")
        return sb

    # ------------------------------------------------------------------------------------------ all-reduce
    def _next_ar_gen(self) -> int:
        return int(self.lib.ddpx_p2p_next_gen())

    def launch_allreduce(self, flat: torch.Tensor, mode: int = 0, opt: int = 0, hyper=None, nesterov=False, first_step=False,
                         slots=None) -> _EventHandle:
        """Enqueue the fused kernel for one symmetric flat buffer on the communication stream."""
        sb = self.symbuf_of(flat)
        is_bf16 = 1 if flat.dtype == torch.bfloat16 else 0
        n = flat.numel()
        gen = self._next_ar_gen()
        cur = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(cur)
        self.stream.wait_event(ready)
        hy = (ctypes.c_float * 9)(*(hyper or [0.0] * 9))
        if slots:
            params, s0, s1, offs, nums = slots
            ns = len(params)
        else:
            params = s0 = s1 = offs = nums = None
            ns = 0
        # small buckets are pure latency: a few CTAs reach the grid-wide hand-over sooner than 64 do
        ctas = max(1, min(self.ctas, (n * flat.element_size()) >> 15))
        tail = (is_bf16, opt, hy, int(nesterov), int(first_step), ns,
                _cvoidp_array(params) if ns else None, _cvoidp_array(s0) if ns and s0 else None,
                _cvoidp_array(s1) if ns and s1 else None,
                (ctypes.c_longlong * max(1, ns))(*(offs or [0])), (ctypes.c_longlong * max(1, ns))(*(nums or [0])),
                ctas, ctypes.c_void_p(self.stream.cuda_stream))
        mc_ptr = getattr(sb, "mc_ptr", 0)
        if mode == 2 and not mc_ptr:
            mode = 0                                        # not a multicast-bound buffer: plain two-shot
        if mode == 2:
            rc = self.lib.ddpx_p2p_allreduce_nvls(
                self.rank, self.world_size, _cvoidp_array(sb.ptrs), ctypes.c_void_p(mc_ptr), self._pads_arr,
                ctypes.c_uint(gen), ctypes.c_longlong(n), *tail)
        else:
            rc = self.lib.ddpx_p2p_allreduce_fused(
                self.rank, self.world_size, _cvoidp_array(sb.ptrs), self._pads_arr, ctypes.c_uint(gen), ctypes.c_longlong(n),
                tail[0], mode, *tail[1:])
        _lib.count_launch(1)
        _lib.check(rc, "ddpx_p2p_allreduce_fused")
        done = torch.cuda.Event()
        done.record(self.stream)
        return _EventHandle(done)

    def all_reduce_sum_(self, flat: torch.Tensor, async_op: bool = True):
        h = self.launch_allreduce(flat, mode=2 if getattr(self.symbuf_of(flat), "mc_ptr", 0) else 0, opt=0)
        if not async_op:
            h.wait()
        return h

    # ------------------------------------------------------------------------------------------ broadcast
    def broadcast_tensors_(self, tensors: Sequence[torch.Tensor], src: int = 0):
        """In-place broadcast of ``src``'s values of ``tensors`` (any dtypes, contiguous) to every rank."""
        if self.world_size == 1:
            return
        pieces = []        # (ptr, nbytes)
        for t in tensors:
            if not t.is_contiguous():
                raise RuntimeError("broadcast_tensors_: tensors must be contiguous")
            ptr, nb = t.data_ptr(), t.numel() * t.element_size()
            while nb > 0:               # split anything larger than a staging slot
                take = min(nb, self.slot_bytes)
                pieces.append((ptr, take))
                ptr += take
                nb -= take
        chunk, used = [], 0
        for ptr, nb in pieces:
            pad = (nb + 15) // 16 * 16
            if chunk and (used + pad > self.slot_bytes or len(chunk) >= _MAX_SLOTS):
                self._bcast_chunk(chunk, src)
                chunk, used = [], 0
            chunk.append((ptr, used, nb))
            used += pad
        if chunk:
            self._bcast_chunk(chunk, src)

    def _bcast_chunk(self, chunk, src):
        gen = int(self.lib.ddpx_p2p_next_bcast_gen())
        n = len(chunk)
        slot = gen & 1
        # the source may refill a staging slot only after every peer consumed what it last put there: every rank tracks,
        # per (source, slot), the generation of that last use (identical on all ranks: same calls in the same order)
        prev = self._bcast_last.get((src, slot), 0)
        self._bcast_last[(src, slot)] = gen
        staging = self.staging.ptrs[src] + slot * self.slot_bytes
        ptrs = _cvoidp_array([c[0] for c in chunk])
        offs = (ctypes.c_longlong * n)(*[c[1] for c in chunk])
        nbs = (ctypes.c_longlong * n)(*[c[2] for c in chunk])
        total = sum(c[2] for c in chunk)
        ctas = max(1, min(self.ctas, (total + (1 << 16) - 1) >> 16))
        rc = self.lib.ddpx_p2p_broadcast(self.rank, self.world_size, self._pads_arr, src, ctypes.c_uint(gen),
                                         ctypes.c_uint(prev), ctypes.c_void_p(staging), n, ptrs, offs, nbs, ctas,
                                         ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.count_launch(1)
        _lib.check(rc, "ddpx_p2p_broadcast")

    def broadcast_(self, flat: torch.Tensor, src: int = 0):
        self.broadcast_tensors_([flat], src)

    # ------------------------------------------------------------------------------------------ misc
    def error_word(self) -> int:
        fn = self.lib.ddpx_p2p_read_error
        fn.restype = ctypes.c_uint
        return int(fn(ctypes.c_void_p(self.pad.local)))

    def close(self):
        try:
            torch.cuda.synchronize()
            self.barrier()
        except Exception:
            pass
        for sb in self._bufs:
            for r, q in enumerate(sb.ptrs):
                if r != self.rank and q:
                    self.lib.ddpx_p2p_unimport(ctypes.c_void_p(q))
        try:
            self.barrier()
        except Exception:
            pass
        for sb in self._bufs:
            sb.tensor = None
            self.lib.ddpx_p2p_free(ctypes.c_void_p(sb.local))
        self._bufs = []
        self._by_ptr = {}
        if getattr(self, "nvls", None) is not None:
            try:
                for sb in getattr(self, "_nvls_bufs", []):
                    sb.tensor = None
                self._nvls_bufs = []
                self.nvls.close()
            except Exception as e:                      # teardown must not mask the caller's own exception
                print(f"[ddpx] NVLS teardown: {e!r}")
            self.nvls = None
        if getattr(self, "store", None) is not None:
            self.store.close()
            self.store = None


class FusedUpdate:
    """Per-bucket fused all-reduce + optimizer step on the P2P backend."""

    def __init__(self, comm: P2PComm, params, optimizer, kind: str):
        self.comm = comm
        self.params = list(params)
        self.optimizer = optimizer
        self.kind = kind                      # "sgd" | "adamw"
        self.group_of = {}
        for g in optimizer.param_groups:
            for p in g["params"]:
                self.group_of[id(p)] = g
        self.tables = {}
        self.modes = {}
        self._reducer = None
        self._sharded = False                 # True once a shard update ran since the last gather_state()
        self._sync_progress_from_state()
        # a snapshot loaded later (checkpoint.load_snapshot -> optimizer.load_state_dict) replaces the state tensors: the
        # slot tables hold raw pointers into them, so they are rebuilt and the step bookkeeping re-derived right after
        if hasattr(optimizer, "register_load_state_dict_post_hook"):
            self._hook = optimizer.register_load_state_dict_post_hook(lambda opt: self.on_state_loaded())
        if hasattr(optimizer, "register_state_dict_pre_hook"):
            self._sd_hook = optimizer.register_state_dict_pre_hook(lambda opt: self.gather_state())

    def _sync_progress_from_state(self):
        """Derive `first step?` and the AdamW step count from the optimizer's own state, so that wrapping a model after a
        snapshot was loaded (or loading one later) continues where the snapshot stopped: momentum buffers that exist mean
        SGD's first-step rule (buf = g) is over; AdamW's bias correction continues from state['step']."""
        steps, has_momentum = 0, False
        for p in self.params:
            st = self.optimizer.state.get(p, {})
            if st.get("momentum_buffer") is not None:
                has_momentum = True
            if "step" in st:
                steps = max(steps, int(float(st["step"])))
        self.first_step_pending = (self.kind == "sgd") and not has_momentum
        self._steps = steps

    def on_state_loaded(self):
        self._sync_progress_from_state()
        self._sharded = False                 # a loaded state is whole on every rank
        if self._reducer is not None:
            self.on_buckets_installed(self._reducer)

    # ---- sharded optimizer state -------------------------------------------------------------------------------
    def shard_range(self, bucket_numel: int, dtype, rank: int):
        """Element range of `rank`'s shard of a bucket - the same arithmetic as allreduce_fused_kernel (csrc/runtime/p2p.cu)."""
        V = 4 if dtype == torch.float32 else 8
        nvec = bucket_numel // V
        per = (nvec + self.comm.world_size - 1) // self.comm.world_size
        v0 = min(per * rank, nvec)
        v1 = min(v0 + per, nvec)
        return v0 * V, v1 * V

    def _state_names(self):
        return ("momentum_buffer",) if self.kind == "sgd" else ("exp_avg", "exp_avg_sq")

    def gather_state(self):
        """Collective.  In two-shot / NVLS mode every rank updates only its shard of a bucket, so momentum / Adam moments
        are current only on their owner.  This re-replicates them: for every owner, its slices of the state tensors are
        broadcast to the other ranks.  Runs before ``optimizer.state_dict()`` (pre-hook) and before the buckets are
        rebuilt; every rank must take part."""
        comm = self.comm
        if comm.world_size == 1 or self._reducer is None or not self._sharded:
            return
        torch.cuda.synchronize()
        for b in self._reducer.buckets:
            if self.tables.get(b.index) is None or self.modes.get(b.index) not in (0, 2):
                continue
            for owner in range(comm.world_size):
                e0, e1 = self.shard_range(b.numel, b.dtype, owner)
                views = []
                for pid, off in zip(b.param_ids, b.offsets):
                    p = self.params[pid]
                    a, z = max(off, e0), min(off + p.numel(), e1)
                    if a >= z:
                        continue
                    st = self.optimizer.state[p]
                    for name in self._state_names():
                        t = st.get(name)
                        if t is not None:
                            views.append(t.view(-1)[a - off:z - off])
                if views:
                    comm.broadcast_tensors_(views, src=owner)
        torch.cuda.synchronize()
        self._sharded = False

    @staticmethod
    def try_create(comm, params, optimizer) -> Optional["FusedUpdate"]:
        from .. import optim as _optim
        if not _optim.supported(optimizer):
            return None
        params = list(params)
        owned = {id(p) for g in optimizer.param_groups for p in g["params"]}
        if not all(id(p) in owned for p in params):
            return None
        dts = {p.dtype for p in params}
        if len(dts) != 1 or next(iter(dts)) not in (torch.float32, torch.bfloat16):
            return None
        if not all(p.is_cuda and p.is_contiguous() for p in params):
            return None
        _optim.fuse(optimizer)
        if not getattr(optimizer, "_ddpx_fused", False):
            return None
        kind = "sgd" if type(optimizer) is torch.optim.SGD else "adamw"
        return FusedUpdate(comm, params, optimizer, kind)

    # state tensors live where optimizer.state_dict() expects them
    def _ensure_state(self, p):
        st = self.optimizer.state[p]
        if self.kind == "sgd":
            g = self.group_of[id(p)]
            if float(g["momentum"]) != 0.0 and st.get("momentum_buffer") is None:
                from ..ops import native
                st["momentum_buffer"] = native.zero_(torch.empty_like(p, memory_format=torch.preserve_format))
        else:
            if len(st) == 0:
                from ..ops import native
                st["step"] = torch.tensor(0.0, dtype=torch.float32)
                st["exp_avg"] = native.zero_(torch.empty_like(p, memory_format=torch.preserve_format))
                st["exp_avg_sq"] = native.zero_(torch.empty_like(p, memory_format=torch.preserve_format))
        return st

    def on_buckets_installed(self, reducer):
        self._reducer = reducer
        fresh = self.first_step_pending          # _ensure_state creates zero buffers: they do not end SGD's first step
        self.tables = {}
        for b in reducer.buckets:
            groups = {id(self.group_of[id(self.params[pid])]) for pid in b.param_ids}
            if len(b.param_ids) > _MAX_SLOTS or len(groups) != 1:
                self.tables[b.index] = None        # falls back to all-reduce only + optimizer.step()
                continue
            ps, s0, s1, offs, nums = [], [], [], [], []
            for pid, off in zip(b.param_ids, b.offsets):
                p = self.params[pid]
                st = self._ensure_state(p)
                ps.append(p.data_ptr())
                if self.kind == "sgd":
                    mb = st.get("momentum_buffer")
                    s0.append(mb.data_ptr() if mb is not None else 0)
                    s1.append(0)
                else:
                    s0.append(st["exp_avg"].data_ptr())
                    s1.append(st["exp_avg_sq"].data_ptr())
                offs.append(off)
                nums.append(p.numel())
            self.tables[b.index] = (ps, s0, s1, offs, nums)
        self.first_step_pending = fresh
        if hasattr(reducer, "on_slots_changed"):
            reducer.on_slots_changed()

    def _hyper(self, group):
        if self.kind == "sgd":
            return [float(group["lr"]), float(group["momentum"]), float(group["dampening"]), float(group["weight_decay"]),
                    0.0, 0.0, 0.0, 1.0, 1.0]
        b1, b2 = group["betas"]
        step = self._steps + 1
        return [float(group["lr"]), 0.0, 0.0, float(group["weight_decay"]), float(b1), float(b2), float(group["eps"]),
                1.0 - b1 ** step, math.sqrt(1.0 - b2 ** step)]

    def plan(self, bucket):
        """(mode, opt, hyper, nesterov, first_step, slot table) for this iteration's kernel of `bucket`.
        mode 0 two-shot, 1 one-shot, 2 NVLS; opt 0 none (the bucket is only averaged), 1 SGD, 2 AdamW."""
        tab = self.tables.get(bucket.index)
        if tab is None:
            self._unfused = True
            return 0, 0, None, False, False, None
        group = self.group_of[id(self.params[bucket.param_ids[0]])]
        nbytes = bucket.numel * (2 if bucket.dtype == torch.bfloat16 else 4)
        has_mc = bool(getattr(self.comm.symbuf_of(bucket.buffers[0]), "mc_ptr", 0))
        mode = select_mode(self.comm.cfg.allreduce_algo, nbytes, self.comm.world_size, nvls=has_mc)
        self.modes[bucket.index] = mode
        if mode in (0, 2) and self.comm.world_size > 1:
            self._sharded = True
        return (mode, 1 if self.kind == "sgd" else 2, self._hyper(group), bool(group.get("nesterov", False)),
                bool(self.first_step_pending), tab)

    def reduce_and_update(self, bucket, buf, parity):
        mode, opt, hyper, nesterov, first_step, tab = self.plan(bucket)
        if opt == 0:
            return self.comm.launch_allreduce(buf, mode=0, opt=0)
        return self.comm.launch_allreduce(buf, mode=mode, opt=opt, hyper=hyper, nesterov=nesterov, first_step=first_step,
                                          slots=tab)

    def after_backward(self, reducer):
        """Called once every bucket's kernel has been waited for: mark this iteration's optimizer step as done."""
        if any(v is None for v in self.tables.values()):
            # mixed: some buckets were only averaged; expose their gradients and let optimizer.step() handle them
            for b in reducer.buckets:
                if self.tables.get(b.index) is None:
                    for pid, v in zip(b.param_ids, b.views[0]):
                        self.params[pid].grad = v
                else:
                    for pid in b.param_ids:
                        self.params[pid].grad = None
            self.first_step_pending = False
            self._steps += 1
            if self.kind == "adamw":         # fused parameters took their step inside the kernel; the others count in step()
                for b in reducer.buckets:
                    if self.tables.get(b.index) is not None:
                        for pid in b.param_ids:
                            self.optimizer.state[self.params[pid]]["step"] += 1
            return
        self.first_step_pending = False
        self._steps += 1
        for b in reducer.buckets:
            # the fused kernels consume the gradients: a two-shot / NVLS bucket now holds updated PARAMETERS (shard
            # update), a one-shot bucket still holds this rank's share only.  Neither is a gradient: .grad is dropped
            # rather than left misleading (DDPX_FUSE_OPTIMIZER=0 keeps averaged gradients for inspection or clipping)
            for pid in b.param_ids:
                self.params[pid].grad = None
        if self.kind == "adamw":
            for p in self.params:
                self.optimizer.state[p]["step"] += 1
        self.optimizer._ddpx_skip_step = True
