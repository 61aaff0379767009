"""Gradient reducer: buckets, autograd hooks, in-order launch, finalisation, one-time rebuild, ``no_sync``.

Counterpart of ``c10d::Reducer`` (c10d/reducer.hpp:45-584) that torch DDP builds for the reference at multigpu.py:89.
Bookkeeping (ready counting, strict bucket order, ready-order recording) is the C++ core in
csrc/runtime/reducer_core.cpp; this file owns the tensors:

* every bucket is ONE flat buffer from the communication backend (peer-mapped symmetric memory on the P2P backend),
  double-buffered by step parity so a fast rank can start filling step t+1 while a slow peer still reads step t;
* a post-accumulate-grad hook per parameter copies the gradient into its bucket slot pre-divided by the world size
  (one small kernel) and asks the core which buckets became launchable;
* launchable buckets are handed to the backend immediately (overlapping the rest of backward); the last one queues a
  finalisation callback on the autograd engine that waits for the collectives and points ``param.grad`` at the bucket
  (no copy back), or - on the fused path - lets the all-reduce kernel apply the optimizer update itself.
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence

import torch

from .. import _lib
from .buckets import DEFAULT_FIRST_BUCKET_BYTES, compute_bucket_assignment_by_size


class Bucket:
    __slots__ = ("index", "param_ids", "offsets", "numel", "dtype", "buffers", "views", "handle", "size_bytes")

    def __init__(self, index, param_ids, params, comm, device, double_buffer):
        self.index = index
        self.param_ids = list(param_ids)
        self.offsets = []
        off = 0
        for pid in self.param_ids:
            self.offsets.append(off)
            n = params[pid].numel()
            off += (n + 7) // 8 * 8          # keep every slot 16-byte aligned for vector loads (fp32 and bf16)
        self.numel = max(off, 8)
        self.dtype = params[self.param_ids[0]].dtype
        self.size_bytes = sum(params[p].numel() * params[p].element_size() for p in self.param_ids)
        nbuf = 2 if double_buffer else 1
        self.buffers = [comm.alloc(self.numel, self.dtype, device) for _ in range(nbuf)]
        self.views = [[buf[o:o + params[pid].numel()].view_as(params[pid]) for pid, o in zip(self.param_ids, self.offsets)]
                      for buf in self.buffers]
        self.handle = None


class Reducer:
    def __init__(self, params: Sequence[torch.nn.Parameter], comm, bucket_cap_bytes: int,
                 first_bucket_bytes: int = DEFAULT_FIRST_BUCKET_BYTES, fused=None):
        self.params = list(params)
        self.comm = comm
        self.world_size = comm.world_size
        self.bucket_cap_bytes = int(bucket_cap_bytes)
        self.first_bucket_bytes = int(first_bucket_bytes)
        self.device = self.params[0].device if self.params else torch.device("cpu")
        self.fused = fused                      # FusedUpdate object (parallel/p2p.py) or None
        self.double_buffer = getattr(comm, "needs_double_buffer", False)
        self._lib = _lib.lib()
        self._core = None
        self.buckets: List[Bucket] = []
        self.armed = False                      # prepare_for_backward() called and not yet finalised
        self.expect_finalize = False
        self.has_rebuilt = False
        self.iteration = 0
        self.step_parity = 0
        self._pid_of = {id(p): i for i, p in enumerate(self.params)}
        self._slot = {}                         # param id -> (bucket, slot)
        self._hooks = []
        self._callback_queued = False
        self.on_first_launch = None             # hook: called once per synchronised backward, after the first bucket launched
        self._launched_this_iter = 0
        self._scratch = (ctypes.c_int * max(1, len(self.params)))()
        # CUDA: gradients are gathered into their bucket by ONE multi-tensor kernel when the bucket becomes launchable
        # instead of one small kernel per parameter from the hook (34 launches per VGG step -> 3)
        self._batch_copy = (self.device.type == "cuda" and
                            all(p.dtype in (torch.float32, torch.bfloat16) for p in self.params))
        self.stats = {"rebuilt_bucket_sizes": [], "num_buckets": 0, "launched": 0}
        # Native path (CUDA + P2P backend): C++ post-hooks on the gradient accumulators feed the native reducer
        # (csrc/torch_ext/autograd_hooks.cpp -> csrc/runtime/reducer.cu); gather, stream hand-over, kernel launch and
        # completion events happen there.  Everything else (gloo, NCCL, CPU) keeps the Python hooks below.
        self._nh = None
        self._hookset = None
        self._native = self._native_possible()
        # first iteration: one bucket holding everything (distributed.py:1224-1225, find_unused_parameters=False)
        idx, _ = compute_bucket_assignment_by_size(self.params, [2 ** 62])
        if self._native:
            from . import hooks_ext
            self._hookset = hooks_ext.module().HookSet(self.params, self._first_launch_cb, self.finalize_backward)
        self._install(list(reversed(idx)))
        if not self._native:
            for i, p in enumerate(self.params):
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))

    def _native_possible(self) -> bool:
        import os
        if self.device.type != "cuda" or not self._batch_copy or self.double_buffer:
            return False
        if os.environ.get("DDPX_NATIVE_REDUCER", "1") in ("0", ""):
            return False
        comm = self.comm
        if not all(hasattr(comm, a) for a in ("pad", "stream", "_pads_arr", "symbuf_of")):
            return False
        from . import hooks_ext
        return hooks_ext.module() is not None

    def _first_launch_cb(self):
        if self.on_first_launch is not None:
            self.on_first_launch()

    def _install_native(self):
        lib, comm = self._lib, self.comm
        if self._nh is not None:
            lib.ddpx_nreducer_destroy(ctypes.c_void_p(self._nh))
            self._nh = None
        n = len(self.params)
        bof = (ctypes.c_int * max(1, n))()
        for b in self.buckets:
            for pid in b.param_ids:
                bof[pid] = b.index
        lib.ddpx_nreducer_create.restype = ctypes.c_void_p
        self._nh = lib.ddpx_nreducer_create(n, len(self.buckets), bof, comm.rank, comm.world_size, comm._pads_arr,
                                            ctypes.c_void_p(comm.stream.cuda_stream), int(comm.ctas))
        if not self._nh:
            raise RuntimeError("ddpx_nreducer_create failed")
        for b in self.buckets:
            sb = comm.symbuf_of(b.buffers[0])
            k = len(b.param_ids)
            ptrs = (ctypes.c_void_p * max(1, comm.world_size))(*[int(q) for q in sb.ptrs])
            _lib.check(lib.ddpx_nreducer_set_bucket(
                ctypes.c_void_p(self._nh), b.index, k, (ctypes.c_int * k)(*b.param_ids),
                (ctypes.c_longlong * k)(*b.offsets), (ctypes.c_longlong * k)(*[self.params[i].numel() for i in b.param_ids]),
                ptrs, ctypes.c_void_p(getattr(sb, "mc_ptr", 0) or 0), ctypes.c_longlong(b.numel),
                1 if b.dtype == torch.bfloat16 else 0), "ddpx_nreducer_set_bucket")
        mark = ctypes.cast(lib.ddpx_nreducer_mark_ready, ctypes.c_void_p).value
        self._hookset.set_native(int(self._nh), int(mark))

    def on_slots_changed(self):
        """FusedUpdate (re)built its slot tables: hand the parameter / optimizer-state pointers to the native reducer."""
        if not self._native or self._nh is None or self.fused is None:
            return
        for b in self.buckets:
            tab = self.fused.tables.get(b.index)
            if tab is None:
                _lib.check(self._lib.ddpx_nreducer_set_slots(ctypes.c_void_p(self._nh), b.index, 0, None, None, None),
                           "ddpx_nreducer_set_slots")
                continue
            ps, s0, s1, _, _ = tab
            k = len(ps)
            arr = lambda v: (ctypes.c_void_p * k)(*[int(x) if x else None for x in v])
            _lib.check(self._lib.ddpx_nreducer_set_slots(ctypes.c_void_p(self._nh), b.index, k, arr(ps), arr(s0), arr(s1)),
                       "ddpx_nreducer_set_slots")

    def bucket_times(self):
        """Per-bucket device time of the last synchronised backward (native path): list of dicts with bytes, ms, the
        kernel flavour that ran and the bus bandwidth it reached (2 (w-1)/w bytes / time).  Waits for the kernels."""
        if not self._native or self._nh is None:
            return []
        nb = len(self.buckets)
        ms = (ctypes.c_float * nb)(); mode = (ctypes.c_int * nb)()
        self._lib.ddpx_nreducer_bucket_times(ctypes.c_void_p(self._nh), ms, mode, nb)
        w = self.world_size
        out = []
        for b in self.buckets:
            nbytes = b.numel * (2 if b.dtype == torch.bfloat16 else 4)
            t = float(ms[b.index])
            out.append({"bucket": b.index, "bytes": nbytes, "ms": t if t >= 0 else None,
                        "algorithm": ("twoshot", "oneshot", "nvls")[int(mode[b.index])] if w > 1 else "local",
                        "bus_gbs": (nbytes * 2 * (w - 1) / w / (t * 1e6)) if (t > 0 and w > 1) else None})
        return out

    # ------------------------------------------------------------------------------------------ structure
    def _install(self, bucket_indices: List[List[int]]):
        if self._core is not None:
            self._lib.ddpx_reducer_destroy(ctypes.c_void_p(self._core))
            self._core = None
        old = self.buckets
        self.buckets = [Bucket(b, ids, self.params, self.comm, self.device, self.double_buffer)
                        for b, ids in enumerate(bucket_indices)]
        # the replaced buckets (the whole-model bucket of the first iteration) go back to the backend: on the P2P backend
        # that unmaps them on every peer and frees the allocation (collective, same order on every rank)
        release = getattr(self.comm, "free", None)
        for ob in old:
            for p in self.params:
                if p.grad is not None and any(p.grad.data_ptr() == v.data_ptr() for vs in ob.views for v in vs):
                    p.grad = None              # a gradient still viewing the old bucket must not outlive it
            bufs, ob.buffers, ob.views = ob.buffers, [], []
            if release is not None:
                for buf in bufs:
                    release(buf)
        del old
        n = len(self.params)
        bof = (ctypes.c_int * max(1, n))()
        self._slot = {}
        for b in self.buckets:
            for s, pid in enumerate(b.param_ids):
                bof[pid] = b.index
                self._slot[pid] = (b, s)
        self._lib.ddpx_reducer_create.restype = ctypes.c_void_p
        self._core = self._lib.ddpx_reducer_create(n, len(self.buckets), bof)
        if not self._core:
            raise RuntimeError("ddpx_reducer_create failed")
        if self._native:
            self._install_native()
        self.stats["num_buckets"] = len(self.buckets)
        if self.fused is not None:
            self.fused.on_buckets_installed(self)

    def bucket_sizes(self) -> List[int]:
        return [b.size_bytes for b in self.buckets]

    def rebuild_buckets(self) -> bool:
        """One-time rebuild in the gradient-ready order observed during the first backward (reducer.hpp:136,466),
        rank 0's order wins (sync_bucket_indices, reducer.hpp:301).  Returns True if a rebuild happened."""
        if self.has_rebuilt or self.iteration == 0:
            return False
        n = len(self.params)
        if self._native:
            cnt = self._lib.ddpx_nreducer_ready_order(ctypes.c_void_p(self._nh), self._scratch, n)
        else:
            cnt = self._lib.ddpx_reducer_ready_order(ctypes.c_void_p(self._core), self._scratch, n)
        order = [self._scratch[i] for i in range(min(cnt, n))]
        if len(order) != n:
            return False
        order = self.comm.broadcast_object(order, src=0)
        tensors = [self.params[i] for i in order]
        idx, _ = compute_bucket_assignment_by_size(tensors, [self.first_bucket_bytes, self.bucket_cap_bytes],
                                                   tensor_indices=order)
        idx = self._split_tail(idx)
        if self.fused is not None and hasattr(self.fused, "gather_state"):
            # the rebuild changes which rank owns which element of the (sharded) optimizer state: make it whole first,
            # while the old bucket geometry is still installed
            self.fused.gather_state()
        self._install(idx)
        self.has_rebuilt = True
        self.stats["rebuilt_bucket_sizes"] = self.bucket_sizes()
        return True

    def _split_tail(self, idx: List[List[int]]) -> List[List[int]]:
        """P2P backend only: give the parameters that become ready LAST a small bucket of their own.  The last bucket's
        all-reduce cannot overlap anything (nothing is left of backward), so what it costs is pure latency: for VGG the
        torch layout ends with a 303 KB bucket {bn1, conv1, bn0, conv0} whose kernel took 26 us (2 GPUs) / 64 us (8 GPUs)
        after the last compute kernel; split, the 296 KB part overlaps the stem's backward and only {bn0, conv0} (7 KB)
        remains exposed.  ``DDPX_SPLIT_TAIL_BYTES=0`` keeps torch's layout bucket for bucket."""
        from .. import config
        limit = int(config.get().split_tail_bytes)
        if limit <= 0 or not getattr(self.comm, "supports_fused_update", False) or not idx or len(idx[-1]) < 2:
            return idx
        last = idx[-1]
        size = lambda i: self.params[i].numel() * self.params[i].element_size()
        cut, acc = len(last), 0
        while cut > 1 and acc + size(last[cut - 1]) <= limit:
            cut -= 1
            acc += size(last[cut])
        if cut == len(last) or acc == 0:
            return idx
        return idx[:-1] + [last[:cut], last[cut:]]

    # ------------------------------------------------------------------------------------------ per iteration
    def prepare_for_backward(self):
        if self.expect_finalize:
            missing = self._unmarked()
            raise RuntimeError(
                "Expected to have finished reduction in the prior iteration before starting a new one. "
                f"Parameters that did not receive a gradient: {missing} (find_unused_parameters is not supported)")
        if self._native:
            self._arm_native()
        else:
            self._lib.ddpx_reducer_reset(ctypes.c_void_p(self._core))
        self.armed = True
        self.expect_finalize = True
        self._callback_queued = False
        self._launched_this_iter = 0
        for b in self.buckets:
            b.handle = None

    def _arm_native(self):
        """Once per synchronised backward: this iteration's algorithm / hyper-parameters per bucket (LambdaLR moves lr
        every step), then arm the native reducer and the C++ hooks with the stream backward will run on."""
        nh = ctypes.c_void_p(self._nh)
        for b in self.buckets:
            if self.fused is not None:
                mode, opt, hyper, nesterov, first_step, _ = self.fused.plan(b)
            else:
                has_mc = bool(getattr(self.comm.symbuf_of(b.buffers[0]), "mc_ptr", 0))
                mode, opt, hyper, nesterov, first_step = (2 if has_mc else 0), 0, None, False, False
            hy = (ctypes.c_float * 9)(*hyper) if hyper else None
            self._lib.ddpx_nreducer_set_step(nh, b.index, int(mode), int(opt), hy, int(nesterov), int(first_step))
        self._lib.ddpx_nreducer_arm(nh, 1)
        self._hookset.arm(int(torch.cuda.current_stream(self.device).cuda_stream))

    def _unmarked(self):
        n = len(self.params)
        if self._native:
            cnt = self._lib.ddpx_nreducer_unmarked(ctypes.c_void_p(self._nh), self._scratch, n)
            return [self._scratch[i] for i in range(min(cnt, n))]
        cnt = self._lib.ddpx_reducer_unmarked(ctypes.c_void_p(self._core), self._scratch, n)
        return [self._scratch[i] for i in range(min(cnt, n))]

    def _make_hook(self, pid: int):
        def hook(param):
            if not self.armed:
                return            # no_sync(): gradients just accumulate in param.grad
            self._on_grad_ready(pid, param)
        return hook

    def _on_grad_ready(self, pid: int, param):
        bucket, slot = self._slot[pid]
        view = bucket.views[self.step_parity if self.double_buffer else 0][slot]
        g = param.grad
        if g is None:
            raise RuntimeError("gradient hook fired without a gradient")
        if not self._batch_copy:
            scale = 1.0 / self.world_size
            if g.data_ptr() != view.data_ptr():
                _scaled_copy(view, g, scale)
            elif scale != 1.0:
                _scaled_copy(view, view, scale)
        out = self._scratch
        k = self._lib.ddpx_reducer_mark_ready(ctypes.c_void_p(self._core), pid, out, len(self.params))
        if k == -1:
            raise RuntimeError(f"parameter {pid} was marked ready twice in one iteration (reducer.hpp:559-568)")
        if k < 0:
            raise RuntimeError("reducer: bad parameter index")
        for j in range(k):
            self._launch(self.buckets[out[j]])
        if k > 0 and self._lib.ddpx_reducer_all_launched(ctypes.c_void_p(self._core)) and not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self.finalize_backward)

    def _gather_bucket(self, bucket: Bucket):
        """All gradients of the bucket -> their slots, pre-divided by the world size, in one launch."""
        views = bucket.views[self.step_parity if self.double_buffer else 0]
        scale = 1.0 / self.world_size
        dsts, srcs = [], []
        for pid, view in zip(bucket.param_ids, views):
            g = self.params[pid].grad
            if g.dtype == view.dtype and g.is_contiguous():
                if g.data_ptr() == view.data_ptr() and scale == 1.0:
                    continue
                dsts.append(view); srcs.append(g)
            else:
                view.copy_(g)
                if scale != 1.0:
                    view.mul_(scale)
        if dsts:
            from ..ops import native
            native.scale_copy_multi(dsts, srcs, scale)

    def _launch(self, bucket: Bucket):
        if self._batch_copy:
            self._gather_bucket(bucket)
        buf = bucket.buffers[self.step_parity if self.double_buffer else 0]
        self.stats["launched"] += 1
        if self.fused is not None:
            bucket.handle = self.fused.reduce_and_update(bucket, buf, self.step_parity)
        else:
            bucket.handle = self.comm.all_reduce_sum_(buf, async_op=True)
        self._launched_this_iter += 1
        if self._launched_this_iter == 1 and self.on_first_launch is not None:
            self.on_first_launch()

    def finalize_backward(self):
        parity = self.step_parity if self.double_buffer else 0
        timing = getattr(self, "measure_exposed", False) and self.device.type == "cuda"
        if timing:      # exposed communication = how long the compute stream stalls here for the last buckets
            t0 = torch.cuda.Event(enable_timing=True)
            t0.record()
        if self._native:
            _lib.check(self._lib.ddpx_nreducer_wait(ctypes.c_void_p(self._nh),
                                                    ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)),
                       "ddpx_nreducer_wait")
            k = int(self._lib.ddpx_nreducer_launched(ctypes.c_void_p(self._nh)))
            self.stats["launched"] += k
            _lib.count_launch(2 * k)            # gather + fused all-reduce kernel per bucket
        for b in self.buckets:
            if b.handle is not None:
                b.handle.wait()
                b.handle = None
            if self.fused is None:
                for pid, v in zip(b.param_ids, b.views[parity]):
                    self.params[pid].grad = v          # averaged gradient lives in the bucket (no copy back)
        if timing:
            t1 = torch.cuda.Event(enable_timing=True)
            t1.record()
            self.stats.setdefault("exposed_events", []).append((t0, t1))
        if self.fused is not None:
            self.fused.after_backward(self)
        self.armed = False
        self.expect_finalize = False
        self.iteration += 1
        if self.double_buffer:
            self.step_parity ^= 1

    def close(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []
        if self._hookset is not None:
            self._hookset.remove()
            self._hookset = None
        if self._nh is not None:
            self._lib.ddpx_nreducer_destroy(ctypes.c_void_p(self._nh))
            self._nh = None
        if self._core is not None:
            self._lib.ddpx_reducer_destroy(ctypes.c_void_p(self._core))
            self._core = None


def _scaled_copy(dst: torch.Tensor, src: torch.Tensor, scale: float):
    """dst = src * scale with the framework's own kernel on CUDA fp32; torch math elsewhere."""
    if dst.is_cuda and dst.dtype == torch.float32 and src.dtype == torch.float32 and src.is_contiguous():
        from ..ops import native
        native.scale_copy(src, dst, dst.numel(), scale, 0)
    elif dst.is_cuda and src.is_contiguous() and dst.dtype == src.dtype and dst.dtype == torch.bfloat16:
        from ..ops import native
        native.scale_copy_bf16(src, dst, dst.numel(), scale)
    else:
        if dst.data_ptr() == src.data_ptr():
            dst.mul_(scale)
        else:
            dst.copy_(src)
            if scale != 1.0:
                dst.mul_(scale)
