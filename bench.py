"""Headline benchmark: training throughput (samples/s) of the reference's own workload.

Config ("reference default"): the VGG conv-BN-ReLU net of /root/reference/multigpu.py:36-71
on CIFAR-shaped synthetic data, per-GPU batch 512 (multigpu.py:259), fp32 parameters/activations with TF32
tensor-core convolutions (what torch's cuDNN default does for the reference), SGD(lr .4, momentum .9, wd 5e-4) +
the triangular LambdaLR stepped per batch (multigpu.py:129-143), data-parallel over N GPUs (weak scaling).

    python bench.py --gpus 1 --steps 20 --warmup 5                  # this framework
    python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 # unmodified reference (baseline/_ref)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N ...

One JSON line on stdout (rank 0). `value` is device-timed (CUDA events, max over ranks) with device-resident
synthetic batches; `e2e` re-measures through the public Trainer API with the batch copied from pinned host memory
every step and the loss read back to the host.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO))

BATCH_PER_GPU = 512
NUM_CLASSES = 10


# ------------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md "clocks DURING the timed region")
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock / power / throttle reasons DURING the timed region.  NVML is polled from a thread every ~4 ms (a 20-step
    timed region is ~0.1 s: `nvidia-smi -lms` gave it one or two samples); nvidia-smi is the fallback."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    # nvmlClocksEventReason* bit masks
    R_SW_POWER_CAP, R_HW_SLOWDOWN, R_SW_THERMAL, R_HW_THERMAL = 0x4, 0x8, 0x20, 0x40

    def __init__(self, gpu_index: int = 0):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines: list[str] = []
        self.thread = None
        self.samples = []          # (sm_mhz, power_w, reasons_mask)
        self.smax = None
        self._stop = False
        self.mode = None

    def _nvml_loop(self, nv, h):
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                rs = int(get_reasons(h))
                self.samples.append((float(sm), pw, rs))
            except Exception:
                pass
            time.sleep(0.004)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it is a plain index list
            idx = self.gpu_index
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                parts = [x.strip() for x in vis.split(",") if x.strip()]
                if idx < len(parts) and parts[idx].isdigit():
                    idx = int(parts[idx])
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.smax = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.mode = "nvml"
            self.thread = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
            self.thread.start()
            return
        except Exception:
            self.mode = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.mode = "smi"
        except Exception:
            self.proc = None
            return
        def pump():
            for line in self.proc.stdout:
                self.lines.append(line.strip())
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self) -> dict:
        if self.mode == "nvml":
            self._stop = True
            self.thread.join(timeout=1)
            if not self.samples:
                return {"sm_mhz": None, "sm_max_mhz": self.smax, "reasons": ["no samples"]}
            sm = sorted(x[0] for x in self.samples)
            mask = 0
            for x in self.samples:
                mask |= x[2]
            reasons = [n for n, bit in (("hw_slowdown", self.R_HW_SLOWDOWN), ("hw_thermal_slowdown", self.R_HW_THERMAL),
                                        ("sw_thermal_slowdown", self.R_SW_THERMAL), ("sw_power_cap", self.R_SW_POWER_CAP))
                       if mask & bit]
            return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.smax, "power_w_max": max(x[1] for x in self.samples),
                    "samples": len(sm), "source": "nvml", "reasons": sorted(reasons)}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        for ln in self.lines:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); smax.append(float(parts[2])); power.append(float(parts[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(smax), "power_w_max": max(power), "samples": len(sm),
                "source": "nvidia-smi", "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# shared helpers
# ------------------------------------------------------------------------------------------------
def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return rank, local_rank, world


def make_host_batches(n: int, batch: int, seed: int):
    """`n` distinct CIFAR-shaped batches in pinned host memory (synthetic: uniform pixels, random labels)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    xs, ys = [], []
    for _ in range(n):
        x = torch.rand(batch, 3, 32, 32, generator=g)
        y = torch.randint(0, NUM_CLASSES, (batch,), generator=g)
        if torch.cuda.is_available():
            x = x.pin_memory(); y = y.pin_memory()
        xs.append(x); ys.append(y)
    return xs, ys


def sync_report(params, loss_value, world: int) -> dict:
    """Driver-visible correctness at N > 1: every rank's fp64 parameter checksums must agree after the timed steps.
    Returns the keys to merge into the JSON line; the caller exits non-zero when replicas diverged."""
    from distributed_data_parallel_experiment_b200.utils import debug as _debug
    ok = bool(_debug.replicas_in_sync(list(params)))
    cs = _debug.replica_checksums(list(params))
    return {"replicas_in_sync": ok if world > 1 else None, "param_checksum": cs[0], "loss": loss_value}


def max_over_ranks(ms: float, world: int) -> float:
    import torch
    if world == 1:
        return ms
    import torch.distributed as dist
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def timed_loop(step_fn, steps: int, world: int) -> float:
    """barrier+sync, CUDA-event time `steps` calls of step_fn(i), barrier+sync; returns ms (max over ranks)."""
    import torch
    barrier(world)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step_fn(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    barrier(world)
    return max_over_ranks(ms, world)


def timed_block(fn, world: int) -> float:
    """Same bracket as timed_loop around ONE callable that runs all the steps itself."""
    import torch
    barrier(world)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    barrier(world)
    return max_over_ranks(ms, world)


def emit(args, world, impl, total_ms, e2e_ms, h2d, d2h, launches, clocks, detail, top=None):
    global_batch = BATCH_PER_GPU * world
    value = global_batch * args.steps / (total_ms / 1e3)
    e2e_value = global_batch * args.steps / (e2e_ms / 1e3) if e2e_ms else None
    baseline = None  # BASELINE.md publishes no number for the reference
    out = {
        "metric": "training throughput, samples/s (VGG/CIFAR-shaped, batch 512 per GPU, SGD+LambdaLR)",
        "value": round(value, 1), "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": baseline, "dtype": "fp32 storage, TF32 tensor-core conv math (reference default: cudnn.allow_tf32)",
        "data": "synthetic (uniform CIFAR-shaped images, random labels; random-init weights)",
        "impl": impl,
        "config": {"model": "VGG (reference multigpu.py:36-71, 9.23M params)", "global_batch": global_batch,
                   "batch_per_gpu": BATCH_PER_GPU, "seq_len": None, "image": "3x32x32",
                   "parallelism": f"dp{world}", "optimizer": "SGD lr0.4 m0.9 wd5e-4 + LambdaLR per step",
                   "l2": "per-step working set (~2.5 GB of activations) exceeds the 126 MB L2; 4 distinct input batches rotate"},
        "detail": detail,
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1) if e2e_value else None, "unit": "samples/s",
                "ms_per_step": round(e2e_ms / args.steps, 4) if e2e_ms else None,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
    }
    out.update(top or {})
    print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------------------
# reference arm: unmodified /root/reference scripts from baseline/_ref, through their own classes
# ------------------------------------------------------------------------------------------------
def run_reference(args):
    ref_dir = REPO / "baseline" / "_ref"
    if not (ref_dir / "multigpu.py").exists() and Path("/root/reference/multigpu.py").exists():
        # not installed yet (fresh checkout): local rank 0 runs the one-off offline install, the others wait for it
        import time
        if int(os.environ.get("LOCAL_RANK", "0")) == 0:
            subprocess.call(["bash", str(REPO / "baseline" / "install_ref.sh")], stdout=subprocess.DEVNULL,
                            stderr=subprocess.DEVNULL)
        else:
            for _ in range(240):
                if (ref_dir / "multigpu.py").exists():
                    break
                time.sleep(0.5)
    if not (ref_dir / "multigpu.py").exists():
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is missing (run baseline/install_ref.sh)"}))
        return
    sys.path.insert(0, str(ref_dir))
    import numpy as np
    import torch
    rank, local_rank, world = dist_env()
    try:
        if world > 1:
            import multigpu as ref   # noqa: the reference's multi-GPU script (DDP + NCCL)
        else:
            import singlegpu as ref  # noqa: the reference's single-GPU script
    except Exception as e:  # pragma: no cover
        print(json.dumps({"impl": "reference", "unavailable": f"import failed: {type(e).__name__}: {e}"}))
        return
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "reference", "unavailable": "no CUDA device"}))
        return
    torch.cuda.set_device(local_rank)
    if world > 1:
        # ddp_setup (multigpu.py:24-33) rewrites MASTER_ADDR/PORT to localhost:12355 for every rank and calls
        # init_process_group("nccl") - the reference's own bootstrap.  torchrun tells workers to reuse the agent's store
        # (TORCHELASTIC_USE_AGENT_STORE) at its own MASTER_PORT; the reference overwrites the port with 12355, where no
        # agent store listens, so let rank 0 host the store itself exactly as under the reference's mp.spawn launch.
        os.environ.pop("TORCHELASTIC_USE_AGENT_STORE", None)
        ref.ddp_setup(rank, world)
    torch.manual_seed(1234 + rank)
    # load_train_objs() (multigpu.py:122-144) minus the CIFAR-10 download (no network): same literals.
    model = ref.VGG().cuda()
    optimizer = torch.optim.SGD(model.parameters(), lr=0.4, momentum=0.9, weight_decay=5e-4)
    num_epochs, steps_per_epoch = 20, (49 if world > 1 else 98)
    lr_lambda = lambda step: np.interp([step / steps_per_epoch], [0, num_epochs * 0.3, num_epochs], [0, 1, 0])[0]
    scheduler = torch.optim.lr_scheduler.LambdaLR(optimizer, lr_lambda)
    trainer = ref.Trainer(model, [], optimizer, local_rank, 1, scheduler=scheduler)

    xs, ys = make_host_batches(4, BATCH_PER_GPU, seed=100 + rank)
    dxs = [x.cuda() for x in xs]; dys = [y.cuda() for y in ys]

    def dev_step(i):
        trainer._run_batch(dxs[i % 4], dys[i % 4])

    probe = trainer.model.module.classifier.bias if world > 1 else trainer.model.classifier.bias

    def e2e_step(i):
        # what Trainer._run_epoch does per batch (multigpu.py:104-107): blocking .to(gpu) then _run_batch;
        # the reference never reads the loss, so the device->host read is one updated parameter element.
        source = xs[i % 4].to(local_rank); targets = ys[i % 4].to(local_rank)
        trainer._run_batch(source, targets)
        probe[0].item()

    for i in range(args.warmup):
        dev_step(i)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    total_ms = timed_loop(dev_step, args.steps, world)
    clocks = sampler.stop() if rank == 0 else {}
    for i in range(2):
        e2e_step(i)
    e2e_ms = timed_loop(e2e_step, args.steps, world)
    with torch.no_grad():
        ref_model = trainer.model.module if world > 1 else trainer.model
        ref_loss = float(torch.nn.functional.cross_entropy(trainer.model(dxs[0]), dys[0]).item())
    rep = sync_report(ref_model.parameters(), ref_loss, world)
    if rank == 0:
        h2d = xs[0].numel() * 4 + ys[0].numel() * 8
        emit(args, world, "reference", total_ms, e2e_ms, h2d, 4, 0, clocks,
             {"stack": "torch DDP + NCCL + cuDNN/cuBLAS" if world > 1 else "torch eager + cuDNN/cuBLAS"}, top=rep)
    if world > 1:
        ref.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# this framework
# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import distributed_data_parallel_experiment_b200 as ddpx
    from distributed_data_parallel_experiment_b200 import _lib
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (CPU runs are covered by tests/)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        ddpx.launch.setup_from_env()
    torch.manual_seed(1234 + rank)
    train_objs = ddpx.recipes.vgg_train_objs(world_size=world, synthetic=True)
    model, optimizer, scheduler = train_objs.model, train_objs.optimizer, train_objs.scheduler
    trainer = ddpx.Trainer(model, [], optimizer, local_rank, 1, scheduler=scheduler)

    xs, ys = make_host_batches(4, BATCH_PER_GPU, seed=100 + rank)
    dxs = [x.cuda() for x in xs]; dys = [y.cuda() for y in ys]

    def dev_step(i):
        trainer._run_batch(dxs[i % 4], dys[i % 4])

    def e2e_step(i):
        loss = trainer.train_step(xs[i % 4], ys[i % 4])   # public API: pinned host batch in, host float out
        return loss

    for i in range(args.warmup):
        dev_step(i)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = _lib.launches()
    total_ms = timed_loop(dev_step, args.steps, world)
    launches = _lib.launches() - l0
    clocks = sampler.stop() if rank == 0 else {}
    # end to end through the public streaming API: every step's batch is copied from pinned host memory and every
    # step's loss is read back to the host; the upload of batch i+1 and the read-back of loss i-1 overlap step i
    def e2e_run(n):
        losses = list(trainer.train_stream((xs[i % 4], ys[i % 4]) for i in range(n)))
        assert len(losses) == n and all(l == l for l in losses)
    e2e_run(3)
    e2e_ms = timed_block(lambda: e2e_run(args.steps), world)
    extra_e2e_sync_ms = timed_loop(e2e_step, args.steps, world)      # unpipelined train_step(), for comparison
    extra = {"stack": "hand-written sm_100a kernels (tcgen05/TMEM/TMA conv+GEMM, fused BN/ReLU/pool/CE/SGD)"
                      + (", P2P fused all-reduce+SGD over NVLink" if world > 1 else ""),
             "e2e_api": "Trainer.train_stream (pipelined upload / loss read-back)",
             "e2e_unpipelined_ms_per_step": round(extra_e2e_sync_ms / args.steps, 4)}
    # outside the timed regions: host time to enqueue one step (no synchronisation inside) - tells launch-bound from
    # GPU-bound when the two arms or two GPU counts differ
    import time as _time
    torch.cuda.synchronize()
    h0 = _time.perf_counter()
    for i in range(10):
        dev_step(i)
    extra["host_enqueue_ms_per_step"] = round((_time.perf_counter() - h0) * 100.0, 3)
    torch.cuda.synchronize()
    if world > 1 and hasattr(trainer.model, "measure_exposed_comm"):
        # outside the timed regions: how long the compute stream stalls for the last gradient buckets, per step
        trainer.model.measure_exposed_comm(True)
        for i in range(8):
            dev_step(i)
        ex = trainer.model.exposed_comm_ms()
        trainer.model.measure_exposed_comm(False)
        if ex:
            ex_t = torch.tensor([sum(ex) / len(ex)], device="cuda", dtype=torch.float64)
            import torch.distributed as dist
            dist.all_reduce(ex_t, op=dist.ReduceOp.MAX)
            extra["exposed_comm_ms_per_step"] = round(float(ex_t.item()), 4)
            extra["exposed_comm_frac_of_step"] = round(float(ex_t.item()) / (total_ms / args.steps), 4)
    last_loss = float(trainer.last_loss.item()) if trainer.last_loss is not None else None
    inner = trainer.model.module if hasattr(trainer.model, "module") else trainer.model
    rep = sync_report(inner.parameters(), last_loss, world)
    if rank == 0:
        h2d = xs[0].numel() * 4 + ys[0].numel() * 8
        emit(args, world, "ddpx_b200", total_ms, e2e_ms, h2d, 4, launches, clocks, extra, top=rep)
    if world > 1:
        ddpx.launch.teardown()
    if world > 1 and rep["replicas_in_sync"] is False:
        raise SystemExit("replicas diverged: parameter checksums differ between ranks")
    if last_loss is None or last_loss != last_loss:
        raise SystemExit("training loss is NaN")


# ------------------------------------------------------------------------------------------------
# the other BASELINE.json configurations (the reference, a CNN script, cannot run them: the bar is its *stack* -
# torch modules + torch DDP + NCCL + cuBLAS / SDPA - measured in the same process tree)
# ------------------------------------------------------------------------------------------------
LM_CONFIGS = {"mlp": "mlp", "gpt2": "gpt2-small", "gpt2-xl": "gpt2-xl", "gpt2-tiny": "gpt2-tiny", "mlp-small": "mlp-small"}


def run_lm(args):
    import torch
    sys.path.insert(0, str(REPO / "bench"))
    import lm_bench
    import distributed_data_parallel_experiment_b200 as ddpx
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (CPU runs are covered by tests/)")
    torch.cuda.set_device(local_rank)
    name = LM_CONFIGS[args.config]
    cfg = lm_bench.CONFIGS[name]
    if world > 1:
        ddpx.launch.setup_from_env()          # P2P backend + a torch process group (gloo / NCCL) for the torch-stack arm
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ours = lm_bench.run_arm(name, "ours", rank, world, args.steps, args.warmup, e2e=True)
    clocks = sampler.stop() if rank == 0 else {}
    torch_arm = None
    if not args.no_torch_stack:
        try:
            torch_arm = lm_bench.run_arm(name, "torch", rank, world, args.steps, args.warmup, e2e=True)
        except Exception as e:      # the baseline arm must not take the product's number down with it
            torch_arm = {"error": f"{type(e).__name__}: {e}"}
    if rank == 0:
        kind = cfg["kind"]
        model = {"mlp": f"MLP {cfg.get('depth')}x{cfg.get('width')} bf16", "gpt2": f"GPT-2 {cfg.get('size')}"}[kind]
        out = {
            "metric": f"training throughput, {ours['unit']} ({model}, data-parallel)",
            "value": round(ours["value"], 1), "unit": ours["unit"], "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ours["ms_per_step"], 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "impl": "ddpx_b200",
            "data": "synthetic (random tokens / features, random-init weights)",
            "config": {"model": model, "global_batch": ours["global_batch"], "seq_len": ours["seq_len"],
                       "parallelism": f"dp{world}", "accum_steps": ours["accum"],
                       "optimizer": "SGD momentum 0.9" if kind == "mlp" else "AdamW lr3e-4 wd0.1",
                       "l2": "weights + activations of a step exceed the 126 MB L2; 4 distinct input batches rotate"},
            "clocks": clocks,
            "e2e": {"value": round(ours["e2e_value"], 1), "unit": ours["unit"], "ms_per_step": round(ours["e2e_ms_per_step"], 4),
                    "h2d_bytes_per_step": ours["h2d_bytes_per_step"], "d2h_bytes_per_step": ours["d2h_bytes_per_step"]},
            "gpu_launches": ours["gpu_launches"],
            "replicas_in_sync": ours.get("replicas_in_sync"), "loss": ours.get("loss"),
            "torch_stack_value": round(torch_arm["value"], 1) if torch_arm and "value" in torch_arm else None,
            "torch_stack_e2e_value": round(torch_arm["e2e_value"], 1) if torch_arm and "e2e_value" in torch_arm else None,
            "detail": {"params": ours["params"], "approx_model_tflops_per_gpu": round(ours["approx_model_tflops_per_gpu"], 1),
                       "max_mem_gb": round(ours["max_mem_gb"], 2), "torch_stack": torch_arm,
                       "torch_stack_is": "torch.nn modules + torch DDP + NCCL + cuBLAS / SDPA / fused AdamW, same dtype and shapes"},
        }
        print(json.dumps(out), flush=True)
    if world > 1:
        ddpx.launch.teardown()
    if ours.get("replicas_in_sync") is False:
        raise SystemExit("replicas diverged: parameter checksums differ between ranks")


def run_sweep_config(args):
    import torch
    sys.path.insert(0, str(REPO / "bench"))
    import allreduce_sweep
    import distributed_data_parallel_experiment_b200 as ddpx
    rank, local_rank, world = dist_env()
    if world < 2:
        print(json.dumps({"metric": "gradient all-reduce bus bandwidth", "unavailable": "needs >= 2 GPUs (--gpus 2|4|8)"}))
        return
    torch.cuda.set_device(local_rank)
    ddpx.launch.setup_from_env()
    sa = argparse.Namespace(max_mb=args.sweep_max_mb, dtype="fp32", ctas=[64], fused=True)
    rows = allreduce_sweep.run_sweep(rank, world, sa)
    if rank == 0:
        top = max(r["bytes"] for r in rows)
        best = max((r for r in rows if r["bytes"] == top and r["impl"].startswith("p2p") and "+" not in r["impl"]),
                   key=lambda r: r["busbw_gbs"])
        nccl = [r for r in rows if r["bytes"] == top and r["impl"] == "nccl"]
        print(json.dumps({
            "metric": "gradient all-reduce bus bandwidth, GB/s (largest size of the 1 KB - 1 GB sweep)",
            "value": round(best["busbw_gbs"], 1), "unit": "GB/s", "n_gpus": world, "steps": None, "warmup": 3,
            "higher_is_better": True, "scaling": "n/a", "vs_baseline": None, "dtype": "fp32", "impl": "ddpx_b200",
            "data": "synthetic", "config": {"model": "all-reduce sweep", "bytes": top, "parallelism": f"dp{world}"},
            "torch_stack_value": round(nccl[0]["busbw_gbs"], 1) if nccl else None,
            "frac_of_900": round(best["busbw_gbs"] / 900.0, 3), "frac_of_measured_770": round(best["busbw_gbs"] / 770.0, 3),
            "detail": {"rows": rows}}), flush=True)
    ddpx.launch.teardown()


def run_parity(args):
    """4 steps of this framework and of the unmodified reference (torch DDP + NCCL at N > 1) from the same rank-0 weights
    on the same batches: cosine and norm ratio of the two weight updates, per step loss of both."""
    import numpy as np
    import torch
    import distributed_data_parallel_experiment_b200 as ddpx
    ref_dir = REPO / "baseline" / "_ref"
    rank, local_rank, world = dist_env()
    if not (ref_dir / "multigpu.py").exists():
        print(json.dumps({"mode": "parity", "unavailable": "baseline/_ref is missing (run baseline/install_ref.sh)"}))
        return
    sys.path.insert(0, str(ref_dir))
    import multigpu as ref
    torch.cuda.set_device(local_rank)
    if world > 1:
        ddpx.launch.setup_from_env()
    torch.manual_seed(7)                                  # same initial weights on every rank and for both stacks
    objs = ddpx.recipes.vgg_train_objs(world_size=world, synthetic=True)
    ours_model = objs.model
    state0 = {k: v.detach().clone() for k, v in ours_model.state_dict().items()}
    ref_model = ref.VGG().cuda()
    ref_model.load_state_dict(state0, strict=True)
    ref_opt = torch.optim.SGD(ref_model.parameters(), lr=0.4, momentum=0.9, weight_decay=5e-4)
    spe = 49 if world > 1 else 98
    lam = lambda step: np.interp([step / spe], [0, 6, 20], [0, 1, 0])[0]
    ref_sched = torch.optim.lr_scheduler.LambdaLR(ref_opt, lam)
    ref_net = torch.nn.parallel.DistributedDataParallel(ref_model, device_ids=[local_rank]) if world > 1 else ref_model
    trainer = ddpx.Trainer(ours_model, [], objs.optimizer, local_rank, 1, scheduler=objs.scheduler)
    xs, ys = make_host_batches(args.parity_steps, BATCH_PER_GPU, seed=300 + rank)
    lo, lr_ = [], []
    for i in range(args.parity_steps):
        x = xs[i].cuda(); y = ys[i].cuda()
        lo.append(float(trainer._run_batch(x, y).item()))
        ref_opt.zero_grad()
        l = torch.nn.functional.cross_entropy(ref_net(x), y)
        l.backward(); ref_opt.step(); ref_sched.step()
        lr_.append(float(l.item()))
    inner = trainer.model.module if hasattr(trainer.model, "module") else trainer.model
    so, sr = inner.state_dict(), ref_model.state_dict()
    du = torch.cat([(so[k].float() - state0[k].float()).flatten() for k in state0 if state0[k].dtype.is_floating_point and "running" not in k])
    dr = torch.cat([(sr[k].float() - state0[k].float()).flatten() for k in state0 if state0[k].dtype.is_floating_point and "running" not in k])
    cos = float(torch.dot(du.double(), dr.double()) / (du.double().norm() * dr.double().norm()).clamp_min(1e-300))
    ratio = float(du.double().norm() / dr.double().norm().clamp_min(1e-300))
    rep = sync_report(inner.parameters(), lo[-1], world)
    if rank == 0:
        print(json.dumps({"mode": "parity", "n_gpus": world, "steps": args.parity_steps, "update_cosine": round(cos, 6),
                          "update_norm_ratio": round(ratio, 6), "loss_ours": lo, "loss_reference": lr_,
                          "replicas_in_sync": rep["replicas_in_sync"],
                          "note": "TF32 tensor-core convolutions on both sides: per-step rounding differs, direction and size of the update agree"}),
              flush=True)
    if world > 1:
        ddpx.launch.teardown()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)      # ~0.5 s timed: enough nvidia-smi samples under load
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="vgg", choices=["vgg", "sweep"] + list(LM_CONFIGS),
                    help="vgg = the reference's own workload (default, the only one with a reference-script arm); the others are "
                         "the remaining BASELINE.json configurations, each printing torch_stack_value beside value")
    ap.add_argument("--no-torch-stack", action="store_true", help="skip the torch-stack arm of the mlp / gpt2 configs")
    ap.add_argument("--sweep-max-mb", type=int, default=1024)
    ap.add_argument("--parity", action="store_true", help="compare 4 steps of this framework with the reference from the same weights")
    ap.add_argument("--parity-steps", type=int, default=4)
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    _, _, world = dist_env()
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun when invoked directly with --gpus N
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29511"), __file__,
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup), "--impl", args.impl,
               "--config", args.config, "--sweep-max-mb", str(args.sweep_max_mb), "--parity-steps", str(args.parity_steps)]
        cmd += (["--parity"] if args.parity else []) + (["--no-torch-stack"] if args.no_torch_stack else [])
        raise SystemExit(subprocess.call(cmd))
    if args.parity:
        run_parity(args)
    elif args.config == "sweep":
        run_sweep_config(args)
    elif args.config != "vgg":
        if args.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": "the reference is a CNN script (multigpu.py:36-71) and cannot run "
                              "this configuration; its torch stack is measured as torch_stack_value in the default arm's line"}))
        else:
            run_lm(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
