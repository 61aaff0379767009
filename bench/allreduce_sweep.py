"""Gradient all-reduce bandwidth sweep 1 KB - 1 GB (BASELINE.json config 5): the P2P two-shot kernel vs NCCL on the
same buffers, device-timed (CUDA events, max over ranks), reported as algorithm and bus bandwidth against the
900 GB/s-per-direction NVLink 5 figure (measured peer copy on this pool: 770 GB/s, B200_PROFILING.md).

    python bench/allreduce_sweep.py --gpus 2 [--max-mb 1024] [--dtype fp32|bf16] [--ctas 32,64]
Writes one JSON line per (size, impl) on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _time(fn, iters, world):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_sweep(rank, world, args):
    """The sweep itself; the caller has set up the P2P backend (ddpx.ddp_setup / setup_from_env).  Returns the rows."""
    import distributed_data_parallel_experiment_b200 as ddpx
    comm = ddpx.launch.get_comm()
    dtype = torch.float32 if args.dtype == "fp32" else torch.bfloat16
    es = 4 if dtype == torch.float32 else 2
    size = 1024
    results = []
    while size <= args.max_mb * 1024 * 1024:
        n = max(8, size // es // 8 * 8)
        buf = comm.alloc(n, dtype, None)
        buf.normal_()
        ref = torch.empty(n, dtype=dtype, device="cuda").normal_()
        iters = 50 if size <= (1 << 22) else (10 if size <= (1 << 28) else 4)
        has_mc = bool(getattr(comm.symbuf_of(buf), "mc_ptr", 0))
        for ctas in args.ctas:
            comm.ctas = ctas
            # plain all-reduce: two-shot on every buffer, and the in-switch reduction where the buffer is multicast-bound
            # (run with DDPX_ALLREDUCE_ALGO=nvls / --nvls so that every size is)
            for mode, label in ((0, "p2p_twoshot"), (2, "p2p_nvls")):
                if mode == 2 and not has_mc:
                    continue
                ms = _time(lambda: comm.launch_allreduce(buf, mode=mode, opt=0).wait(), iters, world)
                algbw = size / ms / 1e6
                results.append({"impl": label, "ctas": ctas, "bytes": size, "us": ms * 1e3, "algbw_gbs": algbw,
                                "busbw_gbs": algbw * 2 * (world - 1) / world})
        if args.fused and size <= (64 << 20):
            # the product path: all-reduce + 1/world + SGD-momentum in ONE kernel (two-shot / one-shot) against the
            # baseline pair NCCL all-reduce + torch's fused SGD step on a parameter of the same size
            prm = torch.zeros(n, dtype=dtype, device="cuda"); mom = torch.zeros(n, dtype=dtype, device="cuda")
            slots = ([prm.data_ptr()], [mom.data_ptr()], [0], [0], [n])
            hyper = [1e-3, 0.9, 0.0, 5e-4, 0.0, 0.0, 0.0, 1.0, 1.0]
            comm.ctas = args.ctas[-1]
            for mode, label in ((0, "p2p_twoshot+sgd"), (1, "p2p_oneshot+sgd"), (2, "p2p_nvls+sgd")):
                if (mode == 1 and size > (16 << 20)) or (mode == 2 and not has_mc):
                    continue
                ms = _time(lambda: comm.launch_allreduce(buf, mode=mode, opt=1, hyper=hyper, slots=slots).wait(), iters, world)
                results.append({"impl": label, "ctas": comm.ctas, "bytes": size, "us": ms * 1e3, "algbw_gbs": size / ms / 1e6,
                                "busbw_gbs": size / ms / 1e6 * 2 * (world - 1) / world})
            tp = torch.nn.Parameter(torch.zeros(n, dtype=dtype, device="cuda")); tp.grad = ref
            topt = torch.optim.SGD([tp], lr=1e-3, momentum=0.9, weight_decay=5e-4, fused=True)

            def base_step():
                dist.all_reduce(ref)
                topt.step()
            ms = _time(base_step, iters, world)
            results.append({"impl": "nccl+torch_fused_sgd", "ctas": None, "bytes": size, "us": ms * 1e3,
                            "algbw_gbs": size / ms / 1e6, "busbw_gbs": size / ms / 1e6 * 2 * (world - 1) / world})
            del prm, mom, tp, topt
        ms = _time(lambda: dist.all_reduce(ref), iters, world)
        algbw = size / ms / 1e6
        results.append({"impl": "nccl", "ctas": None, "bytes": size, "us": ms * 1e3, "algbw_gbs": algbw,
                        "busbw_gbs": algbw * 2 * (world - 1) / world})
        size *= 4
    for r in results:
        r.update({"world": world, "dtype": args.dtype, "frac_of_900": r["busbw_gbs"] / 900.0,
                  "frac_of_measured_770": r["busbw_gbs"] / 770.0})
    return results


def emit_table(rows, world, path):
    """Algorithm-selection thresholds for `world` GPUs from the fused rows of a sweep, merged into the JSON table that
    parallel/p2p.py loads: the largest size for which one-shot+SGD is not slower than two-shot+SGD (and every smaller
    measured size agrees), the smallest size from which NVLS+SGD beats two-shot+SGD by at least 3 % for every larger size
    (differences inside run-to-run noise must not flip the algorithm)."""
    by = {}
    for r in rows:
        by.setdefault(r["bytes"], {})[r["impl"]] = r["us"]
    sizes = sorted(by)
    one = None
    for sz in sizes:
        t = by[sz]
        if "p2p_oneshot+sgd" in t and "p2p_twoshot+sgd" in t and t["p2p_oneshot+sgd"] <= t["p2p_twoshot+sgd"] * 1.02:
            one = sz
        elif "p2p_oneshot+sgd" in t and "p2p_twoshot+sgd" in t:
            break
    nv = None
    for sz in reversed(sizes):
        t = by[sz]
        if "p2p_nvls+sgd" in t and "p2p_twoshot+sgd" in t:
            if t["p2p_nvls+sgd"] <= t["p2p_twoshot+sgd"] * 0.97:
                nv = sz
            else:
                break
    table = json.loads(Path(path).read_text()) if Path(path).exists() else {}
    entry = dict(table.get(str(world), {}))
    if one is not None:
        entry["oneshot_max_bytes"] = one
    if any("p2p_nvls+sgd" in by[sz] for sz in sizes):
        entry["nvls_min_bytes"] = nv
    table[str(world)] = entry
    Path(path).write_text(json.dumps(table, indent=1) + "\n")
    return entry


def worker(rank, world, port, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    if getattr(args, "nvls", False):
        os.environ["DDPX_ALLREDUCE_ALGO"] = "nvls"        # every buffer multicast-bound: both flavours on the same memory
    import distributed_data_parallel_experiment_b200 as ddpx
    ddpx.ddp_setup(rank, world, backend="p2p")
    results = run_sweep(rank, world, args)
    if rank == 0:
        for r in results:
            print(json.dumps(r), flush=True)
        if getattr(args, "emit_table", None):
            print(json.dumps({"table_entry": emit_table(results, world, args.emit_table), "world": world}), flush=True)
    ddpx.teardown()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--ctas", default="64")
    ap.add_argument("--fused", action="store_true", help="also time all-reduce+SGD in one kernel vs NCCL + torch fused SGD")
    ap.add_argument("--nvls", action="store_true", help="allocate every buffer from the multicast allocator and time the NVLS flavour too")
    ap.add_argument("--emit-table", default=None, metavar="JSON",
                    help="merge the thresholds measured by this run (needs --fused) into the algorithm table, e.g. "
                         "distributed_data_parallel_experiment_b200/parallel/algo_table.json")
    args = ap.parse_args()
    args.ctas = [int(c) for c in args.ctas.split(",")]
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(worker, args=(args.gpus, port, args), nprocs=args.gpus, join=True)


if __name__ == "__main__":
    main()
