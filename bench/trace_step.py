"""Kernel timeline of VGG training steps (CUPTI via torch.profiler; diagnostics only - never a benchmark number).
    python bench/trace_step.py                      # 1 GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 bench/trace_step.py
Rank 0 writes gpurun_out/trace_rank0.json (chrome trace) and prints one line per kernel of the last traced step:
start offset (us), duration (us), stream, name."""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import distributed_data_parallel_experiment_b200 as ddpx

world = int(os.environ.get("WORLD_SIZE", "1"))
local_rank = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local_rank)
if world > 1:
    ddpx.launch.setup_from_env()
torch.manual_seed(local_rank)
model_name = sys.argv[1] if len(sys.argv) > 1 else "vgg"
if model_name == "vgg":
    objs = ddpx.recipes.vgg_train_objs(world_size=world, synthetic=True, synthetic_size=0)
    tr = ddpx.Trainer(objs.model, [], objs.optimizer, local_rank, 1, scheduler=objs.scheduler)
    x = torch.rand(512, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (512,), device="cuda")
    first_kernel = "im2col"
else:                                   # gpt2 | gpt2-medium | ... [batch]: seq 1024, bf16, AdamW
    objs, loss_fn = ddpx.recipes.make(model_name, world_size=world, synthetic=True, synthetic_size=0)
    tr = ddpx.Trainer(objs.model, [], objs.optimizer, local_rank, 1, loss_fn=loss_fn)
    V = objs.model.cfg.vocab_size
    Bt = int(sys.argv[2]) if len(sys.argv) > 2 else (2 if "xl" in model_name else 8)     # BASELINE.json micro-batches
    x = torch.randint(0, V, (Bt, 1024), device="cuda"); y = torch.randint(0, V, (Bt, 1024), device="cuda")
    first_kernel = "embedding"
for _ in range(6):
    tr._run_batch(x, y)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        tr._run_batch(x, y)
    torch.cuda.synchronize()
def _short(name: str) -> str:
    import re
    n = name.replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
    n = re.sub(r"^void\s+", "", n)
    n = n.split("(")[0].strip()
    return (n or name)[:64]


if local_rank == 0:
    out = Path("gpurun_out"); out.mkdir(exist_ok=True)
    path = out / f"trace_w{world}_rank0.json"
    prof.export_chrome_trace(str(path))
    ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    ev.sort(key=lambda e: e["ts"])
    # split into steps at the stem im2col kernel (first kernel of a forward)
    starts = [i for i, e in enumerate(ev) if first_kernel in e["name"] and "bwd" not in e["name"]]
    if len(starts) >= 2:
        a, b = starts[-2], starts[-1]
        step = ev[a:b]
        t0 = step[0]["ts"]
        print(f"# world {world}: step span {ev[b]['ts'] - t0:.1f} us, {len(step)} GPU activities")
        busy = {}
        for e in step:
            s = e["args"].get("stream", "?")
            busy[s] = busy.get(s, 0.0) + e["dur"]
            nm = _short(e["name"])
            print(f"{e['ts'] - t0:9.1f} {e['dur']:8.1f}  s{s}  {nm}")
        agg = {}
        for e in step:
            nm = _short(e["name"])
            a = agg.setdefault(nm, [0, 0.0]); a[0] += 1; a[1] += e["dur"]
        print("# per kernel (calls, total us):")
        for nm, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print(f"#   {d:9.1f} us  {c:4d}  {nm}")
        print("# busy us per stream:", {k: round(v, 1) for k, v in busy.items()})
if world > 1:
    ddpx.launch.teardown()
