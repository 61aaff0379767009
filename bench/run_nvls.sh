#!/bin/bash
# NVLS (multimem) all-reduce against two-shot and NCCL, and the VGG step with --allreduce-algo nvls.  $1 = GPUs.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
DDPX_ALLREDUCE_ALGO=nvls timeout 300 python bench/allreduce_sweep.py --gpus $N --max-mb 256 --ctas 64 > gpurun_out/sweep_nvls_n$N.log 2>&1
DDPX_ALLREDUCE_ALGO=nvls timeout 300 $TR --master-port 29821 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_ours_nvls_n$N.log 2>&1
timeout 300 $TR --master-port 29822 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_ours_twoshot_n$N.log 2>&1
python - <<PY
import json
rows=[json.loads(l) for l in open("gpurun_out/sweep_nvls_n$N.log") if l.startswith("{")]
by={}
for r in rows: by.setdefault(r["bytes"],{})[r["impl"]]=r
for b in sorted(by):
    d=by[b]; print(b, {k: (round(v["us"],1), round(v["busbw_gbs"])) for k,v in d.items()})
for f in ("nvls","twoshot"):
    d=json.loads(open(f"gpurun_out/bench_ours_{f}_n$N.log").read().strip().splitlines()[-1])
    print(f, d["ms_per_step"], d["value"], d["config"].get("exposed_comm_ms_per_step"))
PY
