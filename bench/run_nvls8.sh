#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_p2p.py -q -k nvls 2>&1 | grep -v Warning | tail -2 > gpurun_out/pytest_nvls_n8.log
DDPX_ALLREDUCE_ALGO=nvls timeout 200 python bench/allreduce_sweep.py --gpus 8 --max-mb 256 --ctas 64 > gpurun_out/sweep_nvls_n8.log 2>&1
DDPX_ALLREDUCE_ALGO=nvls timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29821 bench.py --gpus 8 --steps 50 --warmup 5 > gpurun_out/bench_ours_nvls_n8.log 2>&1
tail -1 gpurun_out/pytest_nvls_n8.log
python - <<PY
import json
rows=[json.loads(l) for l in open("gpurun_out/sweep_nvls_n8.log") if l.startswith("{")]
by={}
for r in rows: by.setdefault(r["bytes"],{})[r["impl"]]=r
for b in sorted(by):
    d=by[b]; print(b, {k: (round(v["us"],1), round(v["busbw_gbs"])) for k,v in d.items()})
d=json.loads(open("gpurun_out/bench_ours_nvls_n8.log").read().strip().splitlines()[-1])
print("nvls", d["ms_per_step"], d["value"], d["config"].get("exposed_comm_ms_per_step"))
PY
