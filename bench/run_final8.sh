#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29841 bench.py --gpus 8 --steps 50 --warmup 5 > gpurun_out/bench_ours_n8_c.log 2>&1
timeout 200 $TR --nproc-per-node 4 --master-port 29842 bench.py --gpus 4 --steps 50 --warmup 5 > gpurun_out/bench_ours_n4_c.log 2>&1
for f in bench_ours_n8_c bench_ours_n4_c; do tail -1 gpurun_out/$f.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$f', d['ms_per_step'], d['value'], d['e2e']['value'], d['config'].get('exposed_comm_ms_per_step'), d['clocks'])"; done
