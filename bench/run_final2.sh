#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_p2p.py -q --tb=short 2>&1 | grep -v Warning | tail -12 > gpurun_out/pytest_p2p_9.log
timeout 300 $TR --master-port 29831 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/bench_ours_n2_4.log 2>&1
tail -3 gpurun_out/pytest_p2p_9.log; tail -1 gpurun_out/bench_ours_n2_4.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('n2', d['ms_per_step'], d['value'], d['e2e'], d['config'].get('exposed_comm_ms_per_step'))"
