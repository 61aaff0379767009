#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -m gpu -q -x 2>&1 | grep -v Warning | tail -8 > gpurun_out/pytest_gpu_12.log
timeout 300 python bench.py --steps 50 > gpurun_out/bench_ours_13.log 2>&1
timeout 300 python bench/lm_bench.py --config gpt2-small --impl ours > gpurun_out/lm_4.log 2>&1
timeout 300 python bench/lm_bench.py --config mlp --impl ours >> gpurun_out/lm_4.log 2>&1
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
tail -3 gpurun_out/pytest_gpu_12.log; tail -1 gpurun_out/bench_ours_13.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['config'].get('host_enqueue_ms_per_step'), d['gpu_launches'])"
grep "^{" gpurun_out/lm_4.log | cut -c1-200; tail -1 gpurun_out/smoke.log
