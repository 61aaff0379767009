#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -m gpu -q -x 2>&1 | grep -v Warning | tail -8 > gpurun_out/pytest_gpu_11.log
timeout 300 python bench.py --steps 50 > gpurun_out/bench_ours_12.log 2>&1
DDPX_WGRAD_PACKED=0 timeout 300 python bench.py --steps 50 > gpurun_out/bench_ours_12_unpacked.log 2>&1
timeout 300 python bench/profile_step.py 5 > gpurun_out/profile_step_10.log 2>&1
tail -3 gpurun_out/pytest_gpu_11.log
for f in bench_ours_12 bench_ours_12_unpacked; do tail -1 gpurun_out/$f.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$f', d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['clocks'])"; done
grep "wgrad\|unpack\|zero" gpurun_out/profile_step_10.log
