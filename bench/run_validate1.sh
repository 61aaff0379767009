#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -m gpu -q -x 2>&1 | grep -v Warning | tail -8 > gpurun_out/pytest_gpu_9.log
timeout 300 python bench.py > gpurun_out/bench_ours_10.log 2>&1
tail -3 gpurun_out/pytest_gpu_9.log; tail -1 gpurun_out/bench_ours_10.log | cut -c1-1600
