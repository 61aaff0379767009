#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench/check_tc.py 2>&1 | cut -c1-220 > gpurun_out/check_tc_6.log
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -x 2>&1 | grep -v Warning | tail -15 > gpurun_out/pytest_kernels_10.log
timeout 300 python bench.py > gpurun_out/bench_ours_7.log 2>&1
timeout 300 python bench/profile_step.py 5 > gpurun_out/profile_step_8.log 2>&1
grep "TFLOP\|rc=[1-9]" gpurun_out/check_tc_6.log | awk '{print $1,$2,$3,$7,$8,$9,$10,$11,$12,$13,$14,$15}'; tail -3 gpurun_out/pytest_kernels_10.log; tail -1 gpurun_out/bench_ours_7.log | cut -c1-300
