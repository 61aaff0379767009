#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/check_tc.py conv_w > gpurun_out/check_w_big.log 2>&1
DDPX_WGRAD_BIGBOX=0 timeout 300 python bench/check_tc.py conv_w > gpurun_out/check_w_small.log 2>&1
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -x 2>&1 | grep -v Warning | tail -15 > gpurun_out/pytest_kernels_8.log
timeout 300 python bench.py > gpurun_out/bench_ours_5.log 2>&1
timeout 200 python bench/ncu_target.py wgrad > gpurun_out/plain_wgrad.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_kernel -s 1 -c 1 -f -o gpurun_out/prof_wgrad python bench/ncu_target.py wgrad > gpurun_out/ncu_wgrad.log 2>&1
cut -c1-200 gpurun_out/check_w_big.log; cut -c1-200 gpurun_out/check_w_small.log; tail -3 gpurun_out/pytest_kernels_8.log; tail -1 gpurun_out/bench_ours_5.log | cut -c1-400; tail -2 gpurun_out/ncu_wgrad.log
