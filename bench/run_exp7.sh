#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/check_tc.py gemm 2>&1 | cut -c1-220 > gpurun_out/check_gemm_big.log
DDPX_GEMM_BIGBOX=0 timeout 300 python bench/check_tc.py gemm_big 2>&1 | cut -c1-220 > gpurun_out/check_gemm_small.log
timeout 900 python -m pytest tests/ -m gpu -q -x 2>&1 | grep -v Warning | tail -8 > gpurun_out/pytest_gpu_10.log
timeout 600 python bench/lm_bench.py --config gpt2-small --impl ours > gpurun_out/lm_3.log 2>&1
timeout 600 python bench/lm_bench.py --config mlp --impl ours >> gpurun_out/lm_3.log 2>&1
timeout 300 python bench/profile_lm.py gpt2-small > gpurun_out/profile_lm_gpt2_4.log 2>&1
grep "TFLOP\|rc=[1-9]" gpurun_out/check_gemm_big.log | awk '{print $1,$2,$7,$8,$9,$10}'; grep "TFLOP" gpurun_out/check_gemm_small.log | awk '{print "smallbox",$1,$2,$7,$8,$9,$10}'; tail -3 gpurun_out/pytest_gpu_10.log; grep "^{" gpurun_out/lm_3.log | cut -c1-250
