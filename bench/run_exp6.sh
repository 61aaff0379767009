#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/check_tc.py conv_w 2>&1 | cut -c1-220 > gpurun_out/check_w_wave1.log
DDPX_WGRAD_WAVES=2 timeout 300 python bench/check_tc.py conv_w_vgg 2>&1 | cut -c1-220 > gpurun_out/check_w_wave2.log
timeout 900 python -m pytest tests/ -m gpu -q -x 2>&1 | grep -v Warning | tail -15 > gpurun_out/pytest_gpu_8.log
timeout 300 python bench.py > gpurun_out/bench_ours_8.log 2>&1
timeout 300 python bench/profile_step.py 5 > gpurun_out/profile_step_9.log 2>&1
grep "TFLOP\|rc=[1-9]" gpurun_out/check_w_wave1.log | awk '{print $1,$2,$3,$7,$8,$9,$10,$11,$12,$13,$14,$15}'; grep "TFLOP" gpurun_out/check_w_wave2.log | awk '{print "waves2",$1,$2,$7,$8,$9,$10}';  tail -3 gpurun_out/pytest_gpu_8.log; tail -1 gpurun_out/bench_ours_8.log | cut -c1-300
