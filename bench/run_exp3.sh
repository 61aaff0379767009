#!/bin/bash
mkdir -p gpurun_out
timeout 120 ./build/mma_probe > gpurun_out/mma_probe_2.log 2>&1
timeout 300 python bench/check_tc.py conv_w 2>&1 | cut -c1-200 > gpurun_out/check_w_merged.log
for d in 0 1 3; do
  DDPX_DBG=$d timeout 300 python bench/check_tc.py conv_f_vgg 2>&1 | cut -c1-200 | sed "s/^/dbg$d /" >> gpurun_out/fprop_dbg.log
done
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -x 2>&1 | grep -v Warning | tail -15 > gpurun_out/pytest_kernels_9.log
timeout 300 python bench.py > gpurun_out/bench_ours_6.log 2>&1
timeout 300 python bench/profile_step.py 5 > gpurun_out/profile_step_7.log 2>&1
cat gpurun_out/mma_probe_2.log; cat gpurun_out/check_w_merged.log; awk '{print $1,$2,$10,$11,$12,$13,$14}' gpurun_out/fprop_dbg.log; tail -3 gpurun_out/pytest_kernels_9.log; tail -1 gpurun_out/bench_ours_6.log | cut -c1-300
