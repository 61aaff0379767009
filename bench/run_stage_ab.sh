#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/profile_step.py 5 > gpurun_out/profile_step_6.log 2>&1
for cap in 0 2 3 4; do
  for f in gemm_big conv_f_vgg conv_w_vgg; do
    DDPX_MAX_STAGES=$cap timeout 300 python bench/check_tc.py $f 2>&1 | cut -c1-200 | sed "s/^/cap$cap /" >> gpurun_out/stage_ab.log
  done
done
cat gpurun_out/stage_ab.log | awk '{print $1,$2,$8,$9,$10,$11}'
