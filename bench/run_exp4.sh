#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/fprop_dbg2.log
for d in 0 4 8 16; do
  DDPX_DBG=$d timeout 300 python bench/check_tc.py conv_f_vgg 2>&1 | cut -c1-200 | sed "s/^/dbg$d /" >> gpurun_out/fprop_dbg2.log
done
awk '{print $1,$2,$9,$10,$11,$12,$13,$14}' gpurun_out/fprop_dbg2.log
