#!/bin/bash
mkdir -p gpurun_out
timeout 200 python bench/ncu_target.py fprop > gpurun_out/plain_fprop3.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_kernel -s 1 -c 1 -f -o gpurun_out/prof_fprop_final python bench/ncu_target.py fprop > gpurun_out/ncu_fprop3.log 2>&1
tail -2 gpurun_out/ncu_fprop3.log
timeout 300 python bench.py --steps 100 --warmup 5 > gpurun_out/bench_ours_11.log 2>&1
timeout 300 python bench.py --impl reference --steps 100 --warmup 5 > gpurun_out/bench_ref_n1_b.log 2>&1
tail -1 gpurun_out/bench_ours_11.log | cut -c1-250; tail -1 gpurun_out/bench_ref_n1_b.log | cut -c1-250
