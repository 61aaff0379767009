#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_p2p.py -q 2>&1 | grep -v Warning | tail -5 > gpurun_out/pytest_p2p_8.log
timeout 300 $TR --master-port 29811 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/bench_ours_n2_3.log 2>&1
timeout 400 python bench/allreduce_sweep.py --gpus 2 --max-mb 256 --ctas 64 --fused > gpurun_out/sweep_2gpu_3.log 2>&1
tail -3 gpurun_out/pytest_p2p_8.log; tail -1 gpurun_out/bench_ours_n2_3.log | cut -c1-300; tail -3 gpurun_out/sweep_2gpu_3.log | cut -c1-200
