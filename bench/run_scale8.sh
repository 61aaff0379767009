#!/bin/bash
# One 8-GPU box: P2P tests, both bench arms at N=8, ours at N=4, all-reduce sweep.  Logs land in gpurun_out/.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/topo8.log 2>&1
timeout 600 python -m pytest tests/test_gpu_p2p.py -q --tb=long 2>&1 | grep -v Warning | tail -60 > gpurun_out/pytest_p2p_n8.log
timeout 300 $TR --nproc-per-node 8 --master-port 29701 bench.py --gpus 8 --steps 30 --warmup 5 > gpurun_out/bench_ours_n8.log 2>&1
timeout 300 $TR --nproc-per-node 8 --master-port 29702 bench.py --impl reference --gpus 8 --steps 30 --warmup 5 > gpurun_out/bench_ref_n8.log 2>&1
timeout 300 $TR --nproc-per-node 4 --master-port 29703 bench.py --gpus 4 --steps 30 --warmup 5 > gpurun_out/bench_ours_n4.log 2>&1
timeout 400 python bench/allreduce_sweep.py --gpus 8 --max-mb 256 --ctas 32,64 > gpurun_out/sweep_8gpu.log 2>&1
tail -2 gpurun_out/pytest_p2p_n8.log; tail -1 gpurun_out/bench_ours_n8.log; tail -1 gpurun_out/bench_ref_n8.log; tail -1 gpurun_out/bench_ours_n4.log; tail -4 gpurun_out/sweep_8gpu.log
