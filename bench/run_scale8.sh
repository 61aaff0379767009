#!/bin/bash
# One 8-GPU box: P2P tests, both bench arms at N=8, ours at N=4, all-reduce sweeps, a kernel timeline.  Logs -> gpurun_out/.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_p2p.py -q --tb=short 2>&1 | grep -v Warning | tail -30 > gpurun_out/pytest_p2p_n8_b.log
timeout 300 $TR --nproc-per-node 8 --master-port 29701 bench.py --gpus 8 --steps 50 --warmup 5 > gpurun_out/bench_ours_n8_b.log 2>&1
timeout 300 $TR --nproc-per-node 8 --master-port 29702 bench.py --impl reference --gpus 8 --steps 50 --warmup 5 > gpurun_out/bench_ref_n8_b.log 2>&1
timeout 300 $TR --nproc-per-node 4 --master-port 29703 bench.py --gpus 4 --steps 50 --warmup 5 > gpurun_out/bench_ours_n4_b.log 2>&1
timeout 300 $TR --nproc-per-node 8 --master-port 29704 bench/trace_step.py > gpurun_out/trace_w8.log 2>&1
rm -f gpurun_out/trace_w*_rank0.json
timeout 400 python bench/allreduce_sweep.py --gpus 8 --max-mb 256 --ctas 64 --fused > gpurun_out/sweep_8gpu_b.log 2>&1
tail -2 gpurun_out/pytest_p2p_n8_b.log; for f in bench_ours_n8_b bench_ref_n8_b bench_ours_n4_b; do tail -1 gpurun_out/$f.log | cut -c1-260; done; grep "^#" gpurun_out/trace_w8.log; tail -2 gpurun_out/sweep_8gpu_b.log | cut -c1-200
