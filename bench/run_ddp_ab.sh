#!/bin/bash
# 2-GPU A/B of the communication-kernel shape (co-residency with the persistent conv kernels) and of the per-forward
# buffer broadcast.  Logs land in gpurun_out/.
mkdir -p gpurun_out; rm -f gpurun_out/ddp_ab.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 $TR --master-port $((29800 + RANDOM % 100)) bench.py --gpus 2 --steps 30 --warmup 5 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); c=d['config']
print('$name', d['ms_per_step'], d['value'], 'e2e', d['e2e']['ms_per_step'], 'exposed', c.get('exposed_comm_ms_per_step'), 'host', c.get('host_enqueue_ms_per_step'))" >> gpurun_out/ddp_ab.log 2>&1
}
timeout 600 python -m pytest tests/test_gpu_p2p.py -q 2>&1 | grep -v Warning | tail -5 > gpurun_out/pytest_p2p_7.log
run K_t512_c64 DDPX_COMM_THREADS=512 DDPX_COMM_CTAS=64
run A_t512_c32 DDPX_COMM_THREADS=512 DDPX_COMM_CTAS=32
run M_t512_c64_cap8 DDPX_COMM_THREADS=512 DDPX_COMM_CTAS=64 DDPX_BUCKET_CAP_MB=8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 bench/trace_step.py > gpurun_out/trace_w2_b.log 2>&1
rm -f gpurun_out/trace_w*_rank0.json
tail -3 gpurun_out/pytest_p2p_7.log; grep "^#" gpurun_out/trace_w2_b.log
cat gpurun_out/ddp_ab.log
