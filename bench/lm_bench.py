"""Throughput of the BASELINE.json configs the reference itself cannot run (it only has a CNN): the 8x8192 bf16 MLP and
GPT-2 small / XL, data-parallel, synthetic data.  Two arms on the same shapes, dtype and optimizer:

  --impl ours    this framework (tcgen05 GEMMs, native LayerNorm/attention/CE/AdamW, P2P fused all-reduce+optimizer)
  --impl torch   the reference's *stack*: torch.nn modules + torch DDP + NCCL (+ cuBLAS / SDPA), same pure-bf16 setup

    python bench/lm_bench.py --config mlp --gpus 1 --steps 10
    python bench/lm_bench.py --config gpt2-small --gpus 2 --impl torch
Prints one JSON line (rank 0): samples/s (mlp) or tokens/s (gpt2), device-timed, max over ranks.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import socket
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn.functional as F
from torch import nn

CONFIGS = {
    "mlp": dict(kind="mlp", width=8192, depth=8, batch=256),
    "mlp-small": dict(kind="mlp", width=2048, depth=4, batch=256),
    "gpt2-tiny": dict(kind="gpt2", size="tiny", batch=4, seq=128, accum=1),
    "gpt2-small": dict(kind="gpt2", size="small", batch=8, seq=1024, accum=1),
    "gpt2-xl": dict(kind="gpt2", size="xl", batch=2, seq=1024, accum=4),
}


# ---------------------------------------------------------------------------------------------- torch-stack models
class TorchMLP(nn.Module):
    def __init__(self, width, depth):
        super().__init__()
        self.layers = nn.ModuleList([nn.Linear(width, width) for _ in range(depth)])

    def forward(self, x):
        for i, l in enumerate(self.layers):
            x = l(x)
            if i < len(self.layers) - 1:
                x = F.relu(x)
        return x


class TorchBlock(nn.Module):
    def __init__(self, C, H):
        super().__init__()
        self.ln_1 = nn.LayerNorm(C); self.ln_2 = nn.LayerNorm(C)
        self.c_attn = nn.Linear(C, 3 * C); self.c_proj = nn.Linear(C, C)
        self.c_fc = nn.Linear(C, 4 * C); self.c_proj2 = nn.Linear(4 * C, C)
        self.H = H

    def forward(self, x):
        B, T, C = x.shape
        q, k, v = self.c_attn(self.ln_1(x)).split(C, dim=2)
        hd = C // self.H
        q = q.view(B, T, self.H, hd).transpose(1, 2); k = k.view(B, T, self.H, hd).transpose(1, 2); v = v.view(B, T, self.H, hd).transpose(1, 2)
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, C)
        x = x + self.c_proj(y)
        return x + self.c_proj2(F.gelu(self.c_fc(self.ln_2(x)), approximate="tanh"))


class TorchGPT2(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.padded_vocab_size, cfg.n_embd); self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        self.h = nn.ModuleList([TorchBlock(cfg.n_embd, cfg.n_head) for _ in range(cfg.n_layer)])
        self.ln_f = nn.LayerNorm(cfg.n_embd)

    def forward(self, idx):
        T = idx.shape[1]
        x = self.wte(idx) + self.wpe.weight[:T]
        for b in self.h:
            x = b(x)
        return F.linear(self.ln_f(x), self.wte.weight)


# ---------------------------------------------------------------------------------------------- worker
def worker(rank, world, port, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    import distributed_data_parallel_experiment_b200 as ddpx
    from distributed_data_parallel_experiment_b200 import _lib
    from distributed_data_parallel_experiment_b200.ops import functional as FX
    cfg = CONFIGS[args.config]
    torch.cuda.set_device(rank)
    if world > 1:
        if args.impl == "ours":
            ddpx.ddp_setup(rank, world, backend="p2p")
        else:
            dist.init_process_group("nccl", rank=rank, world_size=world)
    torch.manual_seed(1 + rank)
    dtype = torch.bfloat16
    accum = cfg.get("accum", 1)
    if cfg["kind"] == "mlp":
        W, B = cfg["width"], cfg["batch"]
        xs = [torch.randn(B, W, device="cuda").to(dtype) for _ in range(4)]
        ys = [torch.randint(0, W, (B,), device="cuda") for _ in range(4)]
        unit, per_step = "samples/s", B * world
        if args.impl == "ours":
            model = ddpx.models.MLP(W, W, cfg["depth"], W, dtype=dtype).cuda()
            loss_fn = ddpx.ops.cross_entropy
        else:
            model = TorchMLP(W, cfg["depth"]).cuda().to(dtype)
            loss_fn = lambda o, t: F.cross_entropy(o.float(), t)
        opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
    else:
        gcfg = getattr(ddpx.models.GPT2Config, cfg["size"])()
        B, T = cfg["batch"], min(cfg["seq"], gcfg.n_positions)
        xs = [torch.randint(0, gcfg.vocab_size, (B, T), device="cuda") for _ in range(4)]
        ys = [torch.randint(0, gcfg.vocab_size, (B, T), device="cuda") for _ in range(4)]
        unit, per_step = "tokens/s", B * T * world * accum
        if args.impl == "ours":
            model = ddpx.models.GPT2(gcfg, dtype=dtype).cuda()
            loss_fn = lambda o, t: FX.cross_entropy_padded(o, t, gcfg.vocab_size)
        else:
            model = TorchGPT2(gcfg).cuda().to(dtype)
            loss_fn = lambda o, t: F.cross_entropy(o[..., :gcfg.vocab_size].reshape(-1, gcfg.vocab_size).float(), t.reshape(-1))
        opt = torch.optim.AdamW(model.parameters(), lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1,
                                fused=True if args.impl == "torch" else None)
    nparams = sum(p.numel() for p in model.parameters())
    if args.impl == "ours":
        trainer = ddpx.Trainer(model, [], opt, rank, 1, loss_fn=loss_fn, distributed=world > 1, accum_steps=accum)
        step = lambda i: trainer._run_batch(xs[i % 4], ys[i % 4])
    else:
        net = nn.parallel.DistributedDataParallel(model, device_ids=[rank]) if world > 1 else model
        micro = {"i": 0}

        def step(i):
            last = (micro["i"] + 1) % accum == 0
            if micro["i"] % accum == 0:
                opt.zero_grad()
            import contextlib
            ctx = net.no_sync() if (world > 1 and not last) else contextlib.nullcontext()
            with ctx:
                (loss_fn(net(xs[i % 4]), ys[i % 4]) / accum).backward()
            micro["i"] += 1
            if last:
                opt.step()
    n_calls = args.steps * accum
    for i in range(args.warmup * accum):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = _lib.launches()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n_calls):
        step(i)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
    if rank == 0:
        flops = 6.0 * nparams * (per_step if cfg["kind"] == "gpt2" else per_step) * args.steps / (ms / 1e3) / world
        print(json.dumps({"config": args.config, "impl": args.impl, "n_gpus": world, "value": per_step * args.steps / (ms / 1e3),
                          "unit": unit, "ms_per_step": ms / args.steps, "params": nparams, "dtype": "bf16",
                          "approx_model_tflops_per_gpu": flops / 1e12, "gpu_launches": _lib.launches() - l0,
                          "max_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}), flush=True)
    if world > 1:
        if args.impl == "ours":
            ddpx.teardown()
        else:
            dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="mlp", choices=list(CONFIGS))
    ap.add_argument("--impl", default="ours", choices=["ours", "torch"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    if args.gpus == 1:
        worker(0, 1, port, args)
    else:
        mp.spawn(worker, args=(args.gpus, port, args), nprocs=args.gpus, join=True)


if __name__ == "__main__":
    main()
