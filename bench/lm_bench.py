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


# ---------------------------------------------------------------------------------------------- one arm
def _pin(t):
    return t.pin_memory() if torch.cuda.is_available() else t


def run_arm(config: str, impl: str, rank: int, world: int, steps: int, warmup: int, e2e: bool = False) -> dict:
    """Time `steps` optimizer steps of `config` on the calling rank's GPU (process group already set up by the caller:
    ddpx.ddp_setup / setup_from_env for impl='ours', any NCCL group for impl='torch').  Device-timed, max over ranks.
    With e2e=True the same steps are timed again with every micro-batch copied from pinned host memory and the loss
    read back to the host (ours: Trainer.train_stream; torch: .to(device) / .item())."""
    import distributed_data_parallel_experiment_b200 as ddpx
    from distributed_data_parallel_experiment_b200 import _lib
    from distributed_data_parallel_experiment_b200.ops import functional as FX
    cfg = CONFIGS[config]
    torch.manual_seed(1 + rank)
    dtype = torch.bfloat16
    accum = cfg.get("accum", 1)
    g = torch.Generator().manual_seed(100 + rank)
    if cfg["kind"] == "mlp":
        W, B = cfg["width"], cfg["batch"]
        hxs = [_pin(torch.randn(B, W, generator=g).to(dtype)) for _ in range(4)]
        hys = [_pin(torch.randint(0, W, (B,), generator=g)) for _ in range(4)]
        unit, per_step = "samples/s", B * world
        if impl == "ours":
            model = ddpx.models.MLP(W, W, cfg["depth"], W, dtype=dtype).cuda()
            loss_fn = ddpx.ops.cross_entropy
        else:
            model = TorchMLP(W, cfg["depth"]).cuda().to(dtype)
            loss_fn = lambda o, t: F.cross_entropy(o.float(), t)
        opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
    else:
        gcfg = getattr(ddpx.models.GPT2Config, cfg["size"])()
        B, T = cfg["batch"], min(cfg["seq"], gcfg.n_positions)
        hxs = [_pin(torch.randint(0, gcfg.vocab_size, (B, T), generator=g)) for _ in range(4)]
        hys = [_pin(torch.randint(0, gcfg.vocab_size, (B, T), generator=g)) for _ in range(4)]
        unit, per_step = "tokens/s", B * T * world * accum
        if impl == "ours":
            model = ddpx.models.GPT2(gcfg, dtype=dtype).cuda()
            loss_fn = lambda o, t: FX.cross_entropy_padded(o, t, gcfg.vocab_size)
        else:
            model = TorchGPT2(gcfg).cuda().to(dtype)
            loss_fn = lambda o, t: F.cross_entropy(o[..., :gcfg.vocab_size].reshape(-1, gcfg.vocab_size).float(), t.reshape(-1))
        opt = torch.optim.AdamW(model.parameters(), lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1,
                                fused=True if impl == "torch" else None)
    xs = [x.cuda() for x in hxs]; ys = [y.cuda() for y in hys]
    nparams = sum(p.numel() for p in model.parameters())
    trainer = None
    if impl == "ours":
        trainer = ddpx.Trainer(model, [], opt, torch.cuda.current_device(), 1, loss_fn=loss_fn, distributed=world > 1,
                               accum_steps=accum)
        step = lambda i: trainer._run_batch(xs[i % 4], ys[i % 4])
    else:
        net = nn.parallel.DistributedDataParallel(model, device_ids=[torch.cuda.current_device()]) if world > 1 else model
        micro = {"i": 0}
        import contextlib

        def step(i, x=None, y=None):
            last = (micro["i"] + 1) % accum == 0
            if micro["i"] % accum == 0:
                opt.zero_grad()
            ctx = net.no_sync() if (world > 1 and not last) else contextlib.nullcontext()
            with ctx:
                loss = loss_fn(net(xs[i % 4] if x is None else x), ys[i % 4] if y is None else y)
                (loss / accum).backward()
            micro["i"] += 1
            if last:
                opt.step()
            return loss

    def bracket(fn):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
        return ms

    n_calls = steps * accum
    for i in range(warmup * accum):
        step(i)
    l0 = _lib.launches()
    ms = bracket(lambda: [step(i) for i in range(n_calls)])
    launches = _lib.launches() - l0
    out = {"config": config, "impl": impl, "n_gpus": world, "value": per_step * steps / (ms / 1e3), "unit": unit,
           "ms_per_step": ms / steps, "params": nparams, "dtype": "bf16", "gpu_launches": launches,
           "global_batch": cfg["batch"] * world, "seq_len": cfg.get("seq"), "accum": accum,
           "approx_model_tflops_per_gpu": 6.0 * nparams * per_step * steps / (ms / 1e3) / world / 1e12,
           "max_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}
    if e2e:
        if impl == "ours":
            def e2e_run(n):
                losses = list(trainer.train_stream((hxs[i % 4], hys[i % 4]) for i in range(n)))
                assert len(losses) == n and all(l == l for l in losses)
                return losses
            e2e_run(accum)
            holder = {}
            ems = bracket(lambda: holder.setdefault("l", e2e_run(n_calls)))
            out["loss"] = holder["l"][-1]
        else:
            def e2e_run(n):
                last = None
                for i in range(n):
                    x = hxs[i % 4].to("cuda", non_blocking=True); y = hys[i % 4].to("cuda", non_blocking=True)
                    last = float(step(i, x, y).item())
                return last
            e2e_run(accum)
            holder = {}
            ems = bracket(lambda: holder.setdefault("l", e2e_run(n_calls)))
            out["loss"] = holder["l"]
        out["e2e_value"] = per_step * steps / (ems / 1e3)
        out["e2e_ms_per_step"] = ems / steps
        out["h2d_bytes_per_step"] = (hxs[0].numel() * hxs[0].element_size() + hys[0].numel() * hys[0].element_size()) * accum
        out["d2h_bytes_per_step"] = 4 * accum
    if impl == "ours" and world > 1:
        from distributed_data_parallel_experiment_b200.utils import debug as _debug
        out["replicas_in_sync"] = bool(_debug.replicas_in_sync(model.parameters()))
    # drop the arm's memory before the next one runs in this process
    del model, opt, xs, ys
    if trainer is not None:
        if hasattr(trainer.model, "close"):
            trainer.model.close()
        del trainer
    import gc
    gc.collect(); torch.cuda.empty_cache()
    return out


def worker(rank, world, port, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    import distributed_data_parallel_experiment_b200 as ddpx
    torch.cuda.set_device(rank)
    if world > 1:
        if args.impl == "ours":
            ddpx.ddp_setup(rank, world, backend="p2p")
        else:
            dist.init_process_group("nccl", rank=rank, world_size=world)
    out = run_arm(args.config, args.impl, rank, world, args.steps, args.warmup, e2e=args.e2e)
    if rank == 0:
        print(json.dumps(out), flush=True)
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
    ap.add_argument("--e2e", action="store_true", help="also time the steps end to end (pinned H2D every micro-batch, loss read back)")
    args = ap.parse_args()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    if args.gpus == 1:
        worker(0, 1, port, args)
    else:
        mp.spawn(worker, args=(args.gpus, port, args), nprocs=args.gpus, join=True)


if __name__ == "__main__":
    main()
