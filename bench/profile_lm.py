"""Per-entry-point device time of one training step of an LM / MLP config (see bench/lm_bench.py for the configs).
    python bench/profile_lm.py gpt2-small [steps]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import torch
import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200.ops import native, functional as FX
from lm_bench import CONFIGS

name = sys.argv[1] if len(sys.argv) > 1 else "gpt2-small"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg = CONFIGS[name]
torch.manual_seed(0)
if cfg["kind"] == "mlp":
    W, B = cfg["width"], cfg["batch"]
    model = ddpx.models.MLP(W, W, cfg["depth"], W, dtype=torch.bfloat16).cuda()
    x = torch.randn(B, W, device="cuda").bfloat16(); y = torch.randint(0, W, (B,), device="cuda")
    loss_fn = ddpx.ops.cross_entropy
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
else:
    g = getattr(ddpx.models.GPT2Config, cfg["size"])()
    model = ddpx.models.GPT2(g, dtype=torch.bfloat16).cuda()
    B, T = cfg["batch"], min(cfg["seq"], g.n_positions)
    x = torch.randint(0, g.vocab_size, (B, T), device="cuda"); y = torch.randint(0, g.vocab_size, (B, T), device="cuda")
    loss_fn = lambda o, t: FX.cross_entropy_padded(o, t, g.vocab_size)
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4)
tr = ddpx.Trainer(model, [], opt, 0, 1, loss_fn=loss_fn, distributed=False)
for _ in range(3):
    tr._run_batch(x, y)
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    tr._run_batch(x, y)
e1.record(); torch.cuda.synchronize()
print(f"whole step: {e0.elapsed_time(e1) / steps:.3f} ms")
native.profile_begin()
for _ in range(steps):
    tr._run_batch(x, y)
agg = native.profile_end()
tot = 0.0
for i, (nm, (calls, ms)) in enumerate(agg.items()):
    if i < 40:
        print(f"{ms / steps * 1e3:10.1f} us/step  {calls // steps:4d} calls  {nm}")
    tot += ms
print(f"sum of native calls: {tot / steps:.3f} ms/step")
