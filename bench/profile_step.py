"""Per-entry-point device time of one VGG training step (CUDA events around every native call, after warm-up).
    python bench/profile_step.py [steps]
Prints a table sorted by total time; the sum is compared with the event-timed whole step."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200.ops import native

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
torch.manual_seed(0)
objs = ddpx.recipes.vgg_train_objs(world_size=1, synthetic=True, synthetic_size=0)
tr = ddpx.Trainer(objs.model, [], objs.optimizer, 0, 1, scheduler=objs.scheduler, distributed=False)
x = torch.rand(512, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (512,), device="cuda")
for _ in range(5):
    tr._run_batch(x, y)
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    tr._run_batch(x, y)
e1.record(); torch.cuda.synchronize()
print(f"whole step (no per-call events): {e0.elapsed_time(e1) / steps:.3f} ms")
native.profile_begin()
for _ in range(steps):
    tr._run_batch(x, y)
agg = native.profile_end()
tot = 0.0
for name, (calls, ms) in agg.items():
    print(f"{ms / steps * 1e3:10.1f} us/step  {calls // steps:3d} calls  {name}")
    tot += ms
print(f"sum of native calls: {tot / steps:.3f} ms/step")
