"""Where does the HOST time of a step go?  cProfile over a few steps (diagnostics only; the GPU runs ahead or idles).
    python bench/host_profile.py [vgg|gpt2|gpt2-xl] [steps]"""
import cProfile
import pstats
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import distributed_data_parallel_experiment_b200 as ddpx

name = sys.argv[1] if len(sys.argv) > 1 else "gpt2"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
if name == "vgg":
    objs = ddpx.recipes.vgg_train_objs(world_size=1, synthetic=True, synthetic_size=0)
    tr = ddpx.Trainer(objs.model, [], objs.optimizer, 0, 1, scheduler=objs.scheduler, distributed=False)
    x = torch.rand(512, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (512,), device="cuda")
else:
    objs, loss_fn = ddpx.recipes.make(name, synthetic=True, synthetic_size=0)
    tr = ddpx.Trainer(objs.model, [], objs.optimizer, 0, 1, loss_fn=loss_fn, distributed=False)
    V = objs.model.cfg.vocab_size
    B = 2 if "xl" in name else 8
    x = torch.randint(0, V, (B, 1024), device="cuda"); y = torch.randint(0, V, (B, 1024), device="cuda")
for _ in range(3):
    tr._run_batch(x, y)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(steps):
    tr._run_batch(x, y)
host = (time.perf_counter() - t0) / steps
torch.cuda.synchronize()
total = (time.perf_counter() - t0) / steps
print(f"host enqueue {host * 1e3:.2f} ms/step, with sync {total * 1e3:.2f} ms/step")
pr = cProfile.Profile()
pr.enable()
for _ in range(steps):
    tr._run_batch(x, y)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
