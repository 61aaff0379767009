"""Small, fast launch sequence for compute-sanitizer:
    compute-sanitizer --tool memcheck python bench/sanitize_target.py
One VGG forward+backward+step at batch 8 (every conv / BN / pool / loss / optimizer kernel, both CTA flavours where the
shapes allow), one bf16 Linear fwd/bwd, LayerNorm, softmax, embedding, and the multi-tensor kernels."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200.ops import native

torch.manual_seed(0)
m = ddpx.models.VGG().cuda()
opt = torch.optim.SGD(m.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4)
tr = ddpx.Trainer(m, [], opt, 0, 1, distributed=False)
x = torch.rand(8, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (8,), device="cuda")
for _ in range(2):
    loss = tr._run_batch(x, y)
print("vgg loss", float(loss))
cfg = ddpx.models.GPT2Config.tiny()
g = ddpx.models.GPT2(cfg, dtype=torch.bfloat16).cuda()
gopt = torch.optim.AdamW(g.parameters(), lr=1e-3)
ddpx.optim.fuse(gopt)
idx = torch.randint(0, cfg.vocab_size, (2, 64), device="cuda")
for _ in range(2):
    gopt.zero_grad()
    l = g(idx, idx)
    l.backward()
    gopt.step()
print("gpt2 loss", float(l))
torch.cuda.synchronize()
assert native.read_error_word() == 0
print("ok")
