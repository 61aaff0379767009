"""One launch of every non-contraction kernel family (and the fused attention / all-reduce kernels) at the size it has in
the benchmarked steps, for Nsight Compute:
    python bench/ncu_target_families.py                                   # plain run first
    ncu --set full --clock-control none --import-source on -k regex:'bn_|layernorm|ce_|sgd_multi|adamw_multi|attn_|allreduce_fused|pack_conv|unpack_conv|ew_kernel|colsum' \
        -o gpurun_out/prof_families python bench/ncu_target_families.py
Sizes: VGG block 2 (256 ch @ 16x16, batch 512) for BatchNorm, GPT-2 small (8 x 1024 tokens, C 768) for the transformer
kernels, the 27 MB VGG bucket for the all-reduce + SGD kernel (world size 1: no peer traffic, the local phases only).
The script prints the algorithmic bytes of every launch so that DRAM bytes can be set against them."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200.ops import native

torch.manual_seed(0)
dev = "cuda"
algo = {}

# ---- BatchNorm family: 256 channels @ 16x16, batch 512 (134 MB per activation tensor)
N, H, W, C = 512, 16, 16, 256
y = torch.randn(N, H, W, C, device=dev)
stats = torch.stack([y.sum((0, 1, 2)), (y * y).sum((0, 1, 2))]).contiguous()
gamma = torch.rand(C, device=dev) + 0.5; beta = torch.randn(C, device=dev) * 0.1
rm = torch.zeros(C, device=dev); rv = torch.ones(C, device=dev); nbt = torch.zeros((), dtype=torch.int64, device=dev)
mr = torch.empty(2, C, device=dev)
out = torch.empty(N, H, W, C, device=dev); outp = torch.empty(N, H // 2, W // 2, C, device=dev)
lib = ddpx._lib.lib()
import ctypes
P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
act = y.numel() * 4
for pool, o in ((0, out), (1, outp)):
    rc = lib.ddpx_bn_apply_stats(P(y), P(o), P(stats[0]), P(stats[1]), ctypes.c_float(N * H * W), ctypes.c_float(1e-5),
                                 ctypes.c_float(0.1), P(gamma), P(beta), P(rm), P(rv), P(nbt), P(mr), 1, N, H, W, C, 1, pool, st)
    assert rc == 0
    algo[f"bn_apply_stats pool={pool}"] = act + o.numel() * 4
sums = torch.zeros(2, C, device=dev)
dconv = torch.empty_like(y)
for pool, d in ((0, torch.randn_like(out)), (1, torch.randn_like(outp))):
    sums.zero_()
    native.bn_bwd_reduce(y, d, mr[0], mr[1], gamma, beta, N, H, W, C, 1, pool, sums[0], sums[1])
    algo[f"bn_bwd_reduce pool={pool}"] = act + d.numel() * 4
    native.bn_bwd_apply(y, d, mr[0], mr[1], gamma, beta, sums[0], sums[1], N, H, W, C, 1, pool, 1, dconv)
    algo[f"bn_bwd_apply pool={pool}"] = 2 * act + d.numel() * 4

# ---- weight packing / unpacking of all VGG layers
vgg = ddpx.models.VGG().cuda()
ws = ddpx.ops.functional.prepare_convs([getattr(vgg.backbone, f"conv{i}").weight for i in range(8)])
algo["pack_conv_multi"] = sum(p.numel() for n, p in vgg.named_parameters() if "conv" in n) * 4 * 3
w5 = vgg.backbone.conv5.weight
dw = torch.empty_like(w5)
lib.ddpx_unpack_conv_wgrad_rezero(P(ws[5].staging), P(dw), 512, 512, st)
algo["unpack_conv_wgrad_rezero 512x512"] = w5.numel() * 4 * 3

# ---- transformer kernels at GPT-2 small size
M, Cm = 8192, 768
x = torch.randn(M, Cm, device=dev).bfloat16(); dy = torch.randn(M, Cm, device=dev).bfloat16()
g = (torch.rand(Cm, device=dev) + 0.5).bfloat16(); b = torch.zeros(Cm, device=dev).bfloat16()
yy = torch.empty_like(x); sm = torch.empty(2, M, device=dev)
native.layernorm_fwd(x, g, b, yy, sm[0], sm[1], M, Cm, 1e-5)
algo["layernorm_fwd"] = 2 * x.numel() * 2
dgb = torch.zeros(2, Cm, device=dev); dx = torch.empty_like(x)
native.layernorm_bwd(x, dy, g, sm[0], sm[1], dx, dgb[0], dgb[1], M, Cm)
algo["layernorm_bwd"] = 3 * x.numel() * 2
# the same at a GPT-2 XL micro-batch: 2048 x 1600
Mx, Cx = 2048, 1600
xx = torch.randn(Mx, Cx, device=dev).bfloat16(); dyx = torch.randn(Mx, Cx, device=dev).bfloat16()
gx = (torch.rand(Cx, device=dev) + 0.5).bfloat16(); smx = torch.empty(2, Mx, device=dev); yx = torch.empty_like(xx)
native.layernorm_fwd(xx, gx, gx, yx, smx[0], smx[1], Mx, Cx, 1e-5)
dgbx = torch.zeros(2, Cx, device=dev)
native.layernorm_bwd(xx, dyx, gx, smx[0], smx[1], yx, dgbx[0], dgbx[1], Mx, Cx)
algo["layernorm_bwd 2048x1600"] = 3 * xx.numel() * 2
h = torch.randn(M, 4 * Cm, device=dev).bfloat16(); hh = torch.empty_like(h)
native.gelu_fwd(h, hh); algo["gelu_fwd"] = 2 * h.numel() * 2
native.gelu_bwd(h, hh, hh); algo["gelu_bwd"] = 3 * h.numel() * 2
native.add(x, dy, yy); algo["add"] = 3 * x.numel() * 2
db = torch.empty(4 * Cm, device=dev)
native.colsum_any(h, db, M, 4 * Cm); algo["colsum_bf16"] = h.numel() * 2
V = 50304
logits = torch.randn(M, V, device=dev).bfloat16(); tgt = torch.randint(0, 50257, (M,), device=dev)
lse = torch.empty(M, device=dev); loss = torch.empty(1, device=dev); dl = torch.empty_like(logits)
native.ce_fwd(logits, tgt, lse, loss, M, 50257, V); algo["ce_fwd"] = logits.numel() * 2
native.ce_bwd(logits, tgt, lse, torch.ones((), device=dev), dl, M, 50257, V); algo["ce_bwd"] = 2 * logits.numel() * 2

# ---- fused attention, GPT-2 small: B 8, H 12, T 1024
B, T, Hh = 8, 1024, 12
qkv = torch.randn(B, T, 3 * Hh * 64, device=dev).bfloat16()
o = torch.empty(B, T, Hh * 64, device=dev, dtype=torch.bfloat16)
lse2 = torch.empty(B, Hh, T, device=dev); delta = torch.empty_like(lse2); dqkv = torch.empty_like(qkv)
native.attn_fwd(qkv, o, lse2, B, T, Hh, 0.125)
native.attn_bwd(qkv, o, torch.randn_like(o), lse2, delta, dqkv, B, T, Hh, 0.125)
algo["attn_fwd (flops)"] = 4.0 * B * Hh * T * T * 64 / 2
algo["attn_bwd dq+dkv (flops)"] = 2.5 * 4.0 * B * Hh * T * T * 64 / 2

# ---- optimizers: multi-tensor SGD over the VGG parameters, AdamW over a 50 M-element bf16 tensor
ps = [p for p in vgg.parameters()]
gs = [torch.randn_like(p) for p in ps]; bufs = [torch.zeros_like(p) for p in ps]
native.sgd_multi([p.detach() for p in ps], gs, bufs, 0.1, 0.9, 0.0, 5e-4, False, False)
algo["sgd_multi (VGG)"] = sum(p.numel() for p in ps) * 4 * 5
pa = torch.randn(50_000_000, device=dev).bfloat16(); ga = torch.randn_like(pa); m1 = torch.zeros_like(pa); m2 = torch.zeros_like(pa)
native.adamw_multi([pa], [ga], [m1], [m2], 3e-4, 0.9, 0.95, 1e-8, 0.1, 1)
algo["adamw_multi (50M bf16)"] = pa.numel() * 2 * 7

# ---- fused all-reduce + SGD kernel on the 27 MB VGG bucket (one rank: local phases only)
from distributed_data_parallel_experiment_b200.parallel.p2p import P2PComm
comm = P2PComm(staging_mb=8)
n = 27148288 // 4
buf = comm.alloc(n, torch.float32, None); buf.normal_()
prm = torch.randn(n, device=dev); mom = torch.zeros(n, device=dev)
comm.launch_allreduce(buf, mode=0, opt=1, hyper=[0.1, 0.9, 0.0, 5e-4, 0, 0, 0, 1.0, 1.0],
                      slots=([prm.data_ptr()], [mom.data_ptr()], [0], [0], [n])).wait()
algo["allreduce_fused (27 MB, world 1: shard = whole bucket)"] = n * 4 * 5
torch.cuda.synchronize()
assert native.read_error_word() == 0 and native.attn_error_word() == 0 and comm.error_word() == 0
comm.close()
for k, v in algo.items():
    print(f"{k:60s} {v / 1e6:12.1f} M (bytes unless marked flops)")
print("ok")
