"""Numerics + timing of the fused causal attention (csrc/kernels/attn_sm100.cu) against torch SDPA.
python bench/check_flash.py [fwd|all]"""
import math, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.nn.functional as F
from distributed_data_parallel_experiment_b200.ops import native

what = sys.argv[1] if len(sys.argv) > 1 else "all"


def ref(qkv, H, dout=None):
    B, T, C3 = qkv.shape
    C = C3 // 3
    x = qkv.detach().double().requires_grad_()
    q, k, v = x.split(C, dim=2)
    q = q.view(B, T, H, 64).transpose(1, 2); k = k.view(B, T, H, 64).transpose(1, 2); v = v.view(B, T, H, 64).transpose(1, 2)
    s = (q @ k.transpose(-1, -2)) / 8.0
    mask = torch.ones(T, T, device=qkv.device, dtype=torch.bool).tril()
    s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    o = (torch.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B, T, C)
    g = None
    if dout is not None:
        o.backward(dout.double())
        g = x.grad
    return o.detach(), lse.detach(), g


def rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def run(B, T, H, bwd):
    torch.manual_seed(0)
    C = H * 64
    qkv = torch.randn(B, T, 3 * C, device="cuda").bfloat16()
    out = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16)
    Tp = (T + 127) // 128 * 128; lse2 = torch.zeros(B, H, Tp, device="cuda", dtype=torch.float32)
    native.attn_fwd(qkv, out, lse2, B, T, H, 0.125)
    torch.cuda.synchronize()
    dout = torch.randn(B, T, C, device="cuda").bfloat16() if bwd else None
    o_ref, lse_ref, g_ref = ref(qkv, H, dout)
    msg = f"B{B} T{T} H{H}: out rel {rel(out, o_ref):.3e}  lse rel {rel(lse2[:, :, :T] / math.log2(math.e), lse_ref):.3e}"
    if bwd:
        dqkv = torch.zeros_like(qkv)
        delta = torch.empty(B, H, (T + 127) // 128 * 128, device="cuda", dtype=torch.float32)
        native.attn_bwd(qkv, out, dout, lse2, delta, dqkv, B, T, H, 0.125)
        torch.cuda.synchronize()
        dq, dk, dv = dqkv.split(C, dim=2); rq, rk, rv = g_ref.split(C, dim=2)
        msg += f"  dq {rel(dq, rq):.3e} dk {rel(dk, rk):.3e} dv {rel(dv, rv):.3e}"
    print(msg, " err word", hex(native.attn_error_word()), flush=True)


def timeit(name, fn, flops, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{name:28s} {ms * 1e3:9.1f} us  {flops / ms / 1e9:8.1f} TFLOP/s", flush=True)


bwd = what != "fwd"
for (B, T, H) in [(1, 128, 1), (1, 256, 2), (2, 200, 2), (2, 1024, 3), (1, 2048, 2)]:
    try:
        run(B, T, H, bwd)
    except Exception as e:
        print(f"B{B} T{T} H{H}: FAILED {type(e).__name__}: {e}", flush=True)
        raise

for (B, T, H) in [(8, 1024, 12), (2, 1024, 25)]:
    C = H * 64
    qkv = torch.randn(B, T, 3 * C, device="cuda").bfloat16()
    out = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16)
    Tp = (T + 127) // 128 * 128; lse2 = torch.zeros(B, H, Tp, device="cuda", dtype=torch.float32)
    fl = 4.0 * B * H * T * T * 64 / 2
    timeit(f"flash fwd B{B} H{H}", lambda: native.attn_fwd(qkv, out, lse2, B, T, H, 0.125), fl)
    q4 = torch.randn(B, H, T, 64, device="cuda").bfloat16().requires_grad_()
    timeit(f"torch sdpa fwd B{B} H{H}", lambda: F.scaled_dot_product_attention(q4, q4, q4, is_causal=True), fl)
    if bwd:
        dout = torch.randn(B, T, C, device="cuda").bfloat16()
        dqkv = torch.empty_like(qkv); delta = torch.empty(B, H, (T + 127) // 128 * 128, device="cuda", dtype=torch.float32)
        timeit(f"flash bwd B{B} H{H}", lambda: native.attn_bwd(qkv, out, dout, lse2, delta, dqkv, B, T, H, 0.125), 2.5 * fl)
        o4 = F.scaled_dot_product_attention(q4, q4, q4, is_causal=True)
        g4 = torch.randn_like(o4)
        timeit(f"torch sdpa bwd B{B} H{H}", lambda: torch.autograd.grad(o4, q4, g4, retain_graph=True), 2.5 * fl)
