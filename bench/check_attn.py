"""Micro-timings of the attention building blocks at GPT-2-small shape (B=8, H=12, T=1024, hd=64, bf16)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from distributed_data_parallel_experiment_b200.ops import native

B, H, T, hd = 8, 12, 1024, 64
C = H * hd; C3 = 3 * C
qkv = torch.randn(B, T, C3, device="cuda").bfloat16()
q = qkv.view(-1); k = q[C:]; v = q[2 * C:]
P = torch.empty(B, H, T, T, device="cuda", dtype=torch.bfloat16)
out = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16)
in_bs = (hd, T * C3); p_bs = (T * T, H * T * T); o_bs = (hd, T * C)


def t(name, fn, flops=0, bytes_=0, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{name:34s} {ms * 1e3:9.1f} us  {flops / ms / 1e9:8.1f} TFLOP/s  {bytes_ / ms / 1e6:8.1f} GB/s", flush=True)


fl = 2.0 * B * H * T * T * hd
for causal in (0, 1):
    t(f"S=QK^T causal={causal}", lambda: native.gemm_batched(q, k, P, T, T, hd, C3, C3, T, 0, 0, H, B, in_bs, in_bs, p_bs,
                                                          alpha=0.125, causal=causal), fl, P.numel() * 2)
t("softmax fwd", lambda: native.causal_softmax_fwd(P, B * H * T, T), 0, P.numel() * 2 * 1.5)
for causal in (0, 2):
    t(f"O=PV causal={causal}", lambda: native.gemm_batched(P, v, out, T, hd, T, T, C3, C, 0, 1, H, B, p_bs, in_bs, o_bs,
                                                        causal=causal), fl, P.numel() * 2)
dP = torch.empty_like(P)
t("softmax bwd", lambda: native.causal_softmax_bwd(P, dP, B * H * T, T, 0.125), 0, P.numel() * 2 * 2.5)
x = torch.randn(B * T, C, device="cuda").bfloat16()
for (M, N, K) in [(8192, 3072, 768), (8192, 768, 3072), (8192, 768, 768), (8192, 2304, 768)]:
    w = torch.randn(N, K, device="cuda").bfloat16(); xx = torch.randn(M, K, device="cuda").bfloat16()
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16); bias = torch.zeros(N, device="cuda")
    t(f"gemm {M}x{N}x{K} bias", lambda: native.gemm(xx, w, y, M, N, K, K, K, N, 0, 0, bias, 0), 2.0 * M * N * K)
    t(f"torch {M}x{N}x{K}", lambda: torch.nn.functional.linear(xx, w), 2.0 * M * N * K)
from torch.nn.functional import scaled_dot_product_attention as sdpa
qq = torch.randn(B, H, T, hd, device="cuda").bfloat16()
t("torch sdpa fwd", lambda: sdpa(qq, qq, qq, is_causal=True), 2 * fl)
