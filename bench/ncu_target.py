"""Short, deterministic launch sequence for Nsight Compute captures of the tcgen05 kernels (run plain first, then under
`ncu --set full --clock-control none --import-source on -k regex:tc_kernel ...`, see B200_PROFILING.md):
  3x conv3x3 fprop 256->256 @ 16x16 (batch 512), 3x conv3x3 wgrad (same layer), 3x bf16 GEMM 4096^3.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from distributed_data_parallel_experiment_b200.ops import native

which = sys.argv[1] if len(sys.argv) > 1 else "all"
torch.manual_seed(0)
import os
N, H, W, Cin, Cout = (int(v) for v in os.environ.get("NCU_SHAPE", "512,16,16,256,256").split(","))   # N,H,W,Cin,Cout
x = torch.randn(N, H, W, Cin, device="cuda")
w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.05
wp = torch.empty(9 * Cout * Cin, device="cuda"); wd = torch.empty_like(wp)
native.pack_conv_weights(w, wp, wd, Cout, Cin)
y = torch.empty(N, H, W, Cout, device="cuda")
stats = torch.zeros(2, Cout, device="cuda")
dw = torch.zeros(Cout, Cin, 3, 3, device="cuda")
A = torch.randn(4096, 4096, device="cuda").bfloat16(); B = torch.randn(4096, 4096, device="cuda").bfloat16()
D = torch.empty(4096, 4096, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    if which in ("all", "fprop"):
        native.conv3x3_fprop(x, wp, y, N, H, W, Cin, Cout, None, 0, stats[0], stats[1])
    if which in ("all", "wgrad"):
        native.conv3x3_wgrad(y, x, dw, N, H, W, Cin, Cout)
    if which in ("all", "gemm"):
        native.gemm(A, B, D, 4096, 4096, 4096, 4096, 4096, 4096, 0, 0)
if which == "linear":      # small-K Linear shape of GPT-2 (epilogue-paced)
    xx = torch.randn(8192, 768, device="cuda").bfloat16(); ww = torch.randn(3072, 768, device="cuda").bfloat16()
    yy = torch.empty(8192, 3072, device="cuda", dtype=torch.bfloat16); bb = torch.zeros(3072, device="cuda")
    for _ in range(3):
        native.gemm(xx, ww, yy, 8192, 3072, 768, 768, 768, 3072, 0, 0, bb, 0)
if which == "xl":          # GPT-2 XL block shapes (2048 tokens, C = 1600): forward, data gradient, weight gradient
    M, C = 2048, 1600
    xa = torch.randn(M, C, device="cuda").bfloat16(); x4 = torch.randn(M, 4 * C, device="cuda").bfloat16()
    wfc = torch.randn(4 * C, C, device="cuda").bfloat16(); wpr = torch.randn(C, C, device="cuda").bfloat16()
    y4 = torch.empty(M, 4 * C, device="cuda", dtype=torch.bfloat16); y1 = torch.empty(M, C, device="cuda", dtype=torch.bfloat16)
    dwf = torch.empty(4 * C, C, device="cuda", dtype=torch.bfloat16)
    cases = {
        "fc fwd    [2048x6400x1600] K,K": lambda: native.gemm(xa, wfc, y4, M, 4 * C, C, C, C, 4 * C, 0, 0),
        "proj fwd  [2048x1600x1600] K,K": lambda: native.gemm(xa, wpr, y1, M, C, C, C, C, C, 0, 0),
        "fc dgrad  [2048x1600x6400] K,MN": lambda: native.gemm(x4, wfc, y1, M, C, 4 * C, 4 * C, C, C, 0, 1),
        "fc wgrad  [6400x1600x2048] MN,MN": lambda: native.gemm(x4, xa, dwf, 4 * C, C, M, 4 * C, C, C, 1, 1),
    }
    timing = os.environ.get("NCU_TIMING", "0") == "1"
    for name, fn in cases.items():
        fn()
        if timing:
            for _ in range(5): fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record()
            for _ in range(20): fn()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 50.0
            flops = 2.0 * 2048 * 1600 * (6400 if "6400" in name else 1600)
            print(f"{name:36s} {us:7.1f} us  {flops / us * 1e-6:7.1f} TFLOP/s")
torch.cuda.synchronize()
assert native.read_error_word() == 0
print("ok")
