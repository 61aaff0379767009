"""Bring-up diagnostics for the tcgen05 GEMM / conv kernels: every case runs in its own process (a trapped
kernel poisons the CUDA context) under a timeout, and prints max-abs / relative error vs a torch fp32 oracle.

    python bench/check_tc.py            # run all cases
    python bench/check_tc.py gemm_tn_tf32   # one case, in-process
"""
from __future__ import annotations

import ctypes
import subprocess
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def _lib():
    from distributed_data_parallel_experiment_b200 import _lib as L
    return L


def run_gemm(a_mn, b_mn, bf16, M=512, N=768, K=1024, out_bf16=0, bias=False, act=0, split_k=1, timing=False):
    import torch
    L = _lib()
    lib = L.lib()
    torch.manual_seed(0)
    dt = torch.bfloat16 if bf16 else torch.float32
    A = torch.randn(M, K, device="cuda", dtype=torch.float32)
    B = torch.randn(N, K, device="cuda", dtype=torch.float32)
    if bf16:
        A = A.to(dt); B = B.to(dt)
    ref = A.double() @ B.double().t()
    bias_t = None
    if bias:
        bias_t = torch.randn(N, device="cuda", dtype=torch.float32)
        ref = ref + bias_t.double()
    if act == 1:
        ref = ref.clamp_min(0)
    elif act == 2:
        ref = torch.nn.functional.gelu(ref, approximate="tanh")
    A_s = A.t().contiguous() if a_mn else A.contiguous()   # [K][M] if MN-major
    B_s = B.t().contiguous() if b_mn else B.contiguous()
    lda = A_s.stride(0); ldb = B_s.stride(0)
    D = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16 if out_bf16 else torch.float32)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def call():
        return lib.ddpx_gemm(L.ptr(A_s), L.ptr(B_s), L.ptr(D), M, N, K, ctypes.c_longlong(lda), ctypes.c_longlong(ldb),
                             ctypes.c_longlong(N), a_mn, b_mn, bf16, out_bf16, L.ptr(bias_t), act, split_k, 0, st)
    rc = call()
    torch.cuda.synchronize()
    err = (D.double() - ref).abs().max().item()
    scale = ref.abs().max().item()
    msg = f"rc={rc} max_abs_err={err:.4e} ref_max={scale:.3e} rel={err / scale:.3e}"
    if timing and rc == 0:
        for _ in range(3):
            D.zero_(); call()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        msg += f" | {ms * 1e3:.1f} us, {2.0 * M * N * K / ms / 1e9:.1f} TFLOP/s"
    return msg


def run_conv(Nimg, H, W, Cin, Cout, wgrad=False, timing=False):
    import torch
    L = _lib()
    lib = L.lib()
    torch.manual_seed(0)
    x = torch.randn(Nimg, Cin, H, W, device="cuda")
    w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.1
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    x_nhwc = x.permute(0, 2, 3, 1).contiguous()
    torch.backends.cudnn.allow_tf32 = False
    if not wgrad:
        ref = torch.nn.functional.conv2d(x.double(), w.double(), padding=1).permute(0, 2, 3, 1).contiguous()
        wp = w.permute(2, 3, 0, 1).contiguous()   # [3][3][Cout][Cin]
        y = torch.empty(Nimg, H, W, Cout, device="cuda")
        ssum = torch.zeros(Cout, device="cuda"); ssq = torch.zeros(Cout, device="cuda")

        def call():
            return lib.ddpx_conv3x3_fprop(L.ptr(x_nhwc), L.ptr(wp), L.ptr(y), Nimg, H, W, Cin, Cout, None, 0,
                                          L.ptr(ssum), L.ptr(ssq), st)
        rc = call(); torch.cuda.synchronize()
        err = (y.double() - ref).abs().max().item(); scale = ref.abs().max().item()
        serr = (ssum.double() - ref.sum((0, 1, 2))).abs().max().item() / max(ref.sum((0, 1, 2)).abs().max().item(), 1e-9)
        qerr = (ssq.double() - (ref * ref).sum((0, 1, 2))).abs().max().item() / (ref * ref).sum((0, 1, 2)).abs().max().item()
        msg = f"rc={rc} max_abs_err={err:.4e} ref_max={scale:.3e} rel={err / scale:.3e} stat_rel={serr:.2e}/{qerr:.2e}"
        flops = 2.0 * Nimg * H * W * Cout * Cin * 9
    else:
        dy = torch.randn(Nimg, Cout, H, W, device="cuda")
        ref = torch.nn.grad.conv2d_weight(x.double(), w.shape, dy.double(), padding=1)
        dy_nhwc = dy.permute(0, 2, 3, 1).contiguous()
        dw = torch.zeros(Cout, Cin, 3, 3, device="cuda")

        def call():
            return lib.ddpx_conv3x3_wgrad(L.ptr(dy_nhwc), L.ptr(x_nhwc), L.ptr(dw), Nimg, H, W, Cin, Cout, 0, st)
        rc = call(); torch.cuda.synchronize()
        err = (dw.double() - ref).abs().max().item(); scale = ref.abs().max().item()
        msg = f"rc={rc} max_abs_err={err:.4e} ref_max={scale:.3e} rel={err / scale:.3e}"
        flops = 2.0 * Nimg * H * W * Cout * Cin * 9
    if timing and rc == 0:
        for _ in range(2):
            call()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        msg += f" | {ms * 1e3:.1f} us, {flops / ms / 1e9:.1f} TFLOP/s"
        # cuDNN TF32 for comparison
        torch.backends.cudnn.allow_tf32 = True
        xc = x.contiguous(memory_format=torch.channels_last); wc = w.contiguous(memory_format=torch.channels_last)
        if not wgrad:
            f = lambda: torch.nn.functional.conv2d(xc, wc, padding=1)
        else:
            dyc = dy.contiguous(memory_format=torch.channels_last)
            f = lambda: torch.nn.grad.conv2d_weight(xc, w.shape, dyc, padding=1)
        for _ in range(3):
            f()
        torch.cuda.synchronize(); e0.record()
        for _ in range(5):
            f()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        msg += f" | cudnn {ms * 1e3:.1f} us, {flops / ms / 1e9:.1f} TFLOP/s"
    return msg


CASES = {
    "gemm_tn_bf16": lambda: run_gemm(0, 0, 1),
    "gemm_tn_tf32": lambda: run_gemm(0, 0, 0),
    "gemm_nn_bf16": lambda: run_gemm(0, 1, 1),     # B MN-major
    "gemm_nn_tf32": lambda: run_gemm(0, 1, 0),
    "gemm_nt_bf16": lambda: run_gemm(1, 1, 1),     # both MN-major (wgrad form)
    "gemm_nt_tf32": lambda: run_gemm(1, 1, 0),
    "gemm_tn_bf16_epi": lambda: run_gemm(0, 0, 1, M=300, N=1000, K=520, out_bf16=1, bias=True, act=2),
    "gemm_tn_tf32_splitk": lambda: run_gemm(0, 0, 0, M=256, N=256, K=4096, split_k=4),
    "gemm_small_n": lambda: run_gemm(0, 0, 0, M=512, N=10, K=512, bias=True),
    "gemm_big_bf16": lambda: run_gemm(0, 0, 1, M=8192, N=8192, K=8192, out_bf16=1, timing=True),
    "gemm_big_tf32": lambda: run_gemm(0, 0, 0, M=4096, N=4096, K=4096, timing=True),
    "gemm_mlp_bf16": lambda: run_gemm(0, 0, 1, M=256, N=8192, K=8192, out_bf16=1, timing=True),
    "gemm_big_tf32_nn": lambda: run_gemm(0, 1, 0, M=4096, N=4096, K=4096, timing=True),
    "gemm_big_tf32_nt": lambda: run_gemm(1, 1, 0, M=4096, N=4096, K=4096, timing=True),
    "gemm_big_bf16_nt": lambda: run_gemm(1, 1, 1, M=8192, N=8192, K=8192, out_bf16=1, timing=True),
    "conv_f_32": lambda: run_conv(4, 32, 32, 64, 128),
    "conv_f_16": lambda: run_conv(4, 16, 16, 128, 256),
    "conv_f_8": lambda: run_conv(8, 8, 8, 256, 512),
    "conv_f_4": lambda: run_conv(32, 4, 4, 512, 512),
    "conv_f_c64": lambda: run_conv(4, 32, 32, 128, 64),
    "conv_w_32": lambda: run_conv(4, 32, 32, 64, 128, wgrad=True),
    "conv_w_16": lambda: run_conv(4, 16, 16, 128, 256, wgrad=True),
    "conv_w_8": lambda: run_conv(8, 8, 8, 256, 512, wgrad=True),
    "conv_w_4": lambda: run_conv(32, 4, 4, 512, 512, wgrad=True),
    "conv_f_vgg1": lambda: run_conv(512, 32, 32, 64, 128, timing=True),
    "conv_f_vgg3": lambda: run_conv(512, 16, 16, 256, 256, timing=True),
    "conv_f_vgg5": lambda: run_conv(512, 8, 8, 512, 512, timing=True),
    "conv_f_vgg7": lambda: run_conv(512, 4, 4, 512, 512, timing=True),
    "conv_x_vgg2": lambda: run_conv(512, 16, 16, 128, 256, timing=True),        # the narrower / transition layers
    "conv_x_vgg4": lambda: run_conv(512, 8, 8, 256, 512, timing=True),
    "conv_x_dgrad1": lambda: run_conv(512, 32, 32, 128, 64, timing=True),
    "conv_x_dgrad2": lambda: run_conv(512, 16, 16, 256, 128, timing=True),
    "conv_x_dgrad4": lambda: run_conv(512, 8, 8, 512, 256, timing=True),
    "conv_w_vgg1": lambda: run_conv(512, 32, 32, 64, 128, wgrad=True, timing=True),
    "conv_w_vgg3": lambda: run_conv(512, 16, 16, 256, 256, wgrad=True, timing=True),
    "conv_w_vgg5": lambda: run_conv(512, 8, 8, 512, 512, wgrad=True, timing=True),
    "conv_w_vgg7": lambda: run_conv(512, 4, 4, 512, 512, wgrad=True, timing=True),
}


def main():
    if len(sys.argv) > 1 and sys.argv[1] in CASES:
        print(sys.argv[1], CASES[sys.argv[1]](), flush=True)
        return
    names = [n for n in CASES if len(sys.argv) < 2 or sys.argv[1] in n]
    for n in names:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, __file__, n], capture_output=True, text=True, timeout=120)
            out = (r.stdout.strip().splitlines() or ["<no output>"])[-1]
            if r.returncode != 0:
                out += f" [exit {r.returncode}] " + " | ".join(r.stderr.strip().splitlines()[-3:])
        except subprocess.TimeoutExpired:
            out = f"{n} TIMEOUT"
        print(f"{out}   ({time.time() - t0:.1f}s)", flush=True)


if __name__ == "__main__":
    main()
