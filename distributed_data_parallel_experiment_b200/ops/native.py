"""Thin ctypes wrappers over ``libddpx_kernels.so`` - one Python function per C entry point.

Every wrapper takes torch CUDA tensors, launches on the current stream, counts the launch (``_lib.launches()``
feeds bench.py's ``gpu_launches``) and raises on a non-zero return code.  No torch kernel is ever used as a
fallback here: callers on a CUDA device either reach these or fail.
"""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_longlong, c_void_p

import torch

from .. import _lib

_P = _lib.ptr


def _st():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _call(name, *args, launches: int = 1):
    fn = getattr(_lib.lib(), name)
    rc = fn(*args)
    _lib.count_launch(launches)
    if rc != 0:
        _lib.check(rc, name)


# ------------------------------------------------------------------------------------------------ GEMM / conv
def gemm(A, B, D, M, N, K, lda, ldb, ldd, a_mn=0, b_mn=0, bias=None, act=0, split_k=1, accumulate=0):
    """D[M,N] (=|+=) op(A) op(B); dtype taken from A (fp32 -> TF32 tensor cores, bf16 -> bf16)."""
    bf16 = 1 if A.dtype == torch.bfloat16 else 0
    out_bf16 = 1 if D.dtype == torch.bfloat16 else 0
    _call("ddpx_gemm", _P(A), _P(B), _P(D), c_int(M), c_int(N), c_int(K), c_longlong(lda), c_longlong(ldb),
          c_longlong(ldd), c_int(a_mn), c_int(b_mn), c_int(bf16), c_int(out_bf16), _P(bias), c_int(act),
          c_int(split_k), c_int(accumulate), _st())


def conv3x3_fprop(x_nhwc, w_packed, y_nhwc, N, H, W, Cin, Cout, bias=None, act=0, stat_sum=None, stat_sq=None):
    _call("ddpx_conv3x3_fprop", _P(x_nhwc), _P(w_packed), _P(y_nhwc), c_int(N), c_int(H), c_int(W), c_int(Cin),
          c_int(Cout), _P(bias), c_int(act), _P(stat_sum), _P(stat_sq), _st())


def conv3x3_wgrad(dy_nhwc, x_nhwc, dw, N, H, W, Cin, Cout, split_k=0):
    _call("ddpx_conv3x3_wgrad", _P(dy_nhwc), _P(x_nhwc), _P(dw), c_int(N), c_int(H), c_int(W), c_int(Cin), c_int(Cout),
          c_int(split_k), _st())


def pack_conv_weights(w, w_fprop, w_dgrad, Cout, Cin):
    _call("ddpx_pack_conv_weights", _P(w), _P(w_fprop), _P(w_dgrad), c_int(Cout), c_int(Cin), _st())


def conv_stem_fprop(x_nchw, w, bias, y_nhwc, N, H, W, Cout, act=0, stat_sum=None, stat_sq=None):
    _call("ddpx_conv_stem_fprop", _P(x_nchw), _P(w), _P(bias), _P(y_nhwc), c_int(N), c_int(H), c_int(W), c_int(Cout),
          c_int(act), _P(stat_sum), _P(stat_sq), _st())


def conv_stem_wgrad(dy_nhwc, x_nchw, dw, N, H, W, Cout):
    _call("ddpx_conv_stem_wgrad", _P(dy_nhwc), _P(x_nchw), _P(dw), c_int(N), c_int(H), c_int(W), c_int(Cout), _st())


def transpose_batched(src, dst, batch, rows, cols):
    _call("ddpx_transpose_batched", _P(src), _P(dst), c_int(batch), c_int(rows), c_int(cols), _st())


# ------------------------------------------------------------------------------------------------ elementwise / pool
def relu_fwd(x, y):
    _call("ddpx_relu_fwd", _P(x), _P(y), c_longlong(x.numel()), _st())


def relu_bwd(y, dy, dx):
    _call("ddpx_relu_bwd", _P(y), _P(dy), _P(dx), c_longlong(y.numel()), _st())


def maxpool_fwd(x, y, N, H, W, C):
    _call("ddpx_maxpool2x2_fwd", _P(x), _P(y), c_int(N), c_int(H), c_int(W), c_int(C), _st())


def maxpool_bwd(x, dy, dx, N, H, W, C):
    _call("ddpx_maxpool2x2_bwd", _P(x), _P(dy), _P(dx), c_int(N), c_int(H), c_int(W), c_int(C), _st())


def spatial_mean_fwd(x, y, N, HW, C):
    _call("ddpx_spatial_mean_fwd", _P(x), _P(y), c_int(N), c_int(HW), c_int(C), _st())


def spatial_mean_bwd(dy, dx, N, HW, C):
    _call("ddpx_spatial_mean_bwd", _P(dy), _P(dx), c_int(N), c_int(HW), c_int(C), _st())


def colsum(x, out, rows, cols, accumulate=0):
    _call("ddpx_colsum", _P(x), _P(out), c_longlong(rows), c_int(cols), c_int(accumulate), _st())


def scale_copy(src, dst, n, scale=1.0, accumulate=0):
    _call("ddpx_scale_copy", _P(src), _P(dst), c_longlong(n), c_float(scale), c_int(accumulate), _st())


def zero_(t):
    _call("ddpx_zero", _P(t), c_longlong(t.numel() * t.element_size()), _st(), launches=0)
    return t


# ------------------------------------------------------------------------------------------------ batch norm
def bn_finalize(ssum, ssq, count, mean, rstd, running_mean, running_var, nbt, momentum, eps, C):
    _call("ddpx_bn_finalize", _P(ssum), _P(ssq), c_float(count), _P(mean), _P(rstd), _P(running_mean), _P(running_var),
          _P(nbt), c_float(momentum), c_float(eps), c_int(C), _st())


def bn_eval_stats(rm, rv, eps, mean, rstd, C):
    _call("ddpx_bn_eval_stats", _P(rm), _P(rv), c_float(eps), _P(mean), _P(rstd), c_int(C), _st())


def bn_stats(x, P, C, ssum, ssq):
    _call("ddpx_bn_stats", _P(x), c_longlong(P), c_int(C), _P(ssum), _P(ssq), _st())


def bn_apply(x, y, mean, rstd, gamma, beta, N, H, W, C, relu, pool):
    _call("ddpx_bn_apply", _P(x), _P(y), _P(mean), _P(rstd), _P(gamma), _P(beta), c_int(N), c_int(H), c_int(W), c_int(C),
          c_int(relu), c_int(pool), _st())


def bn_bwd_reduce(x, dout, mean, rstd, gamma, beta, N, H, W, C, relu, pool, sum_dz, sum_dzx):
    _call("ddpx_bn_bwd_reduce", _P(x), _P(dout), _P(mean), _P(rstd), _P(gamma), _P(beta), c_int(N), c_int(H), c_int(W),
          c_int(C), c_int(relu), c_int(pool), _P(sum_dz), _P(sum_dzx), _st())


def bn_bwd_apply(x, dout, mean, rstd, gamma, beta, sum_dz, sum_dzx, N, H, W, C, relu, pool, use_batch_stats, dx):
    _call("ddpx_bn_bwd_apply", _P(x), _P(dout), _P(mean), _P(rstd), _P(gamma), _P(beta), _P(sum_dz), _P(sum_dzx), c_int(N),
          c_int(H), c_int(W), c_int(C), c_int(relu), c_int(pool), c_int(use_batch_stats), _P(dx), _st())


# ------------------------------------------------------------------------------------------------ losses
def softmax_ce(logits, targets, dlogits, loss_sum, rows, cols, ld, ignore_index=-100, grad_scale=1.0):
    bf16 = 1 if logits.dtype == torch.bfloat16 else 0
    _call("ddpx_softmax_ce", _P(logits), _P(targets), _P(dlogits), _P(loss_sum), c_longlong(rows), c_int(cols),
          c_longlong(ld), c_int(bf16), c_longlong(ignore_index), c_float(grad_scale), _st())


def scale_by_device_scalar(t, g):
    bf16 = 1 if t.dtype == torch.bfloat16 else 0
    _call("ddpx_scale_by_device_scalar", _P(t), c_longlong(t.numel()), _P(g), c_int(bf16), _st())


def mse(pred, target, dpred, loss_sum):
    _call("ddpx_mse", _P(pred), _P(target), _P(dpred), _P(loss_sum), c_longlong(pred.numel()), _st())


# ------------------------------------------------------------------------------------------------ optimizers
def _ptr_array(tensors):
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr() if t is not None else None
    return arr


def sgd_multi(params, grads, bufs, lr, momentum, dampening, weight_decay, nesterov, first_step, grad_scale=1.0):
    n = len(params)
    numel = (c_longlong * n)(*[p.numel() for p in params])
    _call("ddpx_sgd_multi", c_int(n), _ptr_array(params), _ptr_array(grads), _ptr_array(bufs) if bufs is not None else None,
          numel, c_float(lr), c_float(momentum), c_float(dampening), c_float(weight_decay), c_int(int(nesterov)),
          c_int(int(first_step)), c_float(grad_scale), _st(), launches=(n + 95) // 96)


def adamw_multi(params, grads, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0):
    n = len(params)
    numel = (c_longlong * n)(*[p.numel() for p in params])
    _call("ddpx_adamw_multi", c_int(n), _ptr_array(params), _ptr_array(grads), _ptr_array(exp_avg), _ptr_array(exp_avg_sq),
          numel, c_float(lr), c_float(beta1), c_float(beta2), c_float(eps), c_float(weight_decay), c_longlong(step),
          c_float(grad_scale), _st(), launches=(n + 95) // 96)


def read_error_word() -> int:
    fn = _lib.lib().ddpx_read_error_word
    fn.restype = ctypes.c_uint
    return int(fn())


def gemm_small(A, B, C, M, N, K, sam, sak, sbn, sbk, ldc, bias=None, act=0, accumulate=0):
    _call("ddpx_gemm_small", _P(A), _P(B), _P(C), c_int(M), c_int(N), c_int(K), c_longlong(sam), c_longlong(sak),
          c_longlong(sbn), c_longlong(sbk), c_longlong(ldc), _P(bias), c_int(act), c_int(accumulate), _st())


def relu_bwd_any(y, dy, dx):
    """ReLU backward for fp32 or bf16 activations."""
    if y.dtype == torch.float32:
        relu_bwd(y, dy, dx)
    else:
        _call("ddpx_relu_bwd_bf16", _P(y), _P(dy), _P(dx), c_longlong(y.numel()), _st())
