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


try:                                   # raw current-stream handle without building a torch.cuda.Stream object
    _raw_stream = torch._C._cuda_getCurrentRawStream
except AttributeError:                 # pragma: no cover - other torch builds
    _raw_stream = None


def _st():
    """Current CUDA stream of the current device.  A GPT-2 step makes ~740 native calls and is host-bound, so the
    per-call overhead of this path matters: ~0.3 us instead of ~2 us through ``torch.cuda.current_stream()``."""
    if _raw_stream is not None:
        return c_void_p(_raw_stream(torch.cuda.current_device()))
    return c_void_p(torch.cuda.current_stream().cuda_stream)


_PROFILE = None   # None = off; else list of (name, start_event, end_event)


def profile_begin():
    """Start per-entry-point device timing (CUDA events around every native call; diagnostics only)."""
    global _PROFILE
    _PROFILE = []


def profile_end():
    """Stop timing; returns {name: (calls, total_ms)} sorted by total time."""
    global _PROFILE
    rec, _PROFILE = _PROFILE, None
    torch.cuda.synchronize()
    agg = {}
    for name, e0, e1 in rec or []:
        c, t = agg.get(name, (0, 0.0))
        agg[name] = (c + 1, t + e0.elapsed_time(e1))
    return dict(sorted(agg.items(), key=lambda kv: -kv[1][1]))


_FN = {}


def _call(name, *args, launches: int = 1, tag: str = ""):
    fn = _FN.get(name)
    if fn is None:
        fn = _FN[name] = getattr(_lib.lib(), name)
    if _PROFILE is not None:
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = fn(*args)
        e1.record()
        _PROFILE.append((name + tag, e0, e1))
    else:
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
          c_int(split_k), c_int(accumulate), _st(), tag=f"[{M}x{N}x{K} {'NT'[a_mn]}{'NT'[b_mn]}]" if _PROFILE is not None else "")


def conv3x3_fprop(x_nhwc, w_packed, y_nhwc, N, H, W, Cin, Cout, bias=None, act=0, stat_sum=None, stat_sq=None):
    _call("ddpx_conv3x3_fprop", _P(x_nhwc), _P(w_packed), _P(y_nhwc), c_int(N), c_int(H), c_int(W), c_int(Cin),
          c_int(Cout), _P(bias), c_int(act), _P(stat_sum), _P(stat_sq), _st(), tag=(f"[{Cin}->{Cout}@{H}]" if _PROFILE is not None else ""))


def conv3x3_wgrad(dy_nhwc, x_nhwc, dw, N, H, W, Cin, Cout, split_k=0):
    _call("ddpx_conv3x3_wgrad", _P(dy_nhwc), _P(x_nhwc), _P(dw), c_int(N), c_int(H), c_int(W), c_int(Cin), c_int(Cout),
          c_int(split_k), _st(), tag=(f"[{Cin}->{Cout}@{H}]" if _PROFILE is not None else ""))


def conv3x3_wgrad_packed(dy_nhwc, x_nhwc, dw_packed, N, H, W, Cin, Cout, split_k=0):
    _call("ddpx_conv3x3_wgrad_packed", _P(dy_nhwc), _P(x_nhwc), _P(dw_packed), c_int(N), c_int(H), c_int(W), c_int(Cin),
          c_int(Cout), c_int(split_k), _st(), tag=(f"[{Cin}->{Cout}@{H}]" if _PROFILE is not None else ""))


def unpack_conv_wgrad(dw_packed, dw, Cout, Cin):
    _call("ddpx_unpack_conv_wgrad", _P(dw_packed), _P(dw), c_int(Cout), c_int(Cin), _st())


def im2col_stem(x_nchw, patches, N, H, W):
    _call("ddpx_im2col_stem", _P(x_nchw), _P(patches), c_int(N), c_int(H), c_int(W), _st())


def pack_stem_weights(w, wp, Cout):
    _call("ddpx_pack_stem_weights", _P(w), _P(wp), c_int(Cout), _st())


def conv1x1_fprop(x_nhwc, w_packed, y_nhwc, N, H, W, Cin, Cout, bias=None, act=0, stat_sum=None, stat_sq=None):
    _call("ddpx_conv1x1_fprop", _P(x_nhwc), _P(w_packed), _P(y_nhwc), c_int(N), c_int(H), c_int(W), c_int(Cin),
          c_int(Cout), _P(bias), c_int(act), _P(stat_sum), _P(stat_sq), _st(), tag=(f"[{Cin}->{Cout}@{H}]" if _PROFILE is not None else ""))


def gemm_batched(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, batch1, batch2, a_bs, b_bs, d_bs, alpha=1.0, causal=0,
                 bias=None, act=0):
    """Batched D = act(alpha * op(A) op(B) + bias); *_bs = (stride_b1, stride_b2) in elements."""
    bf16 = 1 if A.dtype == torch.bfloat16 else 0
    out_bf16 = 1 if D.dtype == torch.bfloat16 else 0
    _call("ddpx_gemm_batched", _P(A), _P(B), _P(D), c_int(M), c_int(N), c_int(K), c_longlong(lda), c_longlong(ldb),
          c_longlong(ldd), c_int(a_mn), c_int(b_mn), c_int(bf16), c_int(out_bf16), _P(bias), c_int(act), c_float(alpha),
          c_int(1), c_int(0), c_int(batch1), c_int(batch2), c_longlong(a_bs[0]), c_longlong(a_bs[1]), c_longlong(b_bs[0]),
          c_longlong(b_bs[1]), c_longlong(d_bs[0]), c_longlong(d_bs[1]), c_int(causal), _st(),
          tag=f"[b{batch1 * batch2} {M}x{N}x{K} {'NT'[a_mn]}{'NT'[b_mn]} c{causal}]" if _PROFILE is not None else "")


def conv_bn_fwd(x, wf, gamma, beta, rm, rv, nbt, momentum, eps, N, H, W, Cin, Cout, relu, pool, training, stats, y, out, mr,
                patches=None):
    """conv3x3 -> BatchNorm (finalize fused into the apply pass) -> [ReLU] -> [pool]: 2 launches (3 for the stem), 1 call."""
    _call("ddpx_conv_bn_fwd", _P(x), _P(wf), _P(gamma), _P(beta), _P(rm), _P(rv), _P(nbt), c_float(momentum), c_float(eps),
          c_int(N), c_int(H), c_int(W), c_int(Cin), c_int(Cout), c_int(relu), c_int(pool), c_int(training), _P(stats), _P(y),
          _P(out), _P(mr), _P(patches), _st(), launches=3 if Cin == 3 else 2,
          tag=(f"[{Cin}->{Cout}@{H}]" if _PROFILE is not None else ""))


def conv_bn_bwd(y, dout, xin, wd, mr, gamma, beta, N, H, W, Cin, Cout, relu, pool, training, sums, dconv, staging, dw, dx):
    _call("ddpx_conv_bn_bwd", _P(y), _P(dout), _P(xin), _P(wd), _P(mr), _P(gamma), _P(beta), c_int(N), c_int(H), c_int(W),
          c_int(Cin), c_int(Cout), c_int(relu), c_int(pool), c_int(training), _P(sums), _P(dconv), _P(staging), _P(dw), _P(dx),
          _st(), launches=3 if Cin == 3 else (5 if dx is not None else 4),
          tag=(f"[{Cin}->{Cout}@{H}]" if _PROFILE is not None else ""))


def pack_conv_weights_multi(ws, wfs, wds, zeros, zero_ns, couts, cins):
    n = len(ws)
    _call("ddpx_pack_conv_weights_multi", c_int(n), _ptr_array(ws), _ptr_array(wfs), _ptr_array(wds), _ptr_array(zeros),
          (c_int * n)(*zero_ns), (c_int * n)(*couts), (c_int * n)(*cins), _st())


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
          c_int(relu), c_int(pool), _st(), tag=(f"[{C}@{H}]" if _PROFILE is not None else ""))


def bn_train_fwd(x, y, gamma, beta, running_mean, running_var, nbt, momentum, eps, N, H, W, C, relu, pool, mean_rstd, work):
    _call("ddpx_bn_train_fwd", _P(x), _P(y), _P(gamma), _P(beta), _P(running_mean), _P(running_var), _P(nbt),
          c_float(momentum), c_float(eps), c_int(N), c_int(H), c_int(W), c_int(C), c_int(relu), c_int(pool), _P(mean_rstd),
          _P(work), _st(), launches=3)


def bn_bwd_reduce(x, dout, mean, rstd, gamma, beta, N, H, W, C, relu, pool, sum_dz, sum_dzx):
    _call("ddpx_bn_bwd_reduce", _P(x), _P(dout), _P(mean), _P(rstd), _P(gamma), _P(beta), c_int(N), c_int(H), c_int(W),
          c_int(C), c_int(relu), c_int(pool), _P(sum_dz), _P(sum_dzx), _st(), tag=(f"[{C}@{H}]" if _PROFILE is not None else ""))


def bn_bwd_apply(x, dout, mean, rstd, gamma, beta, sum_dz, sum_dzx, N, H, W, C, relu, pool, use_batch_stats, dx):
    _call("ddpx_bn_bwd_apply", _P(x), _P(dout), _P(mean), _P(rstd), _P(gamma), _P(beta), _P(sum_dz), _P(sum_dzx), c_int(N),
          c_int(H), c_int(W), c_int(C), c_int(relu), c_int(pool), c_int(use_batch_stats), _P(dx), _st(), tag=(f"[{C}@{H}]" if _PROFILE is not None else ""))


# ------------------------------------------------------------------------------------------------ losses
def softmax_ce(logits, targets, dlogits, loss_sum, rows, cols, ld, ignore_index=-100, grad_scale=1.0):
    bf16 = 1 if logits.dtype == torch.bfloat16 else 0
    _call("ddpx_softmax_ce", _P(logits), _P(targets), _P(dlogits), _P(loss_sum), c_longlong(rows), c_int(cols),
          c_longlong(ld), c_int(bf16), c_longlong(ignore_index), c_float(grad_scale), _st())


def ce_fwd(logits, targets, lse, loss_sum, rows, cols, ld, ignore_index=-100):
    _call("ddpx_ce_fwd", _P(logits), _P(targets), _P(lse), _P(loss_sum), c_longlong(rows), c_int(cols), c_longlong(ld),
          c_int(1 if logits.dtype == torch.bfloat16 else 0), c_longlong(ignore_index), _st())


def ce_bwd(logits, targets, lse, g, dlogits, rows, cols, ld, ignore_index=-100):
    _call("ddpx_ce_bwd", _P(logits), _P(targets), _P(lse), _P(g), _P(dlogits), c_longlong(rows), c_int(cols), c_longlong(ld),
          c_int(1 if logits.dtype == torch.bfloat16 else 0), c_longlong(ignore_index), _st())


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
          c_int(int(first_step)), c_float(grad_scale), c_int(1 if params[0].dtype == torch.bfloat16 else 0), _st(),
          launches=(n + 95) // 96)


def scale_copy_multi(dsts, srcs, scale):
    """dsts[i] = srcs[i] * scale, one launch per 96 tensors (all the same dtype, contiguous)."""
    n = len(dsts)
    numel = (c_longlong * n)(*[d.numel() for d in dsts])
    _call("ddpx_scale_copy_multi", c_int(n), _ptr_array(dsts), _ptr_array(srcs), numel, c_float(scale),
          c_int(1 if dsts[0].dtype == torch.bfloat16 else 0), _st(), launches=(n + 95) // 96)


def adamw_multi(params, grads, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0):
    n = len(params)
    numel = (c_longlong * n)(*[p.numel() for p in params])
    _call("ddpx_adamw_multi", c_int(n), _ptr_array(params), _ptr_array(grads), _ptr_array(exp_avg), _ptr_array(exp_avg_sq),
          numel, c_float(lr), c_float(beta1), c_float(beta2), c_float(eps), c_float(weight_decay), c_longlong(step),
          c_float(grad_scale), c_int(1 if params[0].dtype == torch.bfloat16 else 0), _st(), launches=(n + 95) // 96)


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


# ------------------------------------------------------------------------------------------------ transformer kernels
def _bf(t):
    return 1 if t.dtype == torch.bfloat16 else 0


def layernorm_fwd(x, gamma, beta, y, mean, rstd, rows, C, eps):
    _call("ddpx_layernorm_fwd", _P(x), _P(gamma), _P(beta), _P(y), _P(mean), _P(rstd), c_longlong(rows), c_int(C),
          c_float(eps), c_int(_bf(x)), _st())


def layernorm_bwd_scratch_floats(rows, C) -> int:
    """fp32 elements of per-block partial sums ddpx_layernorm_bwd_ws wants: [min(296, ceil(rows / 8))][2C] (the larger
    grid of csrc/kernels/transformer.cu, restated so that sizing a buffer needs no library call)."""
    return min(296, max(1, (rows + 7) // 8)) * 2 * C


def layernorm_bwd(x, dy, gamma, mean, rstd, dx, dgamma, dbeta, rows, C, scratch=None, dres=None):
    """Without scratch dgamma / dbeta are zeroed accumulators; with it they are overwritten.  dres (needs scratch): the
    gradient of the residual branch around the LayerNorm, added to dx."""
    if dres is not None:
        assert scratch is not None and scratch.numel() >= layernorm_bwd_scratch_floats(rows, C)
        _call("ddpx_layernorm_bwd_res", _P(x), _P(dy), _P(gamma), _P(mean), _P(rstd), _P(dx), _P(dgamma), _P(dbeta),
              c_longlong(rows), c_int(C), c_int(_bf(x)), _P(scratch), _P(dres), _st(), launches=3 if rows < 148 * 8 * 4 else 2)
    elif scratch is None:
        _call("ddpx_layernorm_bwd", _P(x), _P(dy), _P(gamma), _P(mean), _P(rstd), _P(dx), _P(dgamma), _P(dbeta),
              c_longlong(rows), c_int(C), c_int(_bf(x)), _st())
    else:
        assert scratch.numel() >= layernorm_bwd_scratch_floats(rows, C) and scratch.dtype == torch.float32
        _call("ddpx_layernorm_bwd_ws", _P(x), _P(dy), _P(gamma), _P(mean), _P(rstd), _P(dx), _P(dgamma), _P(dbeta),
              c_longlong(rows), c_int(C), c_int(_bf(x)), _P(scratch), _st(), launches=3 if rows < 148 * 8 * 4 else 2)


def gelu_fwd(x, y):
    _call("ddpx_gelu_fwd", _P(x), _P(y), c_longlong(x.numel()), c_int(_bf(x)), _st())


def gelu_bwd(x, dy, dx):
    _call("ddpx_gelu_bwd", _P(x), _P(dy), _P(dx), c_longlong(x.numel()), c_int(_bf(x)), _st())


def add(a, b, out):
    _call("ddpx_add", _P(a), _P(b), _P(out), c_longlong(a.numel()), c_int(_bf(a)), _st())


def cast_scale(src, dst, scale=1.0):
    """dst = src * scale across fp32/bf16 (f32->bf16, bf16->f32, bf16->bf16)."""
    if src.dtype == torch.float32 and dst.dtype == torch.bfloat16:
        kind = 0
    elif src.dtype == torch.bfloat16 and dst.dtype == torch.float32:
        kind = 1
    elif src.dtype == torch.bfloat16 and dst.dtype == torch.bfloat16:
        kind = 2
    else:
        raise TypeError("cast_scale: unsupported dtype pair")
    _call("ddpx_cast_scale", _P(src), _P(dst), c_longlong(src.numel()), c_float(scale), c_int(kind), _st())


def scale_copy_bf16(src, dst, n, scale):
    _call("ddpx_cast_scale", _P(src), _P(dst), c_longlong(n), c_float(scale), c_int(2), _st())


def cast(src, dtype):
    dst = torch.empty(src.shape, device=src.device, dtype=dtype)
    cast_scale(src, dst, 1.0)
    return dst


def colsum_any(x, out, rows, cols):
    if x.dtype == torch.float32:
        colsum(x, out, rows, cols)
    else:
        _call("ddpx_colsum_bf16", _P(x), _P(out), c_longlong(rows), c_int(cols), _st())


def dropout(x, y, p, seed):
    _call("ddpx_dropout", _P(x), _P(y), c_longlong(x.numel()), c_float(p), ctypes.c_ulonglong(seed), c_int(_bf(x)), _st())


def embedding_fwd(idx, wte, wpe, out, BT, T, C):
    _call("ddpx_embedding_fwd", _P(idx), _P(wte), _P(wpe), _P(out), c_longlong(BT), c_int(T), c_int(C), c_int(_bf(wte)), _st())


def embedding_bwd(idx, dy, dwte, dwpe, BT, T, C):
    _call("ddpx_embedding_bwd", _P(idx), _P(dy), _P(dwte), _P(dwpe), c_longlong(BT), c_int(T), c_int(C), c_int(_bf(dy)), _st())


def causal_softmax_fwd(s, rows, T):
    _call("ddpx_causal_softmax_fwd", _P(s), c_longlong(rows), c_int(T), c_int(_bf(s)), _st())


def causal_softmax_bwd(p, dp, rows, T, scale):
    _call("ddpx_causal_softmax_bwd", _P(p), _P(dp), c_longlong(rows), c_int(T), c_float(scale), c_int(_bf(p)), _st())


# ------------------------------------------------------------------------------------------------ fused causal attention
def attn_fwd(qkv, out, lse2, B, T, H, scale):
    """Flash-style causal attention forward: qkv [B,T,3*H*64] bf16 -> out [B,T,H*64] bf16, lse2 [B,H,T] fp32."""
    _call("ddpx_attn_fwd", _P(qkv), _P(out), _P(lse2), c_int(B), c_int(T), c_int(H), c_float(scale), _st())


def attn_bwd(qkv, out, dout, lse2, delta, dqkv, B, T, H, scale):
    """Backward of attn_fwd: writes dqkv [B,T,3C] (dq | dk | dv); delta [B,H,T] fp32 is scratch."""
    _call("ddpx_attn_bwd", _P(qkv), _P(out), _P(dout), _P(lse2), _P(delta), _P(dqkv), c_int(B), c_int(T), c_int(H),
          c_float(scale), _st(), launches=3)


def attn_error_word() -> int:
    v = 0
    for name in ("ddpx_attn_read_error_word", "ddpx_attn_bwd_read_error_word"):
        fn = getattr(_lib.lib(), name)
        fn.restype = ctypes.c_uint
        v |= int(fn())
    return v


# ------------------------------------------------------------------------------------------------ transformer block
def gpt2_block_fwd(x, params, acts, stats, tmp, out, B, T, C, H, eps):
    """One GPT-2 block forward in one call: 7 kernels with the default DDPX_BLOCK_FUSE (two LayerNorms, four GEMMs with the
    residual adds / GELU in their epilogues, fused attention).  params / acts: lists of device pointers (ints)."""
    _call("ddpx_gpt2_block_fwd", _P(x), (c_void_p * 12)(*params), (c_void_p * 7)(*acts), _P(stats), _P(tmp), _P(out),
          c_int(B), c_int(T), c_int(C), c_int(H), c_float(eps), _st(), launches=7)


def gpt2_block_bwd(x, dout, params, acts, stats, grads, dx, wsb, ws32, cnt, B, T, C, H):
    """Backward of the block in one call: eight GEMMs, two attention kernels, two LayerNorm backwards (two kernels each, three
    in the few-rows form), one column-sum and one cast launch."""
    ln = 3 if B * T < 148 * 8 * 4 else 2
    _call("ddpx_gpt2_block_bwd", _P(x), _P(dout), (c_void_p * 12)(*params), (c_void_p * 7)(*acts), _P(stats),
          (c_void_p * 12)(*grads), _P(dx), _P(wsb), _P(ws32), _P(cnt), c_int(B), c_int(T), c_int(C), c_int(H), _st(),
          launches=12 + 2 * ln)


def gemm_fused(A, B, D, M, N, K, lda, ldb, ldd, a_mn=0, b_mn=0, bias=None, act=0, aux=None, aux_op=0, out2=None):
    """bf16 GEMM with an elementwise pass folded into the store: aux_op 1 adds `aux` (residual), aux_op 2 multiplies by
    gelu'(aux); out2 (act = 2) also receives the pre-activation.  aux / out2: bf16 [M, ldd]."""
    _call("ddpx_gemm_fused", _P(A), _P(B), _P(D), c_int(M), c_int(N), c_int(K), c_longlong(lda), c_longlong(ldb),
          c_longlong(ldd), c_int(a_mn), c_int(b_mn), _P(bias), c_int(1 if bias is not None and bias.dtype == torch.bfloat16 else 0),
          c_int(act), _P(aux), c_int(aux_op), _P(out2), _st())


def gemm_splitk_bf16(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, split_k, ws, cnt):
    """bf16 split-K GEMM with the fix-up inside the kernel; ws (fp32 [M*N]) and cnt (int32) all zero before and after."""
    _call("ddpx_gemm_splitk_bf16", _P(A), _P(B), _P(D), c_int(M), c_int(N), c_int(K), c_longlong(lda), c_longlong(ldb),
          c_longlong(ldd), c_int(a_mn), c_int(b_mn), c_int(split_k), _P(ws), _P(cnt), _st())
