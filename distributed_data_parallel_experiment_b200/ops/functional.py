"""Differentiable ops.  On a CUDA tensor every op runs the hand-written sm_100a kernels (ops/native.py); on a CPU
tensor it runs plain torch math with identical semantics (unit tests, gloo runs).  There is no torch-kernel fallback on
CUDA: an unsupported shape raises.

Image tensors are logical NCHW like torch's; on CUDA they are kept in channels-last (NHWC) memory, which is what the
implicit-GEMM convolution and the fused BatchNorm kernels read.
"""
from __future__ import annotations


import torch
import torch.nn.functional as F
from torch.autograd import Function

from . import native

# ------------------------------------------------------------------------------------------------ layout helpers


def to_nhwc(x: torch.Tensor) -> torch.Tensor:
    """Logical-NCHW CUDA tensor -> contiguous [N,H,W,C] tensor (no copy when already channels-last)."""
    v = x.permute(0, 2, 3, 1)
    if v.is_contiguous():
        return v
    N, C, H, W = x.shape
    xc = x if x.is_contiguous() else x.contiguous()
    out = torch.empty((N, H, W, C), device=x.device, dtype=x.dtype)
    native.transpose_batched(xc, out, N, C, H * W)
    return out


def from_nhwc(t: torch.Tensor) -> torch.Tensor:
    """[N,H,W,C] contiguous -> logical NCHW view (channels-last strides)."""
    return t.permute(0, 3, 1, 2)


def to_nchw_contiguous(x: torch.Tensor) -> torch.Tensor:
    """Logical-NCHW tensor -> NCHW-contiguous memory (own transpose kernel on CUDA)."""
    if x.is_contiguous():
        return x
    if not x.is_cuda:
        return x.contiguous()
    N, C, H, W = x.shape
    src = to_nhwc(x)
    out = torch.empty((N, C, H, W), device=x.device, dtype=x.dtype)
    native.transpose_batched(src, out, N, H * W, C)
    return out


def _zeros(shape, device, dtype=torch.float32):
    return native.zero_(torch.empty(shape, device=device, dtype=dtype))


def _check_f32(*ts):
    for t in ts:
        if t is not None and t.dtype != torch.float32:
            raise TypeError(f"native conv/BN path is fp32 (TF32 tensor cores); got {t.dtype}")


# ------------------------------------------------------------------------------------------------ conv (+BN+ReLU+pool)
import os as _os
_WGRAD_PACKED = _os.environ.get("DDPX_WGRAD_PACKED", "1") not in ("0", "")   # A/B switch for the staging-layout epilogue
_BLOCK_CALLS = _os.environ.get("DDPX_BLOCK_CALLS", "1") not in ("0", "")      # A/B switch: 0 = one native call per kernel
_STEM_SPLIT_K = 296   # two K-slices per SM for the pixels-long reduction of the stem weight gradient


def _conv_forward(x, w, y, N, H, W_, Cin, Cout, stem, bias, act, ssum, ssq, need_dgrad):
    """Runs the forward convolution; returns (tensor to save for the weight gradient, packed dgrad weights or None).

    RGB stem: the 3x3x3 patches are gathered once into 32-float rows (kept for the backward) and the convolution is a
    1x1 tcgen05 conv over them; everything else is the implicit-GEMM 3x3 kernel straight off the NHWC activations."""
    if stem:
        xc = x if x.is_contiguous() else to_nchw_contiguous(x)
        patches = torch.empty((N, H, W_, 32), device=w.device, dtype=torch.float32)
        native.im2col_stem(xc, patches, N, H, W_)
        wp = torch.empty(Cout * 32, device=w.device, dtype=torch.float32)
        native.pack_stem_weights(w, wp, Cout)
        native.conv1x1_fprop(patches, wp, y, N, H, W_, 32, Cout, bias, act, ssum, ssq)
        return patches, None
    xin = to_nhwc(x)
    wd = None
    wp = torch.empty(9 * Cout * Cin, device=w.device, dtype=torch.float32)
    if need_dgrad:
        wd = torch.empty(9 * Cout * Cin, device=w.device, dtype=torch.float32)
    native.pack_conv_weights(w, wp, wd, Cout, Cin)
    native.conv3x3_fprop(xin, wp, y, N, H, W_, Cin, Cout, bias, act, ssum, ssq)
    return xin, wd


def _conv_backward(ctx_stem, dconv, xin, w, wd, N, H, W_, Cin, Cout, need_dx):
    if ctx_stem:
        dw = _zeros(w.shape, w.device)
        # dw[Cout][27] += dconv^T[Cout][P] @ patches[P][0:27]  (both operands MN-major, split-K, fp32 atomics)
        native.gemm(dconv, xin, dw, Cout, 27, N * H * W_, Cout, 32, 27, 1, 1, None, 0, _STEM_SPLIT_K, 1)
        return None, dw
    if _WGRAD_PACKED:
        # accumulate into the tap-major staging layout with 16-byte vector reductions, then transpose once
        staging = _zeros((9 * Cout * Cin,), w.device)
        native.conv3x3_wgrad_packed(dconv, xin, staging, N, H, W_, Cin, Cout)
        dw = torch.empty(w.shape, device=w.device, dtype=torch.float32)
        native.unpack_conv_wgrad(staging, dw, Cout, Cin)
    else:
        dw = _zeros(w.shape, w.device)
        native.conv3x3_wgrad(dconv, xin, dw, N, H, W_, Cin, Cout)
    dx = None
    if need_dx:
        dx = torch.empty((N, H, W_, Cin), device=w.device, dtype=torch.float32)
        native.conv3x3_fprop(dconv, wd, dx, N, H, W_, Cout, Cin)   # data gradient = conv with flipped/transposed weights
    return dx, dw


def _conv_supported(Cin, Cout, H, W_):
    if Cin == 3:
        return Cout % 32 == 0
    return Cin % 32 == 0 and Cout % 32 == 0


class ConvWorkspace:
    """Persistent per-layer device buffers of the fused conv-BN block: packed forward / data-gradient weights, the
    BatchNorm (sum, sum^2) accumulators and the tap-major weight-gradient staging buffer (kept all-zero between uses: the
    unpack kernel clears what it reads).  Lives as an attribute of the weight Parameter."""
    __slots__ = ("wf", "wd", "stats", "staging", "Cout", "Cin", "fresh", "ptr")

    def __init__(self, w):
        Cout, Cin = int(w.shape[0]), int(w.shape[1])
        dev = w.device
        self.Cout, self.Cin, self.ptr = Cout, Cin, w.data_ptr()
        stem = Cin == 3
        self.wf = torch.empty(Cout * 32 if stem else 9 * Cout * Cin, device=dev, dtype=torch.float32)
        self.wd = None if stem else torch.empty(9 * Cout * Cin, device=dev, dtype=torch.float32)
        self.stats = _zeros((2, Cout), dev)
        self.staging = None if stem else _zeros((9 * Cout * Cin,), dev)
        self.fresh = False


def conv_workspace(w) -> ConvWorkspace:
    ws = getattr(w, "_ddpx_conv_ws", None)
    if ws is None or ws.ptr != w.data_ptr():
        ws = ConvWorkspace(w.detach())
        try:
            w._ddpx_conv_ws = ws
        except AttributeError:      # exotic tensor subclasses without a __dict__: per-call workspace
            pass
    return ws


def prepare_convs(weights):
    """Pack the weights of every listed conv layer (forward and data-gradient layouts) and clear their BatchNorm
    accumulators in ONE launch.  Call once per forward, before the blocks (models.VGG does); a block whose workspace was
    not prepared since its last use prepares itself."""
    wss = [conv_workspace(w) for w in weights]
    native.pack_conv_weights_multi([w.detach() for w in weights], [s.wf for s in wss], [s.wd for s in wss],
                                   [s.stats for s in wss], [2 * s.Cout for s in wss], [s.Cout for s in wss],
                                   [s.Cin for s in wss])
    for s in wss:
        s.fresh = True
    return wss


class _ConvBNBlock(Function):
    """conv3x3(pad 1, no bias) -> BatchNorm2d -> [ReLU] -> [MaxPool2d(2)]: ONE native call per direction
    (csrc/kernels/blocks.cu).  Forward: convolution (+ batch statistics in its epilogue) and a normalise pass that derives
    mean / rstd itself; backward: BN reduce + apply, weight gradient into the persistent staging buffer, unpack, data
    gradient."""

    @staticmethod
    def forward(ctx, x, beta, gamma, w, running_mean, running_var, nbt, training, momentum, eps, relu, pool, ws):
        # input order (x, beta, gamma, w): gradients reach the parameters' accumulators - and the reducer's hooks - in the
        # order torch's unfused modules deliver them (bn.bias, bn.weight, conv.weight), so the rebuilt buckets match
        # torch DDP's layout
        _check_f32(x, w, gamma, beta)
        N, Cin, H, W_ = x.shape
        Cout = w.shape[0]
        stem = Cin == 3
        dev = x.device
        need_dx = ctx.needs_input_grad[0]
        if not ws.fresh:
            prepare_convs([w])
        ws.fresh = False
        y = torch.empty((N, H, W_, Cout), device=dev, dtype=torch.float32)
        Ho, Wo = (H // 2, W_ // 2) if pool else (H, W_)
        out = torch.empty((N, Ho, Wo, Cout), device=dev, dtype=torch.float32)
        mr = torch.empty((2, Cout), device=dev, dtype=torch.float32)
        if stem:
            xin = x if x.is_contiguous() else to_nchw_contiguous(x)
            patches = torch.empty((N, H, W_, 32), device=dev, dtype=torch.float32)
            native.conv_bn_fwd(xin, ws.wf, gamma, beta, running_mean, running_var, nbt, momentum, eps, N, H, W_, Cin, Cout,
                               int(relu), int(pool), int(training), ws.stats, y, out, mr, patches)
            saved_in = patches
        else:
            saved_in = to_nhwc(x)
            native.conv_bn_fwd(saved_in, ws.wf, gamma, beta, running_mean, running_var, nbt, momentum, eps, N, H, W_, Cin,
                               Cout, int(relu), int(pool), int(training), ws.stats, y, out, mr)
        ctx.save_for_backward(saved_in, y, mr, gamma, beta)
        ctx.ws = ws
        ctx.cfg = (N, H, W_, Cin, Cout, bool(training), bool(relu), bool(pool), need_dx, tuple(w.shape))
        return from_nhwc(out)

    @staticmethod
    def backward(ctx, dout):
        xin, y, mr, gamma, beta = ctx.saved_tensors
        N, H, W_, Cin, Cout, training, relu, pool, need_dx, wshape = ctx.cfg
        ws = ctx.ws
        dev = y.device
        d = to_nhwc(dout)
        sums = torch.empty((2, Cout), device=dev, dtype=torch.float32)
        dconv = torch.empty_like(y)
        dw = torch.empty(wshape, device=dev, dtype=torch.float32)
        dx = torch.empty((N, H, W_, Cin), device=dev, dtype=torch.float32) if (need_dx and Cin != 3) else None
        native.conv_bn_bwd(y, d, xin, ws.wd, mr, gamma, beta, N, H, W_, Cin, Cout, int(relu), int(pool), int(training), sums,
                           dconv, ws.staging, dw, dx)
        return ((from_nhwc(dx) if dx is not None else None), sums[0], sums[1], dw,
                None, None, None, None, None, None, None, None, None)


class _ConvBNAct(Function):
    """conv3x3(pad 1, no bias) -> BatchNorm2d -> [ReLU] -> [MaxPool2d(2)] as one differentiable op (VGG block)."""

    @staticmethod
    def forward(ctx, x, w, gamma, beta, running_mean, running_var, nbt, training, momentum, eps, relu, pool, sync=None):
        _check_f32(x, w, gamma, beta)
        N, Cin, H, W_ = x.shape
        Cout = w.shape[0]
        if not _conv_supported(Cin, Cout, H, W_):
            raise NotImplementedError(f"native conv3x3: unsupported channels Cin={Cin} Cout={Cout}")
        stem = Cin == 3
        dev = x.device
        y = torch.empty((N, H, W_, Cout), device=dev, dtype=torch.float32)
        stats = _zeros((2, Cout), dev) if training else None
        need_dx = ctx.needs_input_grad[0]
        xin, wd = _conv_forward(x, w.detach(), y, N, H, W_, Cin, Cout, stem, None, 0,
                                stats[0] if training else None, stats[1] if training else None, need_dx)
        mr = torch.empty((2, Cout), device=dev, dtype=torch.float32)
        sync_key, sync_world = None, 1
        if training and sync is not None and sync[0].world_size > 1:
            # SyncBatchNorm: batch statistics over the GLOBAL batch - sum the (sum x, sum x^2) blocks over the ranks
            sync_key, sync_world = sync[1], sync[0].world_size
            stats = sync[0].reduce(stats, sync_key + "/fwd")
        if training:
            native.bn_finalize(stats[0], stats[1], float(N * H * W_ * sync_world), mr[0], mr[1], running_mean, running_var,
                               nbt, momentum, eps, Cout)
        else:
            native.bn_eval_stats(running_mean, running_var, eps, mr[0], mr[1], Cout)
        Ho, Wo = (H // 2, W_ // 2) if pool else (H, W_)
        out = torch.empty((N, Ho, Wo, Cout), device=dev, dtype=torch.float32)
        native.bn_apply(y, out, mr[0], mr[1], gamma, beta, N, H, W_, Cout, int(relu), int(pool))
        ctx.save_for_backward(xin, y, mr, gamma, beta, w, wd)
        ctx.cfg = (N, H, W_, Cin, Cout, stem, bool(training), bool(relu), bool(pool), need_dx)
        ctx.sync = (sync[0], sync_key) if sync_key is not None else None
        return from_nhwc(out)

    @staticmethod
    def backward(ctx, dout):
        xin, y, mr, gamma, beta, w, wd = ctx.saved_tensors
        N, H, W_, Cin, Cout, stem, training, relu, pool, need_dx = ctx.cfg
        dev = y.device
        d = to_nhwc(dout)
        sums = _zeros((2, Cout), dev)
        native.bn_bwd_reduce(y, d, mr[0], mr[1], gamma, beta, N, H, W_, Cout, int(relu), int(pool), sums[0], sums[1])
        dconv = torch.empty_like(y)
        tot = sums
        if ctx.sync is not None:
            # the input gradient needs the global (sum dy, sum dy*xhat); bn_bwd_apply divides by the LOCAL pixel count,
            # so hand it the global sums divided by the world size.  gamma / beta keep their local sums (DDP averages)
            tot = ctx.sync[0].reduce(sums, ctx.sync[1] + "/bwd")
            native.scale_copy(tot, tot, tot.numel(), 1.0 / ctx.sync[0].world_size)
        native.bn_bwd_apply(y, d, mr[0], mr[1], gamma, beta, tot[0], tot[1], N, H, W_, Cout, int(relu), int(pool),
                            int(training), dconv)
        dx, dw = _conv_backward(stem, dconv, xin, w, wd, N, H, W_, Cin, Cout, need_dx)
        dxo = from_nhwc(dx) if dx is not None else None
        return dxo, dw, sums[1], sums[0], None, None, None, None, None, None, None, None, None


class _ConvAct(Function):
    """conv3x3(pad 1) + bias + [ReLU] (DeepNN blocks, singlegpu.py:21-31)."""

    @staticmethod
    def forward(ctx, x, w, bias, relu):
        _check_f32(x, w, bias)
        N, Cin, H, W_ = x.shape
        Cout = w.shape[0]
        if not _conv_supported(Cin, Cout, H, W_):
            raise NotImplementedError(f"native conv3x3: unsupported channels Cin={Cin} Cout={Cout}")
        stem = Cin == 3
        y = torch.empty((N, H, W_, Cout), device=x.device, dtype=torch.float32)
        need_dx = ctx.needs_input_grad[0]
        xin, wd = _conv_forward(x, w.detach(), y, N, H, W_, Cin, Cout, stem, bias.detach() if bias is not None else None,
                                1 if relu else 0, None, None, need_dx)
        ctx.save_for_backward(xin, y, w, wd)
        ctx.cfg = (N, H, W_, Cin, Cout, stem, bool(relu), need_dx, bias is not None)
        return from_nhwc(y)

    @staticmethod
    def backward(ctx, dout):
        xin, y, w, wd = ctx.saved_tensors
        N, H, W_, Cin, Cout, stem, relu, need_dx, has_bias = ctx.cfg
        d = to_nhwc(dout)
        if relu:
            dz = torch.empty_like(y)
            native.relu_bwd(y, d, dz)
        else:
            dz = d
        dx, dw = _conv_backward(stem, dz, xin, w, wd, N, H, W_, Cin, Cout, need_dx)
        db = None
        if has_bias:
            db = torch.empty(Cout, device=y.device, dtype=torch.float32)
            native.colsum(dz, db, N * H * W_, Cout)
        return (from_nhwc(dx) if dx is not None else None), dw, db, None


def conv_bn_act(x, w, gamma, beta, running_mean, running_var, num_batches_tracked, training: bool, momentum: float = 0.1,
                eps: float = 1e-5, relu: bool = True, pool: bool = False, sync=None):
    """Fused VGG block (multigpu.py:53-59).  CPU tensors use the torch composition with the same semantics.
    ``sync = (StatSync, key)`` turns the BatchNorm into a SyncBatchNorm (statistics over all ranks)."""
    if x.is_cuda:
        Cin, Cout = x.shape[1], w.shape[0]
        use_sync = training and sync is not None and sync[0].world_size > 1
        if _BLOCK_CALLS and not use_sync and x.dtype == torch.float32 and Cout % 32 == 0 and (Cin == 3 or Cin % 32 == 0) \
                and running_mean is not None:
            return _ConvBNBlock.apply(x, beta, gamma, w, running_mean, running_var, num_batches_tracked, training, momentum,
                                      eps, relu, pool, conv_workspace(w))
        return _ConvBNAct.apply(x, w, gamma, beta, running_mean, running_var, num_batches_tracked, training, momentum, eps,
                                relu, pool, sync)
    y = F.conv2d(x, w, None, padding=1)
    y = batch_norm(y, gamma, beta, running_mean, running_var, num_batches_tracked, training, momentum, eps, sync=sync)
    if relu:
        y = F.relu(y)
    if pool:
        y = F.max_pool2d(y, 2)
    return y


def conv3x3(x, w, bias=None, relu: bool = False):
    if x.is_cuda:
        return _ConvAct.apply(x, w, bias, relu)
    y = F.conv2d(x, w, bias, padding=1)
    return F.relu(y) if relu else y


# ------------------------------------------------------------------------------------------------ stand-alone BN / ReLU / pool
class _BatchNorm(Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, nbt, training, momentum, eps, sync=None):
        _check_f32(x, gamma, beta)
        N, C, H, W_ = x.shape
        xin = to_nhwc(x)
        dev = x.device
        mr = torch.empty((2, C), device=dev, dtype=torch.float32)
        out = torch.empty_like(xin)
        ctx.sync = None
        if training and sync is not None and sync[0].world_size > 1:
            # SyncBatchNorm: local (sum, sum^2) -> summed over the ranks -> finalize with the global count -> apply
            stats = _zeros((2, C), dev)
            native.bn_stats(xin, N * H * W_, C, stats[0], stats[1])
            stats = sync[0].reduce(stats, sync[1] + "/fwd")
            native.bn_finalize(stats[0], stats[1], float(N * H * W_ * sync[0].world_size), mr[0], mr[1], running_mean,
                               running_var, nbt, momentum, eps, C)
            native.bn_apply(xin, out, mr[0], mr[1], gamma, beta, N, H, W_, C, 0, 0)
            ctx.sync = sync
        elif training:
            work = torch.empty((2, C), device=dev, dtype=torch.float32)      # scratch: cleared inside the call
            native.bn_train_fwd(xin, out, gamma, beta, running_mean, running_var, nbt, momentum, eps, N, H, W_, C, 0, 0,
                                mr, work)
        else:
            native.bn_eval_stats(running_mean, running_var, eps, mr[0], mr[1], C)
            native.bn_apply(xin, out, mr[0], mr[1], gamma, beta, N, H, W_, C, 0, 0)
        ctx.save_for_backward(xin, mr, gamma, beta)
        ctx.cfg = (N, H, W_, C, bool(training))
        return from_nhwc(out)

    @staticmethod
    def backward(ctx, dout):
        xin, mr, gamma, beta = ctx.saved_tensors
        N, H, W_, C, training = ctx.cfg
        d = to_nhwc(dout)
        sums = _zeros((2, C), xin.device)
        native.bn_bwd_reduce(xin, d, mr[0], mr[1], gamma, beta, N, H, W_, C, 0, 0, sums[0], sums[1])
        dx = torch.empty_like(xin)
        tot = sums
        if ctx.sync is not None:
            tot = ctx.sync[0].reduce(sums, ctx.sync[1] + "/bwd")
            native.scale_copy(tot, tot, tot.numel(), 1.0 / ctx.sync[0].world_size)
        native.bn_bwd_apply(xin, d, mr[0], mr[1], gamma, beta, tot[0], tot[1], N, H, W_, C, 0, 0, int(training), dx)
        return from_nhwc(dx), sums[1], sums[0], None, None, None, None, None, None, None


class _SyncBatchNormRef(Function):
    """SyncBatchNorm in plain torch math (CPU / gloo): the same formulas as the native path, statistics summed over the
    ranks through ``StatSync`` (torch's own SyncBatchNorm refuses CPU tensors)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, momentum, eps, sync):
        C = x.shape[1]
        red = [0, 2, 3]
        n_local = x.numel() // C
        xf = x.float()
        stats = torch.stack([xf.sum(red), (xf * xf).sum(red)])
        stats = sync[0].reduce(_pad8(stats), sync[1] + "/fwd")[:, :C]
        n = n_local * sync[0].world_size
        mean = stats[0] / n
        var = (stats[1] / n - mean * mean).clamp_min(0)
        rstd = torch.rsqrt(var + eps)
        if running_mean is not None:
            with torch.no_grad():
                running_mean.mul_(1 - momentum).add_(momentum * mean)
                running_var.mul_(1 - momentum).add_(momentum * var * (n / max(n - 1, 1)))
        xhat = (xf - mean[None, :, None, None]) * rstd[None, :, None, None]
        ctx.save_for_backward(xhat, rstd, gamma)
        ctx.sync, ctx.n_local = sync, n_local
        return (xhat * gamma[None, :, None, None] + beta[None, :, None, None]).to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        xhat, rstd, gamma = ctx.saved_tensors
        C = xhat.shape[1]
        red = [0, 2, 3]
        dyf = dy.float()
        sums = torch.stack([dyf.sum(red), (dyf * xhat).sum(red)])
        tot = ctx.sync[0].reduce(_pad8(sums), ctx.sync[1] + "/bwd")[:, :C] / (ctx.n_local * ctx.sync[0].world_size)
        dx = (gamma * rstd)[None, :, None, None] * (dyf - tot[0][None, :, None, None] - xhat * tot[1][None, :, None, None])
        return dx.to(dy.dtype), sums[1], sums[0], None, None, None, None, None


def _pad8(t2c):
    """[2, C] -> [2, C'] with C' a multiple of 4, so that the flattened block is a multiple of 8 elements."""
    C = t2c.shape[1]
    if C % 4 == 0:
        return t2c.contiguous()
    return F.pad(t2c, (0, 4 - C % 4)).contiguous()


def batch_norm(x, gamma, beta, running_mean, running_var, num_batches_tracked, training, momentum=0.1, eps=1e-5, sync=None):
    """``sync = (StatSync, key)`` computes the batch statistics over all ranks (SyncBatchNorm)."""
    if x.is_cuda:
        return _BatchNorm.apply(x, gamma, beta, running_mean, running_var, num_batches_tracked, training, momentum, eps, sync)
    if training and num_batches_tracked is not None:
        num_batches_tracked.add_(1)
    if training and sync is not None and sync[0].world_size > 1:
        return _SyncBatchNormRef.apply(x, gamma, beta, running_mean, running_var, momentum, eps, sync)
    return F.batch_norm(x, running_mean, running_var, gamma, beta, training, momentum, eps)


class _ReLU(Function):
    @staticmethod
    def forward(ctx, x):
        xc = x if _dense(x) else x.contiguous()
        y = torch.empty_like(xc)
        native.relu_fwd(xc, y)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dyc = dy if (_dense(dy) and dy.stride() == y.stride()) else _match(dy, y)
        dx = torch.empty_like(y)
        native.relu_bwd(y, dyc, dx)
        return dx


def _dense(t):
    return t.is_contiguous() or (t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last))


def _match(t, like):
    """Materialise ``t`` with the memory layout of ``like`` (own kernels for the 4-D cases)."""
    if like.dim() == 4 and like.permute(0, 2, 3, 1).is_contiguous():
        return from_nhwc(to_nhwc(t))
    if like.dim() == 4:
        return to_nchw_contiguous(t)
    return t.contiguous()


def relu(x):
    if x.is_cuda and x.dtype == torch.float32:
        return _ReLU.apply(x)
    return F.relu(x)


class _MaxPool(Function):
    @staticmethod
    def forward(ctx, x):
        N, C, H, W_ = x.shape
        xin = to_nhwc(x)
        y = torch.empty((N, H // 2, W_ // 2, C), device=x.device, dtype=x.dtype)
        native.maxpool_fwd(xin, y, N, H, W_, C)
        ctx.save_for_backward(xin)
        ctx.cfg = (N, H, W_, C)
        return from_nhwc(y)

    @staticmethod
    def backward(ctx, dy):
        (xin,) = ctx.saved_tensors
        N, H, W_, C = ctx.cfg
        dx = torch.empty_like(xin)
        native.maxpool_bwd(xin, to_nhwc(dy), dx, N, H, W_, C)
        return from_nhwc(dx)


def max_pool2x2(x):
    if x.is_cuda and x.dtype == torch.float32:
        return _MaxPool.apply(x)
    return F.max_pool2d(x, 2)


class _SpatialMean(Function):
    @staticmethod
    def forward(ctx, x):
        N, C, H, W_ = x.shape
        y = torch.empty((N, C), device=x.device, dtype=x.dtype)
        native.spatial_mean_fwd(to_nhwc(x), y, N, H * W_, C)
        ctx.cfg = (N, C, H, W_)
        return y

    @staticmethod
    def backward(ctx, dy):
        N, C, H, W_ = ctx.cfg
        dx = torch.empty((N, H, W_, C), device=dy.device, dtype=dy.dtype)
        native.spatial_mean_bwd(dy.contiguous(), dx, N, H * W_, C)
        return from_nhwc(dx)


def spatial_mean(x):
    """``x.mean([2, 3])`` (multigpu.py:68)."""
    if x.is_cuda and x.dtype == torch.float32:
        return _SpatialMean.apply(x)
    return x.mean([2, 3])


def flatten_nchw(x):
    """``torch.flatten(x, 1)`` with NCHW element order (singlegpu.py:42) regardless of the memory layout."""
    if x.is_cuda and x.dim() == 4:
        return _FlattenNCHW.apply(x)
    return torch.flatten(x, 1)


class _FlattenNCHW(Function):
    @staticmethod
    def forward(ctx, x):
        ctx.shape = x.shape
        return to_nchw_contiguous(x).view(x.shape[0], -1)

    @staticmethod
    def backward(ctx, dy):
        N, C, H, W_ = ctx.shape
        g = dy.contiguous().view(N, C, H, W_)
        return from_nhwc(to_nhwc(g))


# ------------------------------------------------------------------------------------------------ linear
def _tma_ok(*pitches_bytes):
    return all(p % 16 == 0 for p in pitches_bytes)


def _gemm_nt(x2, w, out, bias, act):
    """out[M,N] = act(x2[M,K] @ w[N,K]^T + bias)"""
    M, K = x2.shape
    N = w.shape[0]
    es = x2.element_size()
    if x2.dtype in (torch.float32, torch.bfloat16) and _tma_ok(K * es) and x2.dtype == w.dtype and M >= 32 and N >= 16:
        native.gemm(x2, w, out, M, N, K, K, K, N, 0, 0, bias, act)
    elif x2.dtype == torch.float32:
        native.gemm_small(x2, w, out, M, N, K, K, 1, K, 1, N, bias, act)
    else:
        raise NotImplementedError("native linear: bf16 needs 16-byte aligned rows")


class _Linear(Function):
    """y = act(x W^T + b); act in {none, relu, gelu(tanh)} fused into the GEMM epilogue."""

    @staticmethod
    def forward(ctx, x, w, bias, act):
        xs = x.shape
        K = xs[-1]
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M, N = x2.shape[0], w.shape[0]
        wv = w.detach()
        bf = None
        if bias is not None:
            bf = bias.detach()
            if bf.dtype != torch.float32:
                bf = bf.float()
        out = torch.empty((M, N), device=x.device, dtype=x.dtype)
        pre = None
        if act == 2 and any(ctx.needs_input_grad[:3]):
            # GELU backward needs the pre-activation: keep it (bias fused), apply GELU in a second epilogue-free pass
            pre = torch.empty((M, N), device=x.device, dtype=x.dtype)
            _gemm_nt(x2, wv, pre, bf, 0)
            native.gelu_fwd(pre, out)
        else:
            _gemm_nt(x2, wv, out, bf, act)
        ctx.save_for_backward(x2, w, out if act == 1 else pre)
        ctx.cfg = (xs, act, bias is not None, bias.dtype if bias is not None else None)
        return out.view(*xs[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, w, aux = ctx.saved_tensors
        xs, act, has_bias, bias_dtype = ctx.cfg
        M, K = x2.shape
        N = w.shape[0]
        d = dy.reshape(M, N)
        if not d.is_contiguous():
            d = d.contiguous()
        if act == 1:
            dz = torch.empty_like(d)
            native.relu_bwd_any(aux, d, dz)
            d = dz
        elif act == 2:
            dz = torch.empty_like(d)
            native.gelu_bwd(aux, d, dz)
            d = dz
        es = d.element_size()
        wv = w.detach()
        dx = dw = db = None
        big = d.dtype in (torch.float32, torch.bfloat16) and _tma_ok(N * es, K * es) and M >= 32 and N >= 16 and K >= 16
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K), device=d.device, dtype=d.dtype)
            if big:
                native.gemm(d, wv, dx, M, K, N, N, K, K, 0, 1)            # dx = d[M,N] @ w[N,K]
            else:
                native.gemm_small(d, wv, dx, M, K, N, N, 1, 1, K, K)
            dx = dx.view(*xs)
        if ctx.needs_input_grad[1]:
            if big:
                # dw = d^T[N,M] @ x[M,K] (both operands MN-major).  A weight gradient has few output tiles and a long
                # reduction (the tokens): when 256x256 tiles cannot fill the 74 CTA pairs, split the reduction and
                # accumulate in fp32 (then cast) instead of shrinking the tiles.
                tiles = ((N + 255) // 256) * ((K + 255) // 256)
                split = 1 if tiles >= 60 or M < 2048 else max(1, min(74 // tiles, M // 1024))
                if d.dtype == torch.float32:
                    if split > 1:
                        dw = _zeros((N, K), d.device)
                        native.gemm(d, x2, dw, N, K, M, N, K, K, 1, 1, None, 0, split, 1)
                    else:
                        dw = torch.empty((N, K), device=d.device, dtype=torch.float32)
                        native.gemm(d, x2, dw, N, K, M, N, K, K, 1, 1)
                elif split > 1:
                    dw32 = _zeros((N, K), d.device)
                    native.gemm(d, x2, dw32, N, K, M, N, K, K, 1, 1, None, 0, split, 1)
                    dw = native.cast(dw32, w.dtype)
                else:
                    dw = torch.empty((N, K), device=d.device, dtype=w.dtype)
                    native.gemm(d, x2, dw, N, K, M, N, K, K, 1, 1)
            else:
                dw = torch.empty((N, K), device=d.device, dtype=torch.float32)
                native.gemm_small(d, x2, dw, N, K, M, 1, N, 1, K, K)
        if has_bias and ctx.needs_input_grad[2]:
            if d.dtype == torch.float32:
                db = torch.empty(N, device=d.device, dtype=torch.float32)
                native.colsum(d, db, M, N)
            else:
                db32 = torch.empty(N, device=d.device, dtype=torch.float32)
                native.colsum_any(d, db32, M, N)
                db = db32 if bias_dtype == torch.float32 else native.cast(db32, bias_dtype)
        return dx, dw, db, None


def linear(x, w, bias=None, act: int = 0):
    """``F.linear`` (+ fused activation: 0 none, 1 ReLU, 2 GELU-tanh)."""
    if x.is_cuda:
        return _Linear.apply(x, w, bias, act)
    y = F.linear(x, w, bias)
    if act == 1:
        y = F.relu(y)
    elif act == 2:
        y = F.gelu(y, approximate="tanh")
    return y


# ------------------------------------------------------------------------------------------------ losses
class _CrossEntropy(Function):
    """Mean cross-entropy over the first ``vocab`` columns of [..., C] logits (``vocab`` = C unless the vocabulary is
    padded).  Forward: one pass over the logits (online softmax) -> loss + per-row log-sum-exp.  Backward: one more pass
    writing dlogits = (softmax - onehot) * g / rows, with g read on the device."""

    @staticmethod
    def forward(ctx, logits, targets, ignore_index, vocab):
        C = logits.shape[-1]
        l2 = logits.reshape(-1, C)
        if not l2.is_contiguous():
            l2 = l2.contiguous()
        t = targets.reshape(-1)
        if t.dtype != torch.int64:
            raise TypeError("cross_entropy targets must be int64")
        t = t.contiguous()
        rows = l2.shape[0]
        cols = C if vocab is None else int(vocab)
        loss = torch.empty(1, device=logits.device, dtype=torch.float32)
        lse = torch.empty(rows, device=logits.device, dtype=torch.float32)
        native.ce_fwd(l2, t, lse, loss, rows, cols, C, ignore_index)
        ctx.save_for_backward(l2, t, lse)
        ctx.cfg = (logits.shape, cols, C, ignore_index)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        l2, t, lse = ctx.saved_tensors
        shape, cols, C, ignore_index = ctx.cfg
        gg = g if g.dtype == torch.float32 else g.float()
        dl = torch.empty_like(l2)
        native.ce_bwd(l2, t, lse, gg.contiguous(), dl, l2.shape[0], cols, C, ignore_index)
        return dl.view(shape), None, None, None


def cross_entropy(logits, targets, ignore_index: int = -100):
    """Mean softmax cross-entropy over all rows (``F.cross_entropy`` at multigpu.py:95); loss and dlogits come out of
    one fused kernel.  Note: like torch with no ignored rows; with ignored rows the mean is over *all* rows."""
    if logits.is_cuda:
        return _CrossEntropy.apply(logits, targets, ignore_index, None)
    return F.cross_entropy(logits.reshape(-1, logits.shape[-1]), targets.reshape(-1), ignore_index=ignore_index)


class _MSE(Function):
    @staticmethod
    def forward(ctx, pred, target):
        p = pred.contiguous()
        t = target.contiguous()
        loss = torch.empty(1, device=pred.device, dtype=torch.float32)
        dp = torch.empty_like(p) if ctx.needs_input_grad[0] else None
        native.mse(p, t, dp, loss)
        ctx.save_for_backward(dp)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        (dp,) = ctx.saved_tensors
        if dp is None:
            return None, None
        native.scale_by_device_scalar(dp, g.contiguous())
        return dp, None


def mse_loss(pred, target):
    if pred.is_cuda and pred.dtype == torch.float32:
        return _MSE.apply(pred, target.to(pred.dtype))
    return F.mse_loss(pred, target)


# ------------------------------------------------------------------------------------------------ transformer ops
def layer_norm(x, weight, bias, eps: float = 1e-5):
    if x.is_cuda:
        from . import transformer as T
        return T.LayerNormFn.apply(x, weight, bias, eps)
    return F.layer_norm(x, (x.shape[-1],), weight, bias, eps)


def add(a, b):
    """Residual add (own kernel on CUDA so the bf16 stream never touches a torch kernel)."""
    if a.is_cuda:
        from . import transformer as T
        return T.AddFn.apply(a, b)
    return a + b


def embedding_sum(idx, wte, wpe):
    """tok_emb[idx] + pos_emb[arange(T)] in one pass."""
    if wte.is_cuda:
        from . import transformer as T
        return T.EmbeddingSumFn.apply(idx, wte, wpe)
    T_ = idx.shape[1]
    return F.embedding(idx, wte) + wpe[:T_].unsqueeze(0)


def causal_attention(qkv, n_head: int):
    """qkv [B,T,3C] (q | k | v along the last dim, heads contiguous inside each) -> [B,T,C], causal softmax(QK^T/sqrt(d))V."""
    if qkv.is_cuda:
        from . import transformer as T
        return T.CausalAttentionFn.apply(qkv, n_head)
    B, T_, C3 = qkv.shape
    C = C3 // 3
    hd = C // n_head
    q, k, v = qkv.split(C, dim=2)
    q = q.view(B, T_, n_head, hd).transpose(1, 2)
    k = k.view(B, T_, n_head, hd).transpose(1, 2)
    v = v.view(B, T_, n_head, hd).transpose(1, 2)
    y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return y.transpose(1, 2).reshape(B, T_, C)


def cross_entropy_padded(logits, targets, vocab_size: int):
    """Cross-entropy over the first ``vocab_size`` columns of a vocabulary-padded logits tensor."""
    if logits.is_cuda:
        return _CrossEntropy.apply(logits, targets, -100, vocab_size)
    V = logits.shape[-1]
    l2 = logits.reshape(-1, V)[:, :vocab_size]
    return F.cross_entropy(l2.float(), targets.reshape(-1))


def dropout(x, p: float):
    if x.is_cuda:
        from . import transformer as T
        return T.DropoutFn.apply(x, p)
    return F.dropout(x, p, True)
