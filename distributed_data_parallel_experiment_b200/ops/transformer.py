"""autograd Functions for the transformer ops on CUDA (LayerNorm, residual add, embedding, causal attention, padded
cross-entropy, dropout).  All kernels are the framework's own (csrc/kernels/transformer.cu, attention_aux.cu and the
batched tcgen05 GEMM); activations may be bf16 or fp32.
"""
from __future__ import annotations

import math

import torch
from torch.autograd import Function

import os

from . import native

_FLASH = os.environ.get("DDPX_FLASH_ATTN", "1") not in ("0", "")     # A/B switch: 0 = materialised-score path


def _zeros(shape, device, dtype=torch.float32):
    return native.zero_(torch.empty(shape, device=device, dtype=dtype))


def _to_dtype(t32, dtype):
    return t32 if dtype == torch.float32 else native.cast(t32, dtype)


class LayerNormFn(Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        C = x.shape[-1]
        x2 = x.reshape(-1, C)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        stats = torch.empty((2, rows), device=x.device, dtype=torch.float32)
        native.layernorm_fwd(x2, weight.detach(), bias.detach(), y, stats[0], stats[1], rows, C, eps)
        ctx.save_for_backward(x2, weight, stats)
        ctx.shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, stats = ctx.saved_tensors
        rows, C = x2.shape
        d = dy.reshape(rows, C)
        if not d.is_contiguous():
            d = d.contiguous()
        dx = torch.empty_like(x2)
        dgb = torch.empty((2, C), device=x2.device, dtype=torch.float32)
        scratch = torch.empty(native.layernorm_bwd_scratch_floats(rows, C), device=x2.device, dtype=torch.float32)
        native.layernorm_bwd(x2, d, weight.detach(), stats[0], stats[1], dx, dgb[0], dgb[1], rows, C, scratch)
        return dx.view(ctx.shape), _to_dtype(dgb[0], weight.dtype), _to_dtype(dgb[1], weight.dtype), None


class AddFn(Function):
    @staticmethod
    def forward(ctx, a, b):
        ac = a if a.is_contiguous() else a.contiguous()
        bc = b if b.is_contiguous() else b.contiguous()
        out = torch.empty_like(ac)
        native.add(ac, bc, out)
        return out

    @staticmethod
    def backward(ctx, g):
        return g, g


class EmbeddingSumFn(Function):
    @staticmethod
    def forward(ctx, idx, wte, wpe):
        B, T = idx.shape
        C = wte.shape[1]
        ic = idx.contiguous()
        out = torch.empty((B, T, C), device=wte.device, dtype=wte.dtype)
        native.embedding_fwd(ic, wte.detach(), wpe.detach(), out, B * T, T, C)
        ctx.save_for_backward(ic)
        ctx.cfg = (B, T, C, wte.shape[0], wpe.shape[0], wte.dtype)
        return out

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        B, T, C, V, NP, dtype = ctx.cfg
        d = dy if dy.is_contiguous() else dy.contiguous()
        dwte = _zeros((V, C), d.device)
        dwpe = _zeros((NP, C), d.device)
        native.embedding_bwd(idx, d, dwte, dwpe, B * T, T, C)
        return None, _to_dtype(dwte, dtype), _to_dtype(dwpe, dtype)


class CausalAttentionFn(Function):
    """softmax(causal(Q K^T / sqrt(d))) V from a packed qkv [B,T,3C] tensor.

    bf16 with 64-wide heads (every GPT-2 size): ONE fused tcgen05 kernel forward (scores and probabilities never leave
    tensor memory; output + per-row log-sum-exp) and two fused kernels backward that recompute the probabilities
    (csrc/kernels/attn_sm100.cu, attn_bwd_sm100.cu) - no [B,H,T,T] tensor exists.
    Other head sizes / fp32: batched tcgen05 GEMM for the scores (tiles above the diagonal are never computed), a causal
    softmax in place, batched GEMM for P V; the probabilities are kept for a backward of four more batched GEMMs.
    """

    @staticmethod
    def forward(ctx, qkv, n_head):
        B, T, C3 = qkv.shape
        C = C3 // 3
        hd = C // n_head
        q3 = qkv if qkv.is_contiguous() else qkv.contiguous()
        scale = 1.0 / math.sqrt(hd)
        if hd == 64 and qkv.dtype == torch.bfloat16 and _FLASH:
            Tp = (T + 127) // 128 * 128
            out = torch.empty((B, T, C), device=qkv.device, dtype=qkv.dtype)
            lse2 = torch.empty((B, n_head, Tp), device=qkv.device, dtype=torch.float32)
            native.attn_fwd(q3, out, lse2, B, T, n_head, scale)
            ctx.save_for_backward(q3, out, lse2)
            ctx.cfg = (B, T, C, n_head, hd, scale)
            ctx.flash = True
            return out
        ctx.flash = False
        BH = B * n_head
        P = torch.empty((B, n_head, T, T), device=qkv.device, dtype=qkv.dtype)
        q = q3.view(-1)
        k = q[C:]
        v = q[2 * C:]
        in_bs = (hd, T * C3)                 # batch1 = head (stride hd), batch2 = batch (stride T*3C)
        # S[b,h] = scale * Q K^T : A = Q [T, hd] (ld 3C), B = K [T, hd] (ld 3C), both K-major
        native.gemm_batched(q, k, P, T, T, hd, C3, C3, T, 0, 0, n_head, B, in_bs, in_bs, (T * T, n_head * T * T),
                            alpha=scale, causal=1)
        native.causal_softmax_fwd(P, BH * T, T)
        # O[b,t,h,:] = P V : A = P [T, T] K-major, B = V stored [T(k)][hd] -> MN-major; output into [B,T,C]
        out = torch.empty((B, T, C), device=qkv.device, dtype=qkv.dtype)
        native.gemm_batched(P, v, out, T, hd, T, T, C3, C, 0, 1, n_head, B, (T * T, n_head * T * T), in_bs, (hd, T * C),
                            causal=2)
        ctx.save_for_backward(q3, P)
        ctx.cfg = (B, T, C, n_head, hd, scale)
        return out

    @staticmethod
    def backward(ctx, dout):
        B, T, C, n_head, hd, scale = ctx.cfg
        do = dout if dout.is_contiguous() else dout.contiguous()
        if ctx.flash:
            q3, out, lse2 = ctx.saved_tensors
            dqkv = torch.empty_like(q3)
            delta = torch.empty_like(lse2)
            native.attn_bwd(q3, out, do, lse2, delta, dqkv, B, T, n_head, scale)
            return dqkv, None
        q3, P = ctx.saved_tensors
        C3 = 3 * C
        flat = q3.view(-1)
        q, k, v = flat, flat[C:], flat[2 * C:]
        dqkv = torch.empty_like(q3)
        dflat = dqkv.view(-1)
        dq, dk, dv = dflat, dflat[C:], dflat[2 * C:]
        in_bs = (hd, T * C3)
        p_bs = (T * T, n_head * T * T)
        o_bs = (hd, T * C)
        # dV[b,k,h,:] = P^T dO : A = P as [K=t_q][M=t_k] (MN-major), B = dO [K=t_q][N=hd] (MN-major)
        native.gemm_batched(P, do.view(-1), dv, T, hd, T, T, C, C3, 1, 1, n_head, B, p_bs, o_bs, in_bs)
        # dP = dO V^T : A = dO [T, hd] K-major (ld C), B = V [T, hd] K-major (ld 3C)
        dP = torch.empty_like(P)
        native.gemm_batched(do.view(-1), v, dP, T, T, hd, C, C3, T, 0, 0, n_head, B, o_bs, in_bs, p_bs, causal=1)
        # dS = P * (dP - rowsum(dP * P)) * scale   (in place, zeros above the diagonal)
        native.causal_softmax_bwd(P, dP, B * n_head * T, T, scale)
        # dQ = dS K : A = dS [T, T] K-major, B = K stored [T(k)][hd] -> MN-major
        native.gemm_batched(dP, k, dq, T, hd, T, T, C3, C3, 0, 1, n_head, B, p_bs, in_bs, in_bs, causal=2)
        # dK = dS^T Q : A = dS as [K=t_q][M=t_k] (MN-major), B = Q [K=t_q][N=hd] (MN-major)
        native.gemm_batched(dP, q, dk, T, hd, T, T, C3, C3, 1, 1, n_head, B, p_bs, in_bs, in_bs)
        return dqkv, None


class DropoutFn(Function):
    _counter = 0

    @staticmethod
    def forward(ctx, x, p):
        xc = x if x.is_contiguous() else x.contiguous()
        DropoutFn._counter += 1
        seed = (int(torch.initial_seed()) * 1000003 + DropoutFn._counter) & 0xFFFFFFFFFFFF
        y = torch.empty_like(xc)
        native.dropout(xc, y, p, seed)
        ctx.cfg = (p, seed)
        return y

    @staticmethod
    def backward(ctx, g):
        p, seed = ctx.cfg
        gc = g if g.is_contiguous() else g.contiguous()
        dx = torch.empty_like(gc)
        native.dropout(gc, dx, p, seed)       # same mask (counter-based RNG), same 1/(1-p) scale
        return dx, None


_BLOCK_CALLS = os.environ.get("DDPX_BLOCK_CALLS", "1") not in ("0", "")
_BLOCK_WS = {}       # (device index, M, C, B*H*Tp) -> (bf16 scratch [M, 9C], fp32 scratch): shared by every block, stream-ordered


def _block_workspace(dev, M, C, lse_elems):
    key = (dev.index, M, C, lse_elems)
    ws = _BLOCK_WS.get(key)
    if ws is None:
        _BLOCK_WS.clear()            # one shape at a time: a new (batch, sequence) shape replaces the old scratch
        # the fp32 scratch starts with the split-K accumulation area and comes with arrival counters: both must be zero
        # before the first use and are left zero by the kernels that use them
        ws = (torch.empty(M * 9 * C, device=dev, dtype=torch.bfloat16),
              native.zero_(torch.empty(4 * C * C + 13 * C + lse_elems + native.layernorm_bwd_scratch_floats(M, C),
                                       device=dev, dtype=torch.float32)),
              native.zero_(torch.empty(16 * (4 * C * C // 8192 + 64), device=dev, dtype=torch.int32)))
        _BLOCK_WS[key] = ws
    return ws


class GPT2BlockFn(Function):
    """A whole GPT-2 block (LayerNorm, qkv projection, fused causal attention, output projection, residual, LayerNorm,
    MLP with GELU, residual) as ONE autograd node over two native calls (csrc/kernels/blocks_lm.cu).  bf16, 64-wide
    heads, no dropout.  Activations needed by the backward live in one flat buffer; the backward's scratch is shared by
    all blocks."""

    # element offsets (in units of M*C) of the saved activations inside the flat buffer: LN1, qkv, attention, x1, LN2,
    # fc pre-activation, fc activation
    _OFFS = (0, 1, 4, 5, 6, 7, 11)
    _TOTAL = 15

    @staticmethod
    def forward(ctx, x, n_head, eps, *params):
        B, T, C = x.shape
        M = B * T
        xc = x if x.is_contiguous() else x.contiguous()
        dev = x.device
        Tp = (T + 127) // 128 * 128
        lse_elems = B * n_head * Tp
        acts = torch.empty(GPT2BlockFn._TOTAL * M * C, device=dev, dtype=torch.bfloat16)
        stats = torch.empty(4 * M + lse_elems, device=dev, dtype=torch.float32)
        out = torch.empty((B, T, C), device=dev, dtype=torch.bfloat16)
        tmp = _block_workspace(dev, M, C, lse_elems)[0]          # first M*C elements double as the forward's scratch
        base = acts.data_ptr()
        act_ptrs = [base + o * M * C * 2 for o in GPT2BlockFn._OFFS]
        pp = [p.data_ptr() for p in params]
        native.gpt2_block_fwd(xc, pp, act_ptrs, stats, tmp, out, B, T, C, n_head, eps)
        ctx.save_for_backward(xc, acts, stats, *params)
        ctx.cfg = (B, T, C, n_head, lse_elems)
        return out

    @staticmethod
    def backward(ctx, dout):
        xc, acts, stats, *params = ctx.saved_tensors
        B, T, C, n_head, lse_elems = ctx.cfg
        M = B * T
        dev = xc.device
        do = dout if dout.is_contiguous() else dout.contiguous()
        wsb, ws32, cnt = _block_workspace(dev, M, C, lse_elems)
        grads = [torch.empty_like(p) for p in params]
        dx = torch.empty_like(xc)
        base = acts.data_ptr()
        act_ptrs = [base + o * M * C * 2 for o in GPT2BlockFn._OFFS]
        native.gpt2_block_bwd(xc, do, [p.data_ptr() for p in params], act_ptrs, stats, [g.data_ptr() for g in grads], dx,
                              wsb, ws32, cnt, B, T, C, n_head)
        return (dx, None, None, *grads)


def gpt2_block_supported(x, n_head, params, dropout) -> bool:
    return (_BLOCK_CALLS and x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 3 and x.shape[-1] == n_head * 64
            and dropout == 0.0 and all(p.dtype == torch.bfloat16 and p.is_contiguous() for p in params))
