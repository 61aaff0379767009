"""GPT-2 (small 124M / medium / large / XL 1.5B) for the BASELINE.json language-model configs.

Parameter names follow the Hugging Face GPT-2 layout (``wte``, ``wpe``, ``h.{i}.ln_1``, ``attn.c_attn`` ...) with
``nn.Linear``-style [out, in] weights; the LM head is tied to ``wte``.  Every matmul is the tcgen05 GEMM (bias / GELU
fused in the epilogue), LayerNorm / softmax-cross-entropy / embedding are hand-written kernels, attention is ONE fused
tcgen05 kernel forward and two backward (csrc/kernels/attn_sm100.cu, attn_bwd_sm100.cu; no [B,H,T,T] tensor), and on
CUDA in bf16 a whole block runs as one autograd node over two native calls (csrc/kernels/blocks_lm.cu).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
from torch import nn

from .. import nn as dnn
from .. import ops
from ..ops import functional as FX


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    dropout: float = 0.0
    padded_vocab_size: int = 0     # 0 -> round vocab up to a multiple of 128 for the tensor-core LM head

    def __post_init__(self):
        if self.padded_vocab_size == 0:
            self.padded_vocab_size = (self.vocab_size + 127) // 128 * 128

    @staticmethod
    def small():
        return GPT2Config()

    @staticmethod
    def medium():
        return GPT2Config(n_embd=1024, n_layer=24, n_head=16)

    @staticmethod
    def large():
        return GPT2Config(n_embd=1280, n_layer=36, n_head=20)

    @staticmethod
    def xl():
        return GPT2Config(n_embd=1600, n_layer=48, n_head=25)

    @staticmethod
    def tiny():
        return GPT2Config(vocab_size=512, n_positions=128, n_embd=128, n_layer=2, n_head=2)


class LayerNorm(nn.LayerNorm):
    def forward(self, x):
        return FX.layer_norm(x, self.weight, self.bias, self.eps)


class Attention(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.n_head = cfg.n_head
        self.c_attn = dnn.Linear(cfg.n_embd, 3 * cfg.n_embd)
        self.c_proj = dnn.Linear(cfg.n_embd, cfg.n_embd)

    def forward(self, x):
        B, T, C = x.shape
        qkv = self.c_attn(x)                                   # [B, T, 3C]
        y = FX.causal_attention(qkv, self.n_head)              # [B, T, C]
        return self.c_proj(y)


class MLPBlock(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.c_fc = dnn.Linear(cfg.n_embd, 4 * cfg.n_embd)
        self.c_proj = dnn.Linear(4 * cfg.n_embd, cfg.n_embd)

    def forward(self, x):
        return self.c_proj(self.c_fc(x, act=2))                # GELU(tanh) fused into the first GEMM's epilogue


class Block(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.ln_1 = LayerNorm(cfg.n_embd)
        self.attn = Attention(cfg)
        self.ln_2 = LayerNorm(cfg.n_embd)
        self.mlp = MLPBlock(cfg)

    def _params(self):
        return (self.ln_1.weight, self.ln_1.bias, self.attn.c_attn.weight, self.attn.c_attn.bias, self.attn.c_proj.weight,
                self.attn.c_proj.bias, self.ln_2.weight, self.ln_2.bias, self.mlp.c_fc.weight, self.mlp.c_fc.bias,
                self.mlp.c_proj.weight, self.mlp.c_proj.bias)

    def forward(self, x):
        if x.is_cuda:
            from ..ops import transformer as TR
            params = self._params()
            if TR.gpt2_block_supported(x, self.attn.n_head, params, 0.0) and self.ln_1.eps == self.ln_2.eps:
                # the whole block as one autograd node over two native calls (csrc/kernels/blocks_lm.cu)
                return TR.GPT2BlockFn.apply(x, self.attn.n_head, self.ln_1.eps, *params)
        x = FX.add(x, self.attn(self.ln_1(x)))
        x = FX.add(x, self.mlp(self.ln_2(x)))
        return x


class GPT2(nn.Module):
    def __init__(self, cfg: GPT2Config = None, dtype: torch.dtype = torch.bfloat16):
        super().__init__()
        cfg = cfg or GPT2Config()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.padded_vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        self.h = nn.ModuleList([Block(cfg) for _ in range(cfg.n_layer)])
        self.ln_f = LayerNorm(cfg.n_embd)
        self.apply(self._init)
        for name, p in self.named_parameters():
            if name.endswith("c_proj.weight"):
                nn.init.normal_(p, mean=0.0, std=0.02 / math.sqrt(2 * cfg.n_layer))
        if dtype is not None:
            self.to(dtype)

    @staticmethod
    def _init(m):
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            if m.bias is not None:
                nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)

    def num_params(self) -> int:
        return sum(p.numel() for p in self.parameters())

    def forward(self, idx: torch.Tensor, targets: torch.Tensor = None):
        """idx [B,T] int64 -> logits [B,T,padded_vocab]; with ``targets`` returns the mean cross-entropy instead."""
        B, T = idx.shape
        x = FX.embedding_sum(idx, self.wte.weight, self.wpe.weight)      # tok + pos in one kernel
        for blk in self.h:
            x = blk(x)
        x = self.ln_f(x)
        logits = ops.linear(x, self.wte.weight, None)                     # tied LM head
        if targets is None:
            return logits
        return FX.cross_entropy_padded(logits, targets, self.cfg.vocab_size)
