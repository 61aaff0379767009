"""Wide MLP of BASELINE.json ("8-layer 8192-wide bf16 MLP, batch 256/GPU"): ``depth`` Linear layers of width ``width``
with ReLU between them and a classifier head; every GEMM is the tcgen05 kernel with bias+ReLU fused in the epilogue."""
from __future__ import annotations

import torch
from torch import nn

from .. import nn as dnn


class MLP(nn.Module):
    def __init__(self, in_features: int = 8192, width: int = 8192, depth: int = 8, num_classes: int = 8192,
                 dtype: torch.dtype = torch.bfloat16):
        super().__init__()
        dims = [in_features] + [width] * (depth - 1) + [num_classes]
        self.layers = nn.ModuleList([dnn.Linear(dims[i], dims[i + 1]) for i in range(depth)])
        if dtype is not None:
            self.to(dtype)

    def forward(self, x):
        n = len(self.layers)
        for i, layer in enumerate(self.layers):
            x = layer(x, act=1 if i < n - 1 else 0)
        return x
