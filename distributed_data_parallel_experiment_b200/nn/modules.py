"""Module containers.

Each class subclasses the torch container of the same name, so parameter/buffer names, shapes, initialisation and
``state_dict`` metadata are exactly torch's (checkpoints interchange with the reference), while
``forward`` dispatches to ``ops`` - hand-written sm_100a kernels on CUDA, torch math on CPU.
"""
from __future__ import annotations

import torch
from torch import nn

from .. import ops


class Conv2d(nn.Conv2d):
    """3x3 / stride 1 / padding 1 convolution (the only geometry the reference uses: multigpu.py:53, singlegpu.py:21-30).
    Other geometries run only on CPU tensors."""

    def _native_ok(self) -> bool:
        return (self.kernel_size == (3, 3) and self.stride == (1, 1) and self.padding == (1, 1)
                and self.dilation == (1, 1) and self.groups == 1 and self.padding_mode == "zeros")

    def forward(self, x: torch.Tensor, relu: bool = False) -> torch.Tensor:
        if x.is_cuda:
            if not self._native_ok():
                raise NotImplementedError("native Conv2d supports kernel 3, stride 1, padding 1 only")
            return ops.conv3x3(x, self.weight, self.bias, relu=relu)
        y = super().forward(x)
        return torch.relu(y) if relu else y


class BatchNorm2d(nn.BatchNorm2d):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            if not (self.affine and self.track_running_stats and self.momentum is not None):
                raise NotImplementedError("native BatchNorm2d needs affine=True, track_running_stats=True, fixed momentum")
            return ops.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var,
                                  self.num_batches_tracked, self.training, self.momentum, self.eps)
        return super().forward(x)


class ReLU(nn.ReLU):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.relu(x)


class MaxPool2d(nn.MaxPool2d):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            ks = self.kernel_size if isinstance(self.kernel_size, tuple) else (self.kernel_size,) * 2
            st = self.stride if isinstance(self.stride, tuple) else (self.stride,) * 2
            if ks != (2, 2) or st != (2, 2) or self.padding not in (0, (0, 0)):
                raise NotImplementedError("native MaxPool2d supports 2x2 stride 2 only")
            return ops.max_pool2x2(x)
        return super().forward(x)


class Linear(nn.Linear):
    def forward(self, x: torch.Tensor, act: int = 0) -> torch.Tensor:
        return ops.linear(x, self.weight, self.bias, act)


class Dropout(nn.Dropout):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and self.training and self.p > 0:
            return ops.functional.dropout(x, self.p)
        return super().forward(x)


class Flatten(nn.Module):
    """``torch.flatten(x, 1)`` in NCHW element order whatever the memory layout (singlegpu.py:42)."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.flatten_nchw(x)


class CrossEntropyLoss(nn.Module):
    def forward(self, logits, targets):
        return ops.cross_entropy(logits, targets)


class MSELoss(nn.Module):
    def forward(self, pred, target):
        return ops.mse_loss(pred, target)
