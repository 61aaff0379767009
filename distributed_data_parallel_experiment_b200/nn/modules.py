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


class SyncBatchNorm(BatchNorm2d):
    """BatchNorm2d whose batch statistics span all ranks (the option the reference leaves commented out at
    multigpu.py:127).  Same parameters, buffers and state_dict keys as BatchNorm2d; with one process it IS BatchNorm2d.
    ``sync_key`` names the layer's statistic buffers on the P2P backend (set by ``convert_sync_batchnorm``)."""

    sync_key = "bn"

    @property
    def stat_sync(self):
        """(StatSync, key) while training in a multi-rank job, else None."""
        import torch.distributed as dist
        if not (self.training and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
            return None
        from ..parallel.syncbn import default_sync
        return (default_sync(), self.sync_key)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        sync = self.stat_sync
        if sync is None:
            return super().forward(x)
        if not (self.affine and self.track_running_stats and self.momentum is not None):
            raise NotImplementedError("SyncBatchNorm needs affine=True, track_running_stats=True, fixed momentum")
        return ops.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var, self.num_batches_tracked,
                              self.training, self.momentum, self.eps, sync=sync)


def convert_sync_batchnorm(module: nn.Module, _prefix: str = "") -> nn.Module:
    """Replace every BatchNorm2d in ``module`` by a SyncBatchNorm sharing its parameters and buffers (counterpart of
    ``torch.nn.SyncBatchNorm.convert_sync_batchnorm``).  Call before wrapping in DataParallel, on every rank."""
    for name, child in list(module.named_children()):
        full = f"{_prefix}.{name}" if _prefix else name
        if isinstance(child, nn.BatchNorm2d) and not isinstance(child, SyncBatchNorm):
            new = SyncBatchNorm(child.num_features, child.eps, child.momentum, child.affine, child.track_running_stats)
            if child.affine:
                new.weight, new.bias = child.weight, child.bias
            if child.track_running_stats:
                new.running_mean, new.running_var = child.running_mean, child.running_var
                new.num_batches_tracked = child.num_batches_tracked
            new.training = child.training
            new.sync_key = full
            setattr(module, name, new)
        else:
            convert_sync_batchnorm(child, full)
    return module


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
