"""VGG-style conv-BN-ReLU network of the reference (multigpu.py:36-71 == singlegpu.py:47-82).

Same module tree and names (``backbone.conv{i}/bn{i}/relu{i}/pool{i}``, ``classifier``) so ``state_dict`` keys, shapes,
dtypes and initialisation are identical: 9,228,362 parameters in 26 tensors, 24 buffers.  On CUDA ``forward`` runs each
conv-BN-ReLU(-pool) group as ONE fused op: a tcgen05 implicit-GEMM convolution whose epilogue also accumulates the
BatchNorm batch statistics, followed by a single normalise+ReLU(+2x2 max-pool) pass over NHWC activations.
"""
from __future__ import annotations

from collections import OrderedDict, defaultdict

import torch
from torch import nn

from .. import nn as dnn
from .. import ops


class VGG(nn.Module):
    ARCH = [64, 128, 'M', 256, 256, 'M', 512, 512, 'M', 512, 512, 'M']

    def __init__(self) -> None:
        super().__init__()
        layers = []
        counts = defaultdict(int)

        def add(name: str, layer: nn.Module) -> None:
            layers.append((f"{name}{counts[name]}", layer))
            counts[name] += 1

        in_channels = 3
        self._plan = []  # (conv index, pooled?)
        for i, x in enumerate(self.ARCH):
            if x != 'M':
                add("conv", dnn.Conv2d(in_channels, x, 3, padding=1, bias=False))
                add("bn", dnn.BatchNorm2d(x))
                add("relu", dnn.ReLU(True))
                in_channels = x
                pooled = i + 1 < len(self.ARCH) and self.ARCH[i + 1] == 'M'
                self._plan.append((counts["conv"] - 1, pooled))
            else:
                add("pool", dnn.MaxPool2d(2))
        self.backbone = nn.Sequential(OrderedDict(layers))
        self.classifier = dnn.Linear(512, 10)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            return self._forward_fused(x)
        # backbone: [N, 3, 32, 32] => [N, 512, 2, 2]
        x = self.backbone(x)
        # avgpool: [N, 512, 2, 2] => [N, 512]
        x = x.mean([2, 3])
        # classifier: [N, 512] => [N, 10]
        return self.classifier(x)

    def _forward_fused(self, x: torch.Tensor) -> torch.Tensor:
        bb = self.backbone
        # one launch packs every layer's weights into the tensor-core layouts and clears the BatchNorm accumulators
        ops.functional.prepare_convs([getattr(bb, f"conv{idx}").weight for idx, _ in self._plan])
        for idx, pooled in self._plan:
            conv = getattr(bb, f"conv{idx}")
            bn = getattr(bb, f"bn{idx}")
            x = ops.conv_bn_act(x, conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                                bn.num_batches_tracked, bn.training, bn.momentum, bn.eps, relu=True, pool=pooled,
                                sync=getattr(bn, "stat_sync", None))
        x = ops.spatial_mean(x)
        return self.classifier(x)
