"""Toy regressor of the upstream tutorial the reference was forked from: ``Linear(20, 1)`` trained with MSE on
``datautils.MyTrainDataset`` (BASELINE.json config #1)."""
from __future__ import annotations

from torch import nn

from .. import nn as dnn


class ToyLinear(nn.Module):
    def __init__(self, in_features: int = 20, out_features: int = 1):
        super().__init__()
        self.net = dnn.Linear(in_features, out_features)

    def forward(self, x):
        return self.net(x)
