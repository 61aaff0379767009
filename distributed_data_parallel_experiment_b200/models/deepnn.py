"""``DeepNN`` of singlegpu.py:18-44 (dead code in the reference, kept runnable here): four biased 3x3 convs with ReLU
and two max-pools, then Linear(2048,512)-ReLU-Dropout(0.1)-Linear(512,num_classes).  1,186,986 parameters."""
from __future__ import annotations

import torch
from torch import nn

from .. import nn as dnn


class DeepNN(nn.Module):
    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.features = nn.Sequential(
            dnn.Conv2d(3, 128, kernel_size=3, padding=1),
            dnn.ReLU(),
            dnn.Conv2d(128, 64, kernel_size=3, padding=1),
            dnn.ReLU(),
            dnn.MaxPool2d(kernel_size=2, stride=2),
            dnn.Conv2d(64, 64, kernel_size=3, padding=1),
            dnn.ReLU(),
            dnn.Conv2d(64, 32, kernel_size=3, padding=1),
            dnn.ReLU(),
            dnn.MaxPool2d(kernel_size=2, stride=2),
        )
        self.classifier = nn.Sequential(
            dnn.Linear(2048, 512),
            dnn.ReLU(),
            dnn.Dropout(0.1),
            dnn.Linear(512, num_classes),
        )

    def forward(self, x):
        if x.is_cuda:
            f = self.features
            x = f[0](x, relu=True)       # bias + ReLU live in the conv epilogue
            x = f[2](x, relu=True)
            x = f[4](x)
            x = f[5](x, relu=True)
            x = f[7](x, relu=True)
            x = f[9](x)
            x = dnn.Flatten()(x)
            c = self.classifier
            x = c[0](x, act=1)
            x = c[2](x)
            return c[3](x)
        x = self.features(x)
        x = torch.flatten(x, 1)
        return self.classifier(x)
