"""Layers with torch-compatible parameter names / state_dict layout whose math runs on the native kernels."""
from .modules import (  # noqa: F401
    Conv2d, BatchNorm2d, SyncBatchNorm, convert_sync_batchnorm, ReLU, MaxPool2d, Linear, Dropout, Flatten, CrossEntropyLoss, MSELoss,
)
