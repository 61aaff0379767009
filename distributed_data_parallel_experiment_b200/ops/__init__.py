"""Differentiable ops over the native sm_100a kernels (CUDA) / torch math (CPU)."""
from . import native  # noqa: F401
from .functional import (  # noqa: F401
    conv_bn_act, conv3x3, batch_norm, relu, max_pool2x2, spatial_mean, flatten_nchw, linear, cross_entropy, mse_loss,
    to_nhwc, from_nhwc, to_nchw_contiguous,
)
