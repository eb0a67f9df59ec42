"""dau_conv: B200-native DAU convolution operator (drop-in for the reference's TF plugin package
`dau_conv`, plugins/tensorflow/dau_conv/__init__.py)."""
from . import _lib
from ._lib import DAUError, default_settings, get_kernels
from .functional import (DAULayerHandle, GraphedStep, dau_conv, dau_conv_fused, dau_conv_grad_op, dau_conv_op, get_handle)
from .layers import DAUConv1d, DAUConv2d, dau_conv2d, dau_grid_mean, zero_n_last

__all__ = ["DAUError", "default_settings", "get_kernels", "DAULayerHandle", "GraphedStep", "dau_conv", "dau_conv_fused",
           "dau_conv_op", "dau_conv_grad_op", "get_handle", "DAUConv2d", "DAUConv1d", "dau_conv2d",
           "dau_grid_mean", "zero_n_last"]
