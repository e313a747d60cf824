"""Functional ops: autograd.Functions over the sm_100a kernels, with torch oracles."""
from .bn import bn_act, bn_relu_maxpool, conv_bn_act
from .conv import conv2d, linear
from . import conv_native  # noqa: F401  (registers the tcgen05 path with ops.conv)
from .loss import CrossEntropyLoss, cross_entropy_with_accuracy
from .pool import global_avg_pool, max_pool2d

__all__ = ["bn_act", "bn_relu_maxpool", "conv_bn_act", "conv2d", "linear", "CrossEntropyLoss", "cross_entropy_with_accuracy",
           "global_avg_pool", "max_pool2d"]
