"""Functional ops: autograd.Functions over the sm_100a kernels, with torch oracles."""
from .bn import bn_act, bn_relu_maxpool, conv_bn_act, conv_bn_relu_maxpool
from .conv import conv2d, linear
from . import conv_native  # noqa: F401  (registers the tcgen05 path with ops.conv)
from .loss import CrossEntropyLoss, cross_entropy_with_accuracy
from .pool import global_avg_pool, max_pool2d


def prepare_backward(param):
    """Per-step preparation the backward kernels can share (call at the start of a training forward with any
    parameter of the model): re-lays out every data-gradient filter of the parameter's arena in one launch."""
    from . import _native
    arena = getattr(param, "_pdt_arena", None)
    if arena is not None and _native.available():
        conv_native.prepare_dgrad_weights(arena)


__all__ = ["bn_act", "bn_relu_maxpool", "conv_bn_act", "conv_bn_relu_maxpool", "conv2d", "linear", "CrossEntropyLoss", "cross_entropy_with_accuracy",
           "global_avg_pool", "max_pool2d"]
