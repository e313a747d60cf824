"""tcgen05 implicit-GEMM convolution (``csrc/conv.cu``) behind ``ops.conv2d``.

Forward and the stride-1 data gradient run on the native kernel; shapes it does not cover
(the 3-channel stem, strided data gradients, and for now the weight gradient) go through
``aten::convolution_backward`` with the same low-precision operands.  ``PDT_CONV=torch``
disables the native path (A/B measurements, debugging).
"""
import os

import torch

from . import _native
from . import conv as _conv

_vp, _i = _native._vp, _native._i
_native.register("pdt_conv_fprop", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp])
_native.register("pdt_conv_wgrad", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp])


def supported(x, weight, stride, padding):
    K, C, R, S = weight.shape
    return (x.is_cuda and x.dtype in (torch.bfloat16, torch.float16) and C % 64 == 0 and K % 8 == 0
            and stride[0] == stride[1] and padding[0] == padding[1] and R == S
            and os.environ.get("PDT_CONV", "native") == "native")


def fprop(x, w_lp, stride, pad):
    """x: [N,C,H,W] channels_last, w_lp: [K,C,R,S] channels_last -> y [N,K,P,Q] channels_last."""
    N, C, H, W = x.shape
    K, _, R, S = w_lp.shape
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    y = torch.empty((N, K, P, Q), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    _native.check(_native.lib().pdt_conv_fprop(x.data_ptr(), w_lp.data_ptr(), y.data_ptr(), _native.dtype_code(x.dtype),
                                               N, H, W, C, K, R, S, stride, pad, _native.stream_ptr(x.device)),
                  "pdt_conv_fprop")
    return y


def wgrad(x, dy, weight_shape, stride, pad, out=None):
    """fp32 dW [K,C,R,S] (channels_last, i.e. KRSC) from x [N,C,H,W] and dy [N,K,P,Q], both channels_last."""
    N, C, H, W = x.shape
    K, _, R, S = weight_shape
    if out is None:
        out = torch.empty((K, C, R, S), dtype=torch.float32, device=x.device, memory_format=torch.channels_last)
    out.zero_()
    _native.check(_native.lib().pdt_conv_wgrad(x.data_ptr(), dy.data_ptr(), out.data_ptr(), _native.dtype_code(x.dtype),
                                               N, H, W, C, K, R, S, stride, pad, _native.stream_ptr(x.device)),
                  "pdt_conv_wgrad")
    return out


def wgrad_supported(C, K):
    return C % 64 == 0 and K % 64 == 0 and os.environ.get("PDT_WGRAD", "native") == "native"


class _ConvNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, w_lp, stride, padding):
        ctx.save_for_backward(x, w_lp)
        ctx.geom = (stride, padding)
        return fprop(x, w_lp, stride[0], padding[0])

    @staticmethod
    def backward(ctx, dy):
        x, w_lp = ctx.saved_tensors
        stride, padding = ctx.geom
        dy = dy.contiguous(memory_format=torch.channels_last)
        K, C, R, S = w_lp.shape
        need_dx = ctx.needs_input_grad[0]
        dx = None
        native_dx = need_dx and stride[0] == 1 and K % 64 == 0 and C % 8 == 0 and R - 1 - padding[0] >= 0
        if native_dx:
            # dx = conv(dy, flip(w)^T, pad' = R-1-pad): the same kernel with the roles of C and K swapped
            w_t = w_lp.flip(2, 3).transpose(0, 1).contiguous(memory_format=torch.channels_last)
            dx = fprop(dy, w_t, 1, R - 1 - padding[0])
        native_dw = wgrad_supported(C, K)
        dw = wgrad(x, dy, w_lp.shape, stride[0], padding[0]) if native_dw else None
        if (need_dx and not native_dx) or not native_dw:
            dx2, dw2, _ = torch.ops.aten.convolution_backward(
                dy, x, w_lp, None, stride, padding, (1, 1), False, (0, 0), 1,
                (need_dx and not native_dx, not native_dw, False))
            if not native_dx:
                dx = dx2
            if not native_dw:
                dw = dw2.float()
        return dx, dw, None, None, None


def conv_native(x, weight, w_lp, stride, padding):
    if not supported(x, weight, stride, padding):
        return None
    if not w_lp.is_contiguous(memory_format=torch.channels_last):
        w_lp = w_lp.contiguous(memory_format=torch.channels_last)
    return _ConvNative.apply(x, weight, w_lp, tuple(stride), tuple(padding))


_conv._conv_native = conv_native
