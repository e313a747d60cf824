"""MaxPool2d / AdaptiveAvgPool2d(1) on channels_last activations — bf16 / fp16 under amp, fp32 in the reference's
fp32 modes — (SURVEY K-MAXPOOL, K-AVGPOOL); torch ops elsewhere.  (ATen's NHWC fp32 max-pool pair cost 13 ms of a
68 ms fp32 step at 1200 images on B200.)"""
import torch
import torch.nn.functional as F

from . import _native


def _lp_native(x):
    return (x.is_cuda and x.dim() == 4 and x.dtype in (torch.bfloat16, torch.float16, torch.float32) and x.shape[1] % 8 == 0
            and _native.available())


class _MaxPoolNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, ks, st, pd):
        with _native.guard(x.device):
            return _MaxPoolNative._forward_impl(ctx, x, ks, st, pd)

    @staticmethod
    def _forward_impl(ctx, x, ks, st, pd):
        N, C, H, W = x.shape
        x = x.contiguous(memory_format=torch.channels_last)
        Ho, Wo = (H + 2 * pd - ks) // st + 1, (W + 2 * pd - ks) // st + 1
        y = torch.empty((N, C, Ho, Wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty((N, Ho, Wo, C), dtype=torch.int8, device=x.device)
        _native.check(_native.lib().pdt_maxpool_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), _native.dtype_code(x.dtype),
                                                    N, H, W, C, ks, st, pd, _native.stream_ptr(x.device)), "pdt_maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.geom = (N, C, H, W, ks, st, pd)
        return y

    @staticmethod
    def backward(ctx, dy):
        with _native.guard(dy.device):
            return _MaxPoolNative._backward_impl(ctx, dy)

    @staticmethod
    def _backward_impl(ctx, dy):
        (idx,) = ctx.saved_tensors
        N, C, H, W, ks, st, pd = ctx.geom
        dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((N, C, H, W), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        _native.check(_native.lib().pdt_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), _native.dtype_code(dy.dtype),
                                                    N, H, W, C, ks, st, pd, _native.stream_ptr(dy.device)), "pdt_maxpool_bwd")
        return dx, None, None, None


class _AvgPoolNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        with _native.guard(x.device):
            return _AvgPoolNative._forward_impl(ctx, x)

    @staticmethod
    def _forward_impl(ctx, x):
        N, C, H, W = x.shape
        x = x.contiguous(memory_format=torch.channels_last)
        y = torch.empty((N, C), dtype=x.dtype, device=x.device)
        _native.check(_native.lib().pdt_avgpool_fwd(x.data_ptr(), y.data_ptr(), _native.dtype_code(x.dtype), N, H * W, C,
                                                    _native.stream_ptr(x.device)), "pdt_avgpool_fwd")
        ctx.geom = (N, C, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        with _native.guard(dy.device):
            return _AvgPoolNative._backward_impl(ctx, dy)

    @staticmethod
    def _backward_impl(ctx, dy):
        N, C, H, W = ctx.geom
        dy = dy.contiguous()
        dx = torch.empty((N, C, H, W), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        _native.check(_native.lib().pdt_avgpool_bwd(dy.data_ptr(), dx.data_ptr(), _native.dtype_code(dy.dtype), N, H * W, C,
                                                    _native.stream_ptr(dy.device)), "pdt_avgpool_bwd")
        return dx


def max_pool2d(x, kernel_size=3, stride=2, padding=1):
    if _lp_native(x):
        return _MaxPoolNative.apply(x, kernel_size, stride, padding)
    return F.max_pool2d(x, kernel_size, stride, padding)


def global_avg_pool(x):
    """[N,C,H,W] -> [N,C]"""
    if _lp_native(x):
        return _AvgPoolNative.apply(x)
    return torch.flatten(F.adaptive_avg_pool2d(x, 1), 1)
