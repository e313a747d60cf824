"""Convolution and linear layers on channels_last low-precision activations.

``conv2d`` / ``linear`` take the module (for its fp32 master weight and
geometry).  Inside an ``amp.autocast`` region on a GPU they run in bf16/fp16
from the optimizer-maintained *shadow* weights (no per-forward weight cast,
SURVEY N11) and produce fp32 weight gradients; the implicit-GEMM tcgen05
kernels of ``csrc/conv.cu`` serve the shapes they support, every other shape
goes through ``aten::convolution`` with the same dtypes and layouts.
Outside autocast (fp32 training, validation — the reference validates without
autocast) they are ``F.conv2d`` / ``F.linear``.
"""
import torch
import torch.nn.functional as F

from .. import amp
from . import _native

_seen_fallbacks = set()


def note_fallback(kind, *key):
    """Say ONCE per (kind, shape) that a layer is served by cuDNN / cuBLAS through aten instead of the package's
    sm_100a kernels (groups, dilation, bias, channel counts the implicit-GEMM tiles do not take, non-ResNet
    models), so nobody benchmarks the library path by accident.  PDT_QUIET_FALLBACK=1 silences it."""
    k = (kind,) + tuple(key)
    if k in _seen_fallbacks:
        return
    _seen_fallbacks.add(k)
    import os
    if os.environ.get("PDT_QUIET_FALLBACK", "0") != "1":
        import logging
        logging.getLogger("pdt_b200").warning("[pdt] %s falls back to the torch/cuDNN/cuBLAS path for %s", kind, key)


_conv_native = None   # set by ops.conv_native when the tcgen05 kernels are importable
_stem_native = None   # (supported(x, conv), run(x, conv, w_lp))
_linear_native = None


def _tf32_on():
    from . import conv_native
    return conv_native.tf32_enabled()


def shadow_of(p, dtype):
    arena = getattr(p, "_pdt_arena", None)
    if arena is not None and arena.shadow_dtype == dtype:
        return arena.shadow_view(p)
    return None


class _ConvLPTorch(torch.autograd.Function):
    """Low-precision conv through aten (cuDNN) with an fp32 weight gradient."""

    @staticmethod
    def forward(ctx, x, weight, w_lp, stride, padding):
        ctx.save_for_backward(x, w_lp)
        ctx.geom = (stride, padding)
        return torch.ops.aten.convolution(x, w_lp, None, stride, padding, (1, 1), False, (0, 0), 1)

    @staticmethod
    def backward(ctx, dy):
        x, w_lp = ctx.saved_tensors
        stride, padding = ctx.geom
        dy = dy.contiguous(memory_format=torch.channels_last)
        dx, dw, _ = torch.ops.aten.convolution_backward(
            dy, x, w_lp, None, stride, padding, (1, 1), False, (0, 0), 1,
            (ctx.needs_input_grad[0], True, False))
        return dx, dw.float(), None, None, None


def conv2d(x, conv, bnpack=None, consume=None, produce=None):
    """``conv(x)``.  With ``bnpack`` (see conv_native.conv_native) the native kernel may return
    ``(y, mean, invstd)`` — batch statistics of y for the BatchNorm that follows.  ``consume`` / ``produce``:
    roles of this convolution in a residual block's gradient link (conv_native.GradLink); only the native
    kernels take them, every other path ignores them (the link then stays unclaimed / unfilled)."""
    lp = amp.active_dtype()
    plain = (not x.is_cuda or conv.groups != 1 or conv.dilation != (1, 1) or conv.bias is not None
             or conv.padding_mode != "zeros" or isinstance(conv.padding, str))
    if x.dtype == torch.uint8:
        # raw pixels (data.build_datasets(uint8=True)): the native stem normalises inside its layout kernel,
        # every other path normalises here with the same arithmetic
        if not plain and _stem_native is not None and _native.available() and _stem_native[0](x, conv) \
                and (lp is not None or _tf32_on()):
            w = (shadow_of(conv.weight, lp) if lp is not None else None)
            if w is None:
                w = conv.weight.detach() if lp is None else conv.weight.detach().to(dtype=lp, memory_format=torch.channels_last)
            return _stem_native[1](x, conv, w, bnpack)
        from .input_norm import normalize_u8
        x = normalize_u8(x)
        if x.dim() == 4:
            x = x.contiguous(memory_format=torch.channels_last)
    if lp is None and not plain and x.dtype == torch.float32 and _conv_native is not None and _native.available():
        # fp32 training (DataParallel / plain DDP): tcgen05 kind::tf32 where torch would allow TF32
        w = conv.weight.detach()
        if _stem_native is not None and _stem_native[0](x, conv) and _tf32_on():
            return _stem_native[1](x, conv, w, bnpack)
        if _tf32_on() and x.dim() == 4:
            xc = x if x.is_contiguous(memory_format=torch.channels_last) else x.contiguous(memory_format=torch.channels_last)
            y = _conv_native(xc, conv.weight, w, conv.stride, conv.padding, bnpack, consume, produce)
            if y is not None:
                return y
    if lp is None or plain:
        if x.is_cuda:
            note_fallback("conv2d", tuple(conv.weight.shape), conv.stride, conv.padding, conv.groups, conv.dilation,
                          conv.bias is not None, str(x.dtype))
        return conv._conv_forward(x, conv.weight, conv.bias)
    w_lp = shadow_of(conv.weight, lp)
    if w_lp is None:
        w_lp = conv.weight.detach().to(dtype=lp, memory_format=torch.channels_last)
    if _stem_native is not None and _native.available() and _stem_native[0](x, conv):
        return _stem_native[1](x, conv, w_lp, bnpack)      # fp32 NCHW images: cast + layout change happen in the stem kernel
    if x.dtype != lp or not x.is_contiguous(memory_format=torch.channels_last):
        x = x.to(dtype=lp, memory_format=torch.channels_last)
    if _conv_native is not None and _native.available():
        y = _conv_native(x, conv.weight, w_lp, conv.stride, conv.padding, bnpack, consume, produce)
        if y is not None:
            return y
    note_fallback("conv2d(amp)", tuple(conv.weight.shape), conv.stride, conv.padding, str(x.dtype))
    return _ConvLPTorch.apply(x, conv.weight, w_lp, conv.stride, conv.padding)


class _LinearLP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, w_lp):
        ctx.save_for_backward(x, w_lp)
        ctx.has_bias = bias is not None
        return F.linear(x, w_lp, bias.to(x.dtype) if bias is not None else None)

    @staticmethod
    def backward(ctx, dy):
        x, w_lp = ctx.saved_tensors
        dx = dy @ w_lp if ctx.needs_input_grad[0] else None
        dw = (dy.t() @ x).float()
        db = dy.float().sum(0) if ctx.has_bias else None
        return dx, dw, db, None


def linear(x, fc):
    lp = amp.active_dtype()
    if lp is None and x.is_cuda and x.dtype == torch.float32 and _linear_native is not None and _native.available() \
            and _linear_native[0](x, fc):
        return _linear_native[1](x, fc, fc.weight.detach())
    if lp is None or not x.is_cuda:
        if x.is_cuda and torch.is_grad_enabled():
            note_fallback("linear", tuple(fc.weight.shape), str(x.dtype))
        return F.linear(x, fc.weight, fc.bias)
    if x.dtype != lp:
        x = x.to(lp)
    w_lp = shadow_of(fc.weight, lp)
    if w_lp is None:
        w_lp = fc.weight.detach().to(lp)
    if _linear_native is not None and _native.available() and _linear_native[0](x, fc):
        return _linear_native[1](x, fc, w_lp)
    note_fallback("linear(amp)", tuple(fc.weight.shape), str(x.dtype))
    return _LinearLP.apply(x, fc.weight, fc.bias, w_lp)
