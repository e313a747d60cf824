"""Fused cross entropy + top-1 accuracy (SURVEY K-CE; distributed.py:247,250).

``cross_entropy_with_accuracy(logits, target)`` returns ``(loss, acc1)``: the
mean NLL (differentiable) and the top-1 hit *fraction* (utils.py:111).  One
kernel produces loss, accuracy and the unit gradient; backward is a scale+cast.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _native

_scratch = {}


def _ce_scratch(device, B):
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    need = _native.lib().pdt_ce_scratch_bytes(max(B, 4096))
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.zeros(need, dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


class _CENative(torch.autograd.Function):
    last_pair = None

    @staticmethod
    def forward(ctx, logits, target):
        with _native.guard(logits.device):
            return _CENative._forward_impl(ctx, logits, target)

    @staticmethod
    def _forward_impl(ctx, logits, target):
        B, K = logits.shape
        logits = logits.contiguous()
        out = torch.empty(2, dtype=torch.float32, device=logits.device)
        dunit = torch.empty(B, K, dtype=torch.float32, device=logits.device)
        _native.check(_native.lib().pdt_ce_fwd(
            logits.data_ptr(), _native.dtype_code(logits.dtype), target.data_ptr(), B, K, dunit.data_ptr(),
            _ce_scratch(logits.device, B).data_ptr(), out.data_ptr(), _native.stream_ptr(logits.device)), "pdt_ce_fwd")
        ctx.save_for_backward(dunit)
        ctx.dtype = logits.dtype
        loss, acc = out[0], out[1]
        ctx.mark_non_differentiable(acc)
        _CENative.last_pair = out            # {loss, acc} as one 8-byte device vector (engine.StepRunner's meters)
        return loss, acc

    @staticmethod
    def backward(ctx, gloss, _gacc):
        with _native.guard(gloss.device):
            return _CENative._backward_impl(ctx, gloss, _gacc)

    @staticmethod
    def _backward_impl(ctx, gloss, _gacc):
        (dunit,) = ctx.saved_tensors
        g = gloss.to(torch.float32).contiguous()
        d = torch.empty(dunit.shape, dtype=ctx.dtype, device=dunit.device)
        _native.check(_native.lib().pdt_scale_cast(dunit.data_ptr(), g.data_ptr(), d.data_ptr(),
                                                   _native.dtype_code(ctx.dtype), dunit.numel(),
                                                   _native.stream_ptr(dunit.device)), "pdt_scale_cast")
        return d, None


def cross_entropy_with_accuracy(logits, target):
    if logits.is_cuda and _native.available() and logits.dim() == 2 and target.dtype == torch.int64 \
            and logits.dtype in (torch.float32, torch.bfloat16, torch.float16):
        return _CENative.apply(logits, target.contiguous())
    loss = F.cross_entropy(logits.float(), target)
    with torch.no_grad():
        acc = (logits.argmax(1) == target).float().mean()
    return loss, acc


class CrossEntropyLoss(nn.Module):
    """``nn.CrossEntropyLoss()`` whose forward also leaves ``last_acc1`` behind, so the trainer
    gets utils.accuracy for free instead of a topk + eq + sum chain."""

    def __init__(self):
        super().__init__()
        self.last_acc1 = None
        self.last_pair = None       # float32[2] = {loss, top-1 fraction} of the last forward, one tensor

    def forward(self, logits, target):
        _CENative.last_pair = None
        loss, acc = cross_entropy_with_accuracy(logits, target)
        self.last_acc1 = acc.detach()
        pair = _CENative.last_pair
        self.last_pair = pair if pair is not None else torch.stack([loss.detach().float(), acc.detach().float()])
        return loss
