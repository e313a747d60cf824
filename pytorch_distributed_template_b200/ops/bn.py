"""Fused BatchNorm(+residual)(+ReLU) on channels_last activations.

Native path: ``csrc/bn.cu`` (two kernels forward, two backward; the SyncBN
cross-rank exchange happens inside the statistics kernels).  Oracle / CPU path:
``F.batch_norm`` + add + relu, and for SyncBN an all-reduce of raw moments.
Reference semantics: eps 1e-5, momentum 0.1, unbiased running variance
(SURVEY §2.4 "Math that the fused kernels must match").
"""
import os

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import _native

_scratch = {}


def _bn_scratch(device, C):
    """One scratch per (device, stream): kernels on a stream never overlap."""
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    need = _native.lib().pdt_bn_scratch_bytes(max(C, 512))
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.zeros(need, dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def _param_grad_slots(params, C, device):
    """Where the BN backward kernel writes dgamma / dbeta: straight into the gradient arena when the
    parameters live in one and it is clean (ParamArena.direct_grad), else fresh tensors."""
    out = []
    for p in params:
        arena = getattr(p, "_pdt_arena", None)
        slot = arena.direct_grad(p) if arena is not None else None
        out.append(slot if slot is not None else torch.empty(C, dtype=torch.float32, device=device))
    return out


def _native_ok(x, C):
    if not (x.is_cuda and x.dim() == 4 and _native.available()):
        return False
    if x.dtype not in (torch.bfloat16, torch.float16, torch.float32):
        return False
    v = 4 if x.dtype == torch.float32 else 8
    cg = C // v
    return C % v == 0 and cg >= 1 and (cg & (cg - 1)) == 0 and cg <= 512 and 2 * C <= 4608


import threading

_tls = threading.local()     # DataParallel runs replicas in worker threads: the pending list is per thread


class batched_counters:
    """Inside this context the ``num_batches_tracked += 1`` of every BatchNorm layer is deferred and applied as
    ONE multi-tensor launch on exit (20 one-element launches per ResNet-18 forward otherwise)."""

    def __enter__(self):
        self._outer = getattr(_tls, "pending", None)
        _tls.pending = []
        return self

    def __exit__(self, *exc):
        pending, _tls.pending = _tls.pending, self._outer
        if pending:
            if self._outer is not None:
                self._outer.extend(pending)
            elif len(pending) == 1 or any(not t.is_cuda for t in pending):
                for t in pending:
                    t.add_(1)
            else:
                torch._foreach_add_(pending, 1)
        return False


def _ema_factor(bn):
    """exponential_average_factor of nn.BatchNorm (torch/nn/modules/batchnorm.py): ``momentum``, or with
    ``momentum=None`` the cumulative moving average 1 / num_batches_tracked (counted including this batch).
    Call BEFORE ``_bump``; the cumulative form reads the counter on the host, like torch does."""
    if bn.momentum is not None:
        return bn.momentum
    if bn.num_batches_tracked is None:
        return 0.0
    return 1.0 / float(int(bn.num_batches_tracked.item()) + 1)


def _bump(counter):
    pending = getattr(_tls, "pending", None)
    if pending is None:
        counter.add_(1)
    else:
        pending.append(counter)


class _BNActNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, res, weight, bias, running_mean, running_var, momentum, eps, relu, comm, pre_mean=None,
                pre_invstd=None, res_link=None):
        # res_link: conv_native.GradLink whose consumer (the block's first convolution, fed by the same tensor as
        # ``res``) adds the shortcut gradient in its data-gradient epilogue; backward then deposits it there
        # instead of returning it.
        ctx.res_link = res_link if (res is not None and res_link is not None and res_link.claimed) else None
        with _native.guard(x.device):
            return _BNActNative._forward_impl(ctx, x, res, weight, bias, running_mean, running_var, momentum, eps, relu,
                                              comm, pre_mean, pre_invstd)

    @staticmethod
    def _forward_impl(ctx, x, res, weight, bias, running_mean, running_var, momentum, eps, relu, comm, pre_mean=None,
                      pre_invstd=None):
        N, C, H, W = x.shape
        M = N * H * W
        x = x.contiguous(memory_format=torch.channels_last)
        if res is not None:
            res = res.contiguous(memory_format=torch.channels_last)
        lib, s, dt = _native.lib(), _native.stream_ptr(x.device), _native.dtype_code(x.dtype)
        if pre_mean is not None:       # statistics came out of the producing convolution's epilogue
            mean, invstd = pre_mean, pre_invstd
        else:
            mean = torch.empty(C, dtype=torch.float32, device=x.device)
            invstd = torch.empty_like(mean)
            scratch = _bn_scratch(x.device, C)
            _native.check(lib.pdt_bn_fwd_stats(x.data_ptr(), M, C, dt, scratch.data_ptr(), comm,
                                               _native.ptr(running_mean), _native.ptr(running_var), momentum, eps,
                                               mean.data_ptr(), invstd.data_ptr(), s), "pdt_bn_fwd_stats")
        y = torch.empty_like(x)
        _native.check(lib.pdt_bn_fwd_apply(x.data_ptr(), _native.ptr(res), y.data_ptr(), M, C, dt, mean.data_ptr(),
                                           invstd.data_ptr(), weight.data_ptr(), bias.data_ptr(), int(relu), 0, eps, s),
                      "pdt_bn_fwd_apply")
        # the ReLU mask is recomputed from x in backward unless a residual entered the ReLU
        ctx.save_for_backward(x, y if (relu and res is not None) else None, weight, bias, mean, invstd)
        ctx.relu, ctx.has_res, ctx.comm = relu, res is not None, comm
        ctx.params = (weight, bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        with _native.guard(dy.device):
            return _BNActNative._backward_impl(ctx, dy)

    @staticmethod
    def _backward_impl(ctx, dy):
        x, y, weight, bias, mean, invstd = ctx.saved_tensors
        N, C, H, W = x.shape
        M = N * H * W
        dy = dy.contiguous(memory_format=torch.channels_last)
        lib, s, dt = _native.lib(), _native.stream_ptr(x.device), _native.dtype_code(x.dtype)
        dgamma, dbeta = _param_grad_slots(ctx.params, C, x.device)
        k = torch.empty(2 * C, dtype=torch.float32, device=x.device)
        scratch = _bn_scratch(x.device, C)
        _native.check(lib.pdt_bn_bwd_reduce(dy.data_ptr(), _native.ptr(y), x.data_ptr(), M, C, dt, mean.data_ptr(),
                                            invstd.data_ptr(), weight.data_ptr(), bias.data_ptr(), int(ctx.relu),
                                            scratch.data_ptr(), ctx.comm, dgamma.data_ptr(), dbeta.data_ptr(),
                                            k.data_ptr(), s), "pdt_bn_bwd_reduce")
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if ctx.has_res else None
        _native.check(lib.pdt_bn_bwd_apply(dy.data_ptr(), _native.ptr(y), x.data_ptr(), dx.data_ptr(), _native.ptr(dres),
                                           M, C, dt, mean.data_ptr(), invstd.data_ptr(), weight.data_ptr(), bias.data_ptr(),
                                           k.data_ptr(), int(ctx.relu), s), "pdt_bn_bwd_apply")
        if ctx.res_link is not None and dres is not None and ctx.needs_input_grad[1]:
            ctx.res_link.grad, dres = dres, None
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None, None, None, None


class _BNReluPoolNative(torch.autograd.Function):
    """maxpool(relu(bn(x))) without ever writing the normalised tensor (ResNet stem tail)."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps, ks, st, pd, comm, pre_mean=None,
                pre_invstd=None):
        with _native.guard(x.device):
            return _BNReluPoolNative._forward_impl(ctx, x, weight, bias, running_mean, running_var, momentum, eps, ks, st, pd,
                                                   comm, pre_mean, pre_invstd)

    @staticmethod
    def _forward_impl(ctx, x, weight, bias, running_mean, running_var, momentum, eps, ks, st, pd, comm, pre_mean=None,
                      pre_invstd=None):
        N, C, H, W = x.shape
        x = x.contiguous(memory_format=torch.channels_last)
        lib, s, dt = _native.lib(), _native.stream_ptr(x.device), _native.dtype_code(x.dtype)
        if pre_mean is not None:       # statistics (and running-stat update) came out of the convolution's epilogue
            mean, invstd = pre_mean, pre_invstd
        else:
            mean = torch.empty(C, dtype=torch.float32, device=x.device)
            invstd = torch.empty_like(mean)
            scratch = _bn_scratch(x.device, C)
            _native.check(lib.pdt_bn_fwd_stats(x.data_ptr(), N * H * W, C, dt, scratch.data_ptr(), comm,
                                               _native.ptr(running_mean), _native.ptr(running_var), momentum, eps,
                                               mean.data_ptr(), invstd.data_ptr(), s), "pdt_bn_fwd_stats")
        Ho, Wo = (H + 2 * pd - ks) // st + 1, (W + 2 * pd - ks) // st + 1
        y = torch.empty((N, C, Ho, Wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty((N, Ho, Wo, C), dtype=torch.int8, device=x.device)
        xmax = torch.empty_like(y)       # conv value at each window's arg-max: lets backward reduce over pooled tensors
        _native.check(lib.pdt_bn_pool_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), xmax.data_ptr(), dt, N, H, W, C, ks, st,
                                          pd, mean.data_ptr(), invstd.data_ptr(), weight.data_ptr(), bias.data_ptr(), s),
                      "pdt_bn_pool_fwd")
        ctx.save_for_backward(x, idx, weight, bias, mean, invstd, y, xmax)
        ctx.geom, ctx.comm = (ks, st, pd), comm
        ctx.params = (weight, bias)
        return y

    @staticmethod
    def backward(ctx, dpool):
        with _native.guard(dpool.device):
            return _BNReluPoolNative._backward_impl(ctx, dpool)

    @staticmethod
    def _backward_impl(ctx, dpool):
        x, idx, weight, bias, mean, invstd, y, xmax = ctx.saved_tensors
        ks, st, pd = ctx.geom
        N, C, H, W = x.shape
        dpool = dpool.contiguous(memory_format=torch.channels_last)
        lib, s, dt = _native.lib(), _native.stream_ptr(x.device), _native.dtype_code(x.dtype)
        dgamma, dbeta = _param_grad_slots(ctx.params, C, x.device)
        k = torch.empty(2 * C, dtype=torch.float32, device=x.device)
        scratch = _bn_scratch(x.device, C)
        _native.check(lib.pdt_bn_pool_bwd_reduce(dpool.data_ptr(), y.data_ptr(), xmax.data_ptr(), dt, N, H, W, C, ks, st, pd,
                                                 mean.data_ptr(), invstd.data_ptr(), weight.data_ptr(), bias.data_ptr(),
                                                 scratch.data_ptr(), ctx.comm, dgamma.data_ptr(), dbeta.data_ptr(),
                                                 k.data_ptr(), s), "pdt_bn_pool_bwd_reduce")
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            _native.check(lib.pdt_bn_pool_bwd_apply(dpool.data_ptr(), idx.data_ptr(), x.data_ptr(), dx.data_ptr(), dt, N, H, W,
                                                    C, ks, st, pd, mean.data_ptr(), invstd.data_ptr(), weight.data_ptr(),
                                                    bias.data_ptr(), k.data_ptr(), s), "pdt_bn_pool_bwd_apply")
        return dx, dgamma, dbeta, None, None, None, None, None, None, None, None, None, None


class _SyncStatsTorch(torch.autograd.Function):
    """CPU/oracle SyncBN: all-reduce of raw moments forward, of (sum dy, sum dy*xhat) backward."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps, group):
        C = x.shape[1]
        xf = x.float()
        dims = [0, 2, 3]
        n = torch.tensor([float(x.numel() // C)], device=x.device)
        stats = torch.cat([xf.sum(dims), (xf * xf).sum(dims), n])
        dist.all_reduce(stats, group=group)
        Ntot = stats[-1]
        mean = stats[:C] / Ntot
        var = (stats[C:2 * C] / Ntot - mean * mean).clamp_min(0)
        invstd = torch.rsqrt(var + eps)
        if running_mean is not None:
            running_mean.mul_(1 - momentum).add_(mean, alpha=momentum)
            running_var.mul_(1 - momentum).add_(var * Ntot / (Ntot - 1), alpha=momentum)
        xhat = (xf - mean.view(1, C, 1, 1)) * invstd.view(1, C, 1, 1)
        ctx.save_for_backward(xhat, weight, invstd, Ntot)
        ctx.group = group
        return (xhat * weight.view(1, C, 1, 1) + bias.view(1, C, 1, 1)).to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        xhat, weight, invstd, Ntot = ctx.saved_tensors
        C = xhat.shape[1]
        dyf = dy.float()
        dims = [0, 2, 3]
        sum_dy, sum_dy_xhat = dyf.sum(dims), (dyf * xhat).sum(dims)
        dgamma, dbeta = sum_dy_xhat.clone(), sum_dy.clone()
        red = torch.cat([sum_dy, sum_dy_xhat])
        dist.all_reduce(red, group=ctx.group)
        k1 = (red[:C] / Ntot).view(1, C, 1, 1)
        k2 = (red[C:] / Ntot).view(1, C, 1, 1)
        dx = (dyf - k1 - xhat * k2) * (invstd * weight).view(1, C, 1, 1)
        return dx.to(dy.dtype), dgamma, dbeta, None, None, None, None, None


def bn_act(x, bn, residual=None, relu=False, comm=None, group=None, stats=None, res_link=None):
    """``relu?(bn(x) + residual?)`` for an ``nn.BatchNorm2d``-like module ``bn``.

    ``comm`` (native CommDesc) or ``group`` (process group) turn it into SyncBatchNorm.
    ``stats = (mean, invstd)``: batch statistics already produced (and running statistics already
    updated) by the convolution that made ``x``.
    """
    C = x.shape[1]
    if stats is not None:
        if bn.track_running_stats and bn.num_batches_tracked is not None:
            _bump(bn.num_batches_tracked)
        # (comm: the statistics were already exchanged inside the convolution; backward still has to sum dy moments)
        return _BNActNative.apply(x, residual, bn.weight, bn.bias, None, None, 0.1, bn.eps, relu, comm,
                                  stats[0], stats[1], res_link)
    training = bn.training or bn.running_mean is None
    sync = training and (comm is not None or group is not None)
    momentum = _ema_factor(bn) if training else 0.0
    if training and bn.track_running_stats and bn.num_batches_tracked is not None:
        _bump(bn.num_batches_tracked)
    # (a SyncBatchNorm in train() under no_grad still synchronises its statistics: the native kernels need no autograd)
    if training and _native_ok(x, C) and (torch.is_grad_enabled() or (sync and comm is not None)) and (not sync or comm is not None):
        return _BNActNative.apply(x, residual, bn.weight, bn.bias, bn.running_mean, bn.running_var, momentum, bn.eps,
                                  relu, comm if sync else None, None, None, res_link)
    if not training and _native_ok(x, C) and x.dtype != torch.float32 and not torch.is_grad_enabled():
        xc = x.contiguous(memory_format=torch.channels_last)
        rc = residual.contiguous(memory_format=torch.channels_last) if residual is not None else None
        y = torch.empty_like(xc)
        N, _, H, W = x.shape
        with _native.guard(x.device):
            _native.check(_native.lib().pdt_bn_fwd_apply(
                xc.data_ptr(), _native.ptr(rc), y.data_ptr(), N * H * W, C, _native.dtype_code(x.dtype),
                bn.running_mean.data_ptr(), bn.running_var.data_ptr(), bn.weight.data_ptr(), bn.bias.data_ptr(),
                int(relu), 1, bn.eps, _native.stream_ptr(x.device)), "pdt_bn_fwd_apply(eval)")
        return y
    if sync and group is not None and dist.is_initialized() and dist.get_world_size(group) > 1:
        y = _SyncStatsTorch.apply(x, bn.weight, bn.bias, bn.running_mean, bn.running_var, momentum, bn.eps, group)
    else:
        y = F.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, training, momentum, bn.eps)
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y


def _pool_fusable(x, ks, st, pd):
    """Shapes the fused BN + ReLU + max-pool kernels serve: any window in 16 bit; fp32 has the 3x3 / 2 / 1 form (the
    ResNet stem) over even maps only.  PDT_BN_POOL_F32=0 sends fp32 through the separate kernels."""
    if x.dtype != torch.float32:
        return True
    return ((ks, st, pd) == (3, 2, 1) and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0
            and os.environ.get("PDT_BN_POOL_F32", "1") == "1")


def bn_relu_maxpool(x, bn, kernel_size=3, stride=2, padding=1, comm=None, group=None):
    """``max_pool2d(relu(bn(x)))``; one fused kernel set on the native training path."""
    C = x.shape[1]
    training = bn.training or bn.running_mean is None
    sync = training and (comm is not None or group is not None)
    if (training and torch.is_grad_enabled() and _native_ok(x, C) and _pool_fusable(x, kernel_size, stride, padding)
            and (not sync or comm is not None) and x.numel() // 4 < 4_000_000_000):
        momentum = _ema_factor(bn)
        if bn.track_running_stats and bn.num_batches_tracked is not None:
            _bump(bn.num_batches_tracked)
        return _BNReluPoolNative.apply(x, bn.weight, bn.bias, bn.running_mean, bn.running_var, momentum, bn.eps,
                                       kernel_size, stride, padding, comm if sync else None)
    from .pool import max_pool2d
    return max_pool2d(bn_act(x, bn, None, True, comm=comm, group=group), kernel_size, stride, padding)


def conv_bn_relu_maxpool(x, conv, bn, kernel_size=3, stride=2, padding=1, comm=None):
    """``max_pool2d(relu(bn(conv(x))))`` (the ResNet stem).  On the native training path the convolution's
    epilogue produces the BatchNorm statistics, so the 112x112x64 conv output is read once (by the fused
    BN+ReLU+pool kernel) instead of twice."""
    from .conv import conv2d
    training = bn.training or bn.running_mean is None
    fuse = (training and torch.is_grad_enabled() and x.is_cuda and _native.available() and bn.track_running_stats
            and bn.affine and os.environ.get("PDT_BNSTATS_FUSION", "1") == "1")
    if fuse:
        momentum = _ema_factor(bn)
        out = conv2d(x, conv, bnpack=(bn.running_mean, bn.running_var, momentum, bn.eps, comm))
        if isinstance(out, tuple):
            y, mean, invstd = out
            C = y.shape[1]
            if _native_ok(y, C) and _pool_fusable(y, kernel_size, stride, padding) and y.numel() // 4 < 4_000_000_000:
                if bn.num_batches_tracked is not None:
                    _bump(bn.num_batches_tracked)
                return _BNReluPoolNative.apply(y, bn.weight, bn.bias, None, None, momentum, bn.eps, kernel_size, stride,
                                               padding, comm, mean, invstd)
            if comm is None:
                raise RuntimeError("conv epilogue produced statistics for a shape the fused pool kernels do not serve")
            from .pool import max_pool2d        # fp32 SyncBatchNorm stem: statistics are done, normalise + pool separately
            if bn.num_batches_tracked is not None:
                _bump(bn.num_batches_tracked)
            return max_pool2d(_BNActNative.apply(y, None, bn.weight, bn.bias, None, None, momentum, bn.eps, True, comm,
                                                 mean, invstd, None), kernel_size, stride, padding)
        return bn_relu_maxpool(out, bn, kernel_size, stride, padding, comm=comm)
    return bn_relu_maxpool(conv2d(x, conv), bn, kernel_size, stride, padding, comm=comm)


def conv_bn_act(x, conv, bn, residual=None, relu=False, comm=None, group=None, consume=None, produce=None, res_link=None):
    """``relu?(bn(conv(x)) + residual?)``.  On the native training path the convolution's epilogue
    produces the BatchNorm statistics (no separate pass over the activation); SyncBatchNorm keeps its
    own statistics kernel because the cross-rank exchange lives there."""
    from .conv import conv2d
    training = bn.training or bn.running_mean is None
    # (comm: SyncBatchNorm on the native fabric — the exchange of the moments happens in the convolution's last CTA)
    fuse = (training and group is None and torch.is_grad_enabled() and x.is_cuda
            and _native.available() and bn.track_running_stats and bn.affine)
    if fuse:
        momentum = _ema_factor(bn)
        out = conv2d(x, conv, bnpack=(bn.running_mean, bn.running_var, momentum, bn.eps, comm), consume=consume, produce=produce)
        if isinstance(out, tuple):
            y, mean, invstd = out
            if _native_ok(y, y.shape[1]):
                return bn_act(y, bn, residual, relu, comm=comm, stats=(mean, invstd), res_link=res_link)
            raise RuntimeError("conv epilogue produced statistics for a shape the BN kernels do not serve")
        return bn_act(out, bn, residual, relu, comm=comm, group=group, res_link=res_link)
    return bn_act(conv2d(x, conv, consume=consume, produce=produce), bn, residual, relu, comm=comm, group=group, res_link=res_link)
