"""SyncBatchNorm: cross-replica batch statistics.

``SyncBatchNorm.convert_sync_batchnorm(model)`` has the contract of
``nn.SyncBatchNorm.convert_sync_batchnorm`` (distributed_syncBN_amp.py:143-145):
every ``BatchNorm*d`` becomes a SyncBatchNorm with the same parameters, buffers
and state_dict keys.  In training the statistics kernels exchange raw moments
with all peers over NVLink from inside the kernel (``csrc/bn.cu``) — the
reference's 20 all_gathers + 20 all_reduces + 20 host syncs per step are gone.
In eval it is plain BatchNorm on running statistics, like torch.
"""
import os

import torch
import torch.distributed as dist
import torch.nn as nn

from .. import ops


class SyncBatchNorm(nn.BatchNorm2d):
    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True,
                 process_group=None):
        super().__init__(num_features, eps, momentum, affine, track_running_stats)
        self.process_group = process_group
        self._pdt_fabric = None   # set by DistributedDataParallel

    def _check_input_dim(self, input):
        if input.dim() < 2:
            raise ValueError("expected at least 2D input (got {}D input)".format(input.dim()))

    def forward(self, x):
        return fused_bn(self, x)

    @classmethod
    def convert_sync_batchnorm(cls, module, process_group=None):
        out = module
        if isinstance(module, nn.modules.batchnorm._BatchNorm) and not isinstance(module, SyncBatchNorm):
            out = SyncBatchNorm(module.num_features, module.eps, module.momentum, module.affine,
                                module.track_running_stats, process_group)
            if module.affine:
                with torch.no_grad():
                    out.weight = module.weight
                    out.bias = module.bias
            out.running_mean = module.running_mean
            out.running_var = module.running_var
            out.num_batches_tracked = module.num_batches_tracked
            out.training = module.training
            if hasattr(module, "qconfig"):
                out.qconfig = module.qconfig
        for name, child in module.named_children():
            out.add_module(name, cls.convert_sync_batchnorm(child, process_group))
        del module
        return out


def fused_bn(bn, x, residual=None, relu=False, res_link=None):
    """Dispatch one BatchNorm-like module through the fused op (SyncBN aware)."""
    if isinstance(bn, SyncBatchNorm) and bn.training:
        fabric = bn._pdt_fabric
        if fabric is not None and fabric.native and fabric.world > 1 and x.is_cuda:
            return ops.bn_act(x, bn, residual, relu, comm=fabric.comm, res_link=res_link)
        group = bn.process_group if bn.process_group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        if group is not None and dist.get_world_size(group) > 1:
            return ops.bn_act(x, bn, residual, relu, group=group)
    if x.dim() != 4:
        y = nn.functional.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, bn.training,
                                     0.1 if bn.momentum is None else bn.momentum, bn.eps)
        if residual is not None:
            y = y + residual
        return torch.relu(y) if relu else y
    return ops.bn_act(x, bn, residual, relu, res_link=res_link)


def fused_bn_relu_maxpool(bn, x, kernel_size=3, stride=2, padding=1):
    """Stem tail ``maxpool(relu(bn(x)))`` through the fused op (SyncBN aware)."""
    if isinstance(bn, SyncBatchNorm) and bn.training:
        fabric = bn._pdt_fabric
        if fabric is not None and fabric.native and fabric.world > 1 and x.is_cuda:
            return ops.bn_relu_maxpool(x, bn, kernel_size, stride, padding, comm=fabric.comm)
        group = bn.process_group if bn.process_group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        if group is not None and dist.get_world_size(group) > 1:
            return ops.bn_relu_maxpool(x, bn, kernel_size, stride, padding, group=group)
    return ops.bn_relu_maxpool(x, bn, kernel_size, stride, padding)


def _native_comm(bn, x):
    """The fabric descriptor for a training SyncBatchNorm on the native NVLink path, else None."""
    if isinstance(bn, SyncBatchNorm) and bn.training:
        fabric = bn._pdt_fabric
        if fabric is not None and fabric.native and fabric.world > 1 and x.is_cuda:
            return fabric.comm
    return None


def fused_conv_bn_relu_maxpool(conv, bn, x, kernel_size=3, stride=2, padding=1):
    """Stem ``maxpool(relu(bn(conv(x))))`` (SyncBN aware).  The batch statistics come out of the stem convolution's
    epilogue; for a SyncBatchNorm on the native fabric the kernel's last CTA also exchanges them with the peers.
    (A SyncBatchNorm over a torch process group keeps its own statistics pass.)"""
    if isinstance(bn, SyncBatchNorm) and bn.training:
        comm = _native_comm(bn, x)
        if comm is None or os.environ.get("PDT_SYNCBN_EPILOGUE", "1") != "1":
            return fused_bn_relu_maxpool(bn, ops.conv2d(x, conv), kernel_size, stride, padding)
        return ops.conv_bn_relu_maxpool(x, conv, bn, kernel_size, stride, padding, comm=comm)
    return ops.conv_bn_relu_maxpool(x, conv, bn, kernel_size, stride, padding)


def fused_conv_bn(conv, bn, x, residual=None, relu=False, consume=None, produce=None, res_link=None):
    """``relu?(bn(conv(x)) + residual?)`` (SyncBN aware): plain BatchNorm gets its statistics from the
    convolution's epilogue, a SyncBatchNorm on the native fabric too — the convolution's last CTA exchanges the
    moments with the peers (PDT_SYNCBN_EPILOGUE=0: its own exchanging statistics kernel).
    consume / produce / res_link: the block's shortcut-gradient link (ops.conv_native.GradLink)."""
    if isinstance(bn, SyncBatchNorm) and bn.training:
        comm = _native_comm(bn, x)
        if comm is None or os.environ.get("PDT_SYNCBN_EPILOGUE", "1") != "1":
            return fused_bn(bn, ops.conv2d(x, conv, consume=consume, produce=produce), residual, relu, res_link=res_link)
        return ops.conv_bn_act(x, conv, bn, residual, relu, comm=comm, consume=consume, produce=produce, res_link=res_link)
    return ops.conv_bn_act(x, conv, bn, residual, relu, consume=consume, produce=produce, res_link=res_link)
