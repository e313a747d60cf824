"""ResNet family with torchvision's parameter names and initialisation, built from fused ops.

``state_dict()`` keys, shapes and the random init match
``torchvision.models.resnet{18,34,50,101,152}`` (what the reference trains,
/root/reference/distributed.py:134-137), so checkpoints interchange in both
directions.  The forward differs: activations are channels_last, every
conv -> BN -> (+identity) -> ReLU chain is ``ops.conv2d`` followed by ONE fused
``ops.bn_act`` (statistics + normalise + residual + ReLU), pooling and the
classifier use the package's kernels, and under ``amp.autocast`` activations
are born in bf16/fp16 from shadow weights.
"""
import torch
import torch.nn as nn

from .. import amp, ops
from ..ops import bn as _bn_ops
from ..ops.conv_native import GradLink, link_enabled
from ..parallel.syncbn import fused_bn, fused_bn_relu_maxpool, fused_conv_bn, fused_conv_bn_relu_maxpool


def conv3x3(cin, cout, stride=1):
    return nn.Conv2d(cin, cout, 3, stride, 1, bias=False)


def conv1x1(cin, cout, stride=1):
    return nn.Conv2d(cin, cout, 1, stride, 0, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = conv3x3(inplanes, planes, stride)
        self.bn1 = nn.BatchNorm2d(planes)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = nn.BatchNorm2d(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x
        # The gradient of x is conv1's data gradient plus the shortcut's gradient: the link lets conv1's backward
        # kernel add the latter in its epilogue (ops.conv_native.GradLink) instead of a separate elementwise pass.
        link = GradLink() if link_enabled(x) else None
        out = fused_conv_bn(self.conv1, self.bn1, x, relu=True, consume=link)
        if link is not None and not link.claimed:
            link = None
        if self.downsample is not None:
            identity = fused_conv_bn(self.downsample[0], self.downsample[1], x, produce=link)
            link = None
        return fused_conv_bn(self.conv2, self.bn2, out, residual=identity, relu=True, res_link=link)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = conv1x1(inplanes, planes)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = conv3x3(planes, planes, stride)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = conv1x1(planes, planes * self.expansion)
        self.bn3 = nn.BatchNorm2d(planes * self.expansion)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x
        link = GradLink() if link_enabled(x) else None       # see BasicBlock.forward
        out = fused_conv_bn(self.conv1, self.bn1, x, relu=True, consume=link)
        if link is not None and not link.claimed:
            link = None
        out = fused_conv_bn(self.conv2, self.bn2, out, relu=True)
        if self.downsample is not None:
            identity = fused_conv_bn(self.downsample[0], self.downsample[1], x, produce=link)
            link = None
        return fused_conv_bn(self.conv3, self.bn3, out, residual=identity, relu=True, res_link=link)


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes=1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():   # torchvision/models/resnet.py init
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(conv1x1(self.inplanes, planes * block.expansion, stride),
                                       nn.BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        layers += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    accepts_uint8 = True      # raw uint8 NCHW images are normalised on the device (ops/input_norm.py)

    def forward(self, x):
        # Weights live KRSC in the arena; keep activations channels_last to match.  (The native
        # amp stem consumes fp32 NCHW images directly and does the layout change itself.)
        if x.dim() == 4 and not (x.is_cuda and amp.active_dtype() is not None) and x.dtype != torch.uint8 \
                and not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        if self.training and torch.is_grad_enabled() and x.is_cuda:
            ops.prepare_backward(self.conv1.weight)     # data-gradient filter layouts of every conv, one launch
        with _bn_ops.batched_counters():       # one launch for all num_batches_tracked increments
            x = fused_conv_bn_relu_maxpool(self.conv1, self.bn1, x, 3, 2, 1)
            x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = ops.global_avg_pool(x)
        return ops.linear(x, self.fc)


_CFG = {"resnet18": (BasicBlock, [2, 2, 2, 2]), "resnet34": (BasicBlock, [3, 4, 6, 3]),
        "resnet50": (Bottleneck, [3, 4, 6, 3]), "resnet101": (Bottleneck, [3, 4, 23, 3]),
        "resnet152": (Bottleneck, [3, 8, 36, 3])}


def native_archs():
    return sorted(_CFG)


def build_resnet(arch, num_classes=1000):
    block, layers = _CFG[arch]
    return ResNet(block, layers, num_classes)
