"""pytorch_distributed_template_b200 — a B200-native data-parallel training framework.

Capabilities of xiezheng-cs/PyTorch_Distributed_Template (DataParallel, DDP,
DDP+amp, DDP+amp+SyncBN ResNet training; /root/reference/*.py), re-designed
for Blackwell: PyTorch is the framework layer (tensors, autograd, data
loading); the hot paths are hand-written sm_100a kernels in ``csrc/`` reached
through a C ABI (``include/pdt_b200.h``).

Sub-packages
  utils/     logging, meters, checkpoint IO         (reference utils.py)
  models/    state_dict-compatible fused ResNets    (torchvision resnet*)
  ops/       autograd.Functions over the kernels + torch oracles
  parallel/  fabric, DistributedDataParallel, DataParallel, SyncBatchNorm
"""
__version__ = "0.1.0"

from . import utils  # noqa: F401
