"""Parallel wrappers: fabric, DistributedDataParallel, DataParallel, SyncBatchNorm."""
from .fabric import Fabric
from .arena import ParamArena
from .ddp import DistributedDataParallel
from .syncbn import SyncBatchNorm
from .dp import DataParallel

__all__ = ["Fabric", "ParamArena", "DistributedDataParallel", "SyncBatchNorm", "DataParallel"]
