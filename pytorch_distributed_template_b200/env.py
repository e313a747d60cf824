"""Rank / world-size / device resolution.

The reference reads ``--local_rank`` only (distributed.py:68) and sizes the
job from ``torch.cuda.device_count()`` (distributed.py:123); both break under
torchrun on torch >= 2.0.  Here every spelling works:
``--local_rank``, ``--local-rank`` and the ``LOCAL_RANK`` environment variable,
and the world size is whatever the process group says.
"""
import os

import torch
import torch.distributed as dist


def env_int(name, default):
    v = os.environ.get(name)
    return default if v in (None, "") else int(v)


def env_flag(name, default=False):
    v = os.environ.get(name)
    if v is None:
        return default
    return v.strip().lower() not in ("", "0", "false", "no", "off")


def resolve_local_rank(args=None):
    if "LOCAL_RANK" in os.environ:
        return int(os.environ["LOCAL_RANK"])
    if args is not None and getattr(args, "local_rank", None) is not None:
        return int(args.local_rank)
    return 0


def use_cuda():
    if os.environ.get("PDT_DEVICE", "").lower() == "cpu":
        return False
    return torch.cuda.is_available()


def pick_device(local_rank):
    if use_cuda():
        torch.cuda.set_device(local_rank)
        return torch.device("cuda", local_rank)
    return torch.device("cpu")


def init_distributed(local_rank):
    """env:// rendezvous (distributed.py:124).  nccl on GPUs, gloo on CPU.

    A bare ``python distributed.py`` (no launcher) becomes a world of one.
    """
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "23334")  # start.sh:3 uses this port
    os.environ.setdefault("RANK", str(local_rank))
    os.environ.setdefault("WORLD_SIZE", "1")
    device = pick_device(local_rank)
    if not dist.is_initialized():
        if device.type == "cuda":
            dist.init_process_group(backend="nccl", device_id=device)
        else:
            dist.init_process_group(backend="gloo")
    return device


def shutdown_distributed():
    """The reference never tears the group down."""
    if dist.is_available() and dist.is_initialized():
        try:
            dist.barrier()
        except Exception:
            pass
        dist.destroy_process_group()


def world_size():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def rank():
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
