"""Command-line surface of the three entrypoints.

Flag names, dests, types and defaults are those of
/root/reference/dataparallel.py:40-67, distributed.py:43-73 and
distributed_syncBN_amp.py:42-75 — including the ``type=bool`` quirk (any
non-empty string is True) because scripts in the wild depend on it.  Knobs the
reference does not have are environment variables (``PDT_*``, see
INTEGRATION.md) so the CLI stays byte-compatible.
"""
import argparse

_DEFAULT_OUT = {"dp": "./output", "ddp": "./output_ddp_test", "ddp_amp": "./output_ddp_amp"}
_DEFAULT_GPUS = {"dp": "5,6,7", "ddp": "0,1,2", "ddp_amp": "0,1,2"}
LOGGER_NAME = {"dp": "DataParallel", "ddp": "DistributedDataParallel",
               "ddp_amp": "DistributedDataParallel_amp"}


def model_names():
    """Every lowercase callable in torchvision.models (distributed.py:39-40)."""
    import torchvision.models as models
    return sorted(n for n in models.__dict__
                  if n.islower() and not n.startswith("__") and callable(models.__dict__[n]))


def build_parser(mode):
    assert mode in _DEFAULT_OUT, mode
    names = model_names()
    p = argparse.ArgumentParser(description="PyTorch ImageNet Training")
    p.add_argument("--data", metavar="DIR", default="/mnt/cephfs/mixed/dataset/imagenet/", help="path to dataset")
    p.add_argument("-a", "--arch", metavar="ARCH", default="resnet18", choices=names,
                   help="model architecture: " + " | ".join(names) + "(default: resnet18)")
    p.add_argument("-j", "--workers", default=8, type=int, metavar="N",
                   help="number of data loading workers (default: 4)")
    p.add_argument("--epochs", default=5, type=int, metavar="N", help="number of total epochs to run")
    p.add_argument("--step", default=[3, 4], metavar="step decay", help="lr decay by step")
    p.add_argument("--start-epoch", default=0, type=int, metavar="N", help="manual epoch number ()")
    p.add_argument("-b", "--batch-size", default=1200, type=int, metavar="N",
                   help="mini-batch size (default: 3200), this is the total batch size of all GPUs on the "
                        "current node when using Data Parallel or Distributed Data Parallel")
    p.add_argument("--lr", "--learning-rate", default=0.1, type=float, metavar="LR",
                   help="initial learning rate", dest="lr")
    p.add_argument("--momentum", default=0.9, type=float, metavar="M", help="momentum")
    p.add_argument("--wd", "--weight-decay", default=1e-4, type=float, metavar="W",
                   help="weight decay (default: 1e-4)", dest="weight_decay")
    p.add_argument("-p", "--print-freq", default=10, type=int, metavar="N", help="print frequency (default: 10)")
    p.add_argument("-e", "--evaluate", dest="evaluate", default=False, type=bool,
                   help="evaluate model on validation set")
    p.add_argument("--pretrained", dest="pretrained", default=False, type=bool, help="use pre-trained model")
    p.add_argument("--seed", default=None, type=int, help="seed for initializing training")
    p.add_argument("--gpus", default=_DEFAULT_GPUS[mode], metavar="gpus_id", help="N gpus for training")
    p.add_argument("--outpath", metavar="DIR", default=_DEFAULT_OUT[mode], help="path to output")
    p.add_argument("--lr-scheduler", metavar="LR scheduler", default="steplr", help="LR scheduler",
                   dest="lr_scheduler")
    p.add_argument("--gamma", default=0.1, type=float, metavar="gamma", help="gamma")
    if mode in ("ddp", "ddp_amp"):
        # torch >= 2.0 launchers pass the dashed spelling; accept both.
        p.add_argument("--local_rank", "--local-rank", default=0, type=int, dest="local_rank",
                       help="node rank for distributed training")
    if mode == "ddp_amp":
        p.add_argument("--use_amp", dest="use_amp", default=True, type=bool,
                       help="use automatic mixed precision (amp)")
        p.add_argument("--sync_batchnorm", dest="sync_batchnorm", default=False, type=bool,
                       help="use sync batchnorm")
    return p


def normalise_steps(step):
    """``--step`` has no type (distributed.py:52): a CLI value arrives as a str."""
    if isinstance(step, str):
        return [int(s) for s in step.replace("[", "").replace("]", "").replace(",", " ").split()]
    return [int(s) for s in step]
