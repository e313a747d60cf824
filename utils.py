#!/usr/bin/env python
"""Drop-in for the reference's ``utils.py`` (/root/reference/utils.py:1-118).

Same eight public names; the implementation lives in the package so the three
entrypoints and library users share one copy.
"""
from pytorch_distributed_template_b200.utils import (  # noqa: F401
    get_logger, output_process, write_settings, get_learning_rate, ddp_print,
    AverageMeter, accuracy, save_checkpoint, load_checkpoint, DeviceMeter,
)
