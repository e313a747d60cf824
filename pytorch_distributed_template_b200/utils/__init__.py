"""Observability / IO helpers (SURVEY §2.1 C4–C11, layer L8).

Same public names and on-disk formats as the reference's ``utils.py``
(/root/reference/utils.py:17-118) so logs, ``settings.log`` and
``checkpoint.pth.tar`` interchange with it.  Behavioural deviations are
deliberate and listed next to each function.
"""
from .logging import get_logger, ddp_print, write_settings, output_process, make_summary_writer
from .meters import AverageMeter, DeviceMeter, accuracy, get_learning_rate, reduce_mean
from .checkpoint import save_checkpoint, load_checkpoint, strip_module_prefix, add_module_prefix

__all__ = [
    "get_logger", "ddp_print", "write_settings", "output_process", "make_summary_writer",
    "AverageMeter", "DeviceMeter", "accuracy", "get_learning_rate", "reduce_mean",
    "save_checkpoint", "load_checkpoint", "strip_module_prefix", "add_module_prefix",
]
