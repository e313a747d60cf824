"""Logger, settings dump, output-directory handling, TensorBoard writer.

Parity targets: /root/reference/utils.py:17-37 (get_logger), :40-51
(output_process), :54-62 (write_settings), :72-74 (ddp_print).
"""
import logging
import os
import shutil
import sys

_FILE_FMT = "%(asctime)s %(levelname)s: %(message)s"
_CONSOLE_FMT = "%(message)s"


def get_logger(save_path, logger_name):
    """INFO logger writing ``<save_path>/experiment.log`` and stdout.

    File lines carry time + level, console lines the bare message, exactly as
    /root/reference/utils.py:22-36.  Unlike the reference, handlers of a
    previous call with the same name are dropped first, so a second run in
    one interpreter (tests, notebooks) does not duplicate every line.
    """
    logger = logging.getLogger(logger_name)
    for h in list(logger.handlers):
        logger.removeHandler(h)
        h.close()
    fh = logging.FileHandler(os.path.join(save_path, "experiment.log"))
    fh.setFormatter(logging.Formatter(_FILE_FMT))
    ch = logging.StreamHandler(sys.stdout)
    ch.setFormatter(logging.Formatter(_CONSOLE_FMT))
    logger.addHandler(fh)
    logger.addHandler(ch)
    logger.setLevel(logging.INFO)
    logger.propagate = False
    return logger


def output_process(output_path):
    """Create the run directory; ask before deleting an existing one.

    /root/reference/utils.py:40-51 always calls ``input()``, which hangs or
    raises EOFError under torchrun.  Here the prompt is kept only
    on an interactive TTY.  Otherwise ``PDT_OUTPUT_EXISTS`` decides:
    ``delete`` (default) removes the directory, ``quit`` raises the same
    OSError the reference raises, ``keep`` reuses it.
    """
    if os.path.exists(output_path):
        print("{} file exist!".format(output_path))
        if sys.stdin is not None and sys.stdin.isatty() and "PDT_OUTPUT_EXISTS" not in os.environ:
            act = input("Select Action: d (delete) / q (quit):").lower().strip()
        else:
            act = {"delete": "d", "quit": "q", "keep": "k"}.get(
                os.environ.get("PDT_OUTPUT_EXISTS", "delete").lower(), "q")
        if act == "d":
            shutil.rmtree(output_path)
        elif act == "k":
            pass
        else:
            raise OSError("Directory {} exits!".format(output_path))
    os.makedirs(output_path, exist_ok=True)


def write_settings(settings):
    """``key: value`` per line into ``<outpath>/settings.log`` (utils.py:54-62)."""
    with open(os.path.join(settings.outpath, "settings.log"), "w") as f:
        for k, v in vars(settings).items():
            f.write("{}: {}\n".format(k, v))


def ddp_print(output, logger, local_rank):
    """Rank-0-only ``logger.info`` (utils.py:72-74); logger is None elsewhere."""
    if local_rank == 0 and logger is not None:
        logger.info(output)


class _NullWriter:
    """Stand-in when no TensorBoard writer can be imported."""

    def add_scalar(self, *a, **k):
        pass

    def close(self):
        pass


def make_summary_writer(logdir):
    """tensorboardX if present (what the reference imports, distributed.py:33),
    else torch.utils.tensorboard, else a no-op writer."""
    try:
        from tensorboardX import SummaryWriter  # type: ignore
        return SummaryWriter(logdir)
    except Exception:
        pass
    try:
        from torch.utils.tensorboard import SummaryWriter
        return SummaryWriter(logdir)
    except Exception:
        return _NullWriter()
