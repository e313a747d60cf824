"""Checkpoint save/load with the reference's dict layout.

Parity: /root/reference/utils.py:114-118 and the dict built at
/root/reference/distributed.py:212-218:
``{'epoch', 'arch', 'state_dict' (no 'module.' prefix), 'best_acc1'}``.
The reference has no load path; ``load_checkpoint`` is the
missing half and accepts files written by either code base, with or without
the ``module.`` prefix.
"""
import os
import shutil

import torch


def save_checkpoint(state, is_best, outpath):
    filename = os.path.join(outpath, "checkpoint.pth.tar")
    torch.save(state, filename)
    if is_best:
        shutil.copyfile(filename, os.path.join(outpath, "model_best.pth.tar"))


def strip_module_prefix(state_dict):
    return {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state_dict.items()}


def add_module_prefix(state_dict):
    return {(k if k.startswith("module.") else "module." + k): v for k, v in state_dict.items()}


def load_checkpoint(path, model=None, map_location="cpu"):
    """Load a checkpoint dict; if ``model`` is given, restore its weights.

    ``model`` may be a bare network or one of our wrappers (which expose the
    wrapped net as ``.module`` like torch's, distributed.py:215).
    """
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    if model is not None:
        target = model.module if hasattr(model, "module") else model
        target.load_state_dict(strip_module_prefix(ckpt["state_dict"]))
    return ckpt
