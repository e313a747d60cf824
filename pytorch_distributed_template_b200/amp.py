"""Mixed precision: autocast policy and a device-resident GradScaler.

Reference behaviour (distributed_syncBN_amp.py:196,259-261,275-278;
torch/amp/grad_scaler.py): scale 65536, x2 every 2000 clean steps, x0.5 and a
skipped step when any gradient is non-finite.  The reference's
``scaler.step`` synchronises the host (``found_inf.item()``); here the whole
decision lives in 8 words of device memory that the gradient kernels read and
write (``csrc/comm.cu``), so a step never waits for the CPU.

``autocast`` wraps ``torch.autocast`` (for layers that still run on torch ops)
and additionally publishes the active dtype for the native layers, which read
the bf16/fp16 *shadow weights* the optimizer maintains instead of casting the
fp32 weights every forward.
"""
import contextlib
import os
import threading

import torch
import torch.distributed as dist

from .ops import _native

_tls = threading.local()


def default_dtype():
    """fp16 is torch's CUDA autocast default and what the reference runs; PDT_AMP_DTYPE=bf16 switches."""
    return {"bf16": torch.bfloat16, "bfloat16": torch.bfloat16, "fp16": torch.float16,
            "float16": torch.float16}[os.environ.get("PDT_AMP_DTYPE", "fp16").lower()]


def active_dtype():
    """Low-precision dtype of the enclosing ``autocast`` region, or None."""
    return getattr(_tls, "dtype", None)


@contextlib.contextmanager
def autocast(enabled=True, dtype=None, device_type="cuda"):
    prev = active_dtype()
    dtype = dtype or default_dtype()
    if device_type == "cpu" and dtype == torch.float16:
        dtype = torch.bfloat16
    _tls.dtype = dtype if enabled else None
    try:
        with torch.autocast(device_type=device_type, dtype=dtype, enabled=enabled):
            yield
    finally:
        _tls.dtype = prev


class GradScaler:
    """Drop-in for ``torch.cuda.amp.GradScaler`` as the reference uses it.

    ``bind(optimizer)`` (or the ``optimizer=`` argument) connects the scaler to
    our fused SGD so that unscale / inf-check / skip happen inside the
    gradient kernels.  With a foreign optimizer it falls back to torch ops.
    """

    def __init__(self, init_scale=65536.0, growth_factor=2.0, backoff_factor=0.5, growth_interval=2000,
                 enabled=True, device=None, optimizer=None):
        self._enabled = bool(enabled)
        self.growth_factor, self.backoff_factor = float(growth_factor), float(backoff_factor)
        self.growth_interval = int(growth_interval)
        self._init_scale = float(init_scale)
        self.state = None          # float32[8] device tensor, layout in include/pdt_b200.h
        self._optimizer = None
        self._device = device
        if optimizer is not None:
            self.bind(optimizer)

    def is_enabled(self):
        return self._enabled

    def _ensure_state(self, device):
        if self.state is None:
            st = torch.zeros(8, dtype=torch.float32, device=device)
            st[0] = self._init_scale if self._enabled else 1.0
            self.state = st
        return self.state

    def bind(self, optimizer):
        self._optimizer = optimizer
        arena = getattr(optimizer, "arena", None)
        device = arena.device if arena is not None else (
            self._device or optimizer.param_groups[0]["params"][0].device)
        self._ensure_state(device)
        if hasattr(optimizer, "attach_scaler"):
            optimizer.attach_scaler(self)
        return self

    def scale(self, loss):
        if not self._enabled:
            return loss
        st = self._ensure_state(loss.device)
        return loss * st[0]

    def step(self, optimizer, *args, **kwargs):
        if not self._enabled:
            return optimizer.step(*args, **kwargs)
        if hasattr(optimizer, "attach_scaler"):
            if optimizer.scaler is not self:
                raise RuntimeError("GradScaler.bind(optimizer) must be called before the first backward: the fused "
                                   "gradient kernels read the loss scale while backward is still running")
            return optimizer.step(*args, **kwargs)   # unscale, inf check and skip happen in-kernel
        # foreign optimizer: torch ops, one host sync like the reference
        st = self.state
        inv = 1.0 / st[0]
        grads = [p.grad for g in optimizer.param_groups for p in g["params"] if p.grad is not None]
        torch._foreach_mul_(grads, inv)
        bad = torch.stack([(~torch.isfinite(g)).any() for g in grads]).any().float()
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(bad, op=dist.ReduceOp.MAX)
        st.view(torch.int32)[4] = int(bad.item() > 0)
        if bad.item() > 0:
            return None
        return optimizer.step(*args, **kwargs)

    def update(self, new_scale=None):
        if not self._enabled or self.state is None:
            return
        st = self.state
        if new_scale is not None:
            st[0] = float(new_scale)
            return
        if st.is_cuda and _native.available():
            _native.check(_native.lib().pdt_scale_update(st.data_ptr(), self.growth_factor, self.backoff_factor,
                                                         self.growth_interval, _native.stream_ptr(st.device)),
                          "pdt_scale_update")
            return
        si = st.view(torch.int32)
        if int(si[4]) != 0:
            st[0] *= self.backoff_factor
            si[1] = 0
            si[3] += 1
        else:
            si[1] += 1
            if int(si[1]) == self.growth_interval:
                st[0] *= self.growth_factor
                si[1] = 0
        si[2] = 0
        si[4] = 0

    def get_scale(self):
        return float(self.state[0]) if (self._enabled and self.state is not None) else 1.0

    def skipped_steps(self):
        return int(self.state.view(torch.int32)[3]) if self.state is not None else 0

    def state_dict(self):
        if not self._enabled or self.state is None:
            return {}
        si = self.state.view(torch.int32)
        return {"scale": float(self.state[0]), "growth_factor": self.growth_factor,
                "backoff_factor": self.backoff_factor, "growth_interval": self.growth_interval,
                "_growth_tracker": int(si[1])}

    def load_state_dict(self, sd):
        if not sd:
            return
        self.growth_factor, self.backoff_factor = sd["growth_factor"], sd["backoff_factor"]
        self.growth_interval = sd["growth_interval"]
        if self.state is not None:
            self.state[0] = sd["scale"]
            self.state.view(torch.int32)[1] = sd["_growth_tracker"]
        else:
            self._init_scale = sd["scale"]
