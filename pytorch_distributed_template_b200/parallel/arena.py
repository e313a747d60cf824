"""Flat parameter / gradient arenas.

All parameters of a model live back to back in one fp32 arena, all gradients in
a second arena with the same offsets, momentum in a third and (for AMP) a
bf16/fp16 shadow of the weights in a fourth.  Parameters are laid out in
REVERSE registration order because that is (to first order) the order in which
backward produces gradients, so "the gradients that are ready first" is always
a prefix and a communication segment is a contiguous range.

This is the counterpart of c10d Reducer buckets (reducer.hpp:30-31: first
bucket 1 MiB, then 25 MiB) with ``gradient_as_bucket_view=True``: ``p.grad`` is
a view into the arena, so the collective needs no staging copy.

Conv weights are stored KRSC (channels_last of the logical OIHW tensor): that
is what the implicit-GEMM kernels and TMA want, and ``state_dict`` IO is
layout-agnostic.
"""
import os

import torch

from .fabric import Fabric

_ALIGN = 64   # elements: 256 B in fp32, 128 B in the 16-bit shadow


def _round_up(x, m):
    return (x + m - 1) // m * m


class Segment:
    __slots__ = ("index", "start", "end", "params", "pending")

    def __init__(self, index, start):
        self.index, self.start, self.end, self.params, self.pending = index, start, start, [], 0

    @property
    def length(self):
        return self.end - self.start


_DIRECT = os.environ.get("PDT_DIRECT_GRAD", "1") == "1"     # 0: backward kernels never write into the arena themselves


class ParamArena:
    def __init__(self, params, fabric, first_bucket_mb=1.0, bucket_cap_mb=25.0, shadow_dtype=None, shard_world=None):
        """``shard_world``: how many ways a segment is split between reducing ranks / devices (defaults to the
        fabric's world; DataParallel passes its device count).  Any value 1..8 works: segments are padded to a
        multiple of lcm(4 * shard_world, 64) elements, so every shard is a whole number of 16-byte vectors."""
        self.fabric = fabric
        self.device = fabric.device
        self.shadow_dtype = shadow_dtype
        params = [p for p in params if p.requires_grad]
        seen, uniq = set(), []
        for p in params:
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append(p)
        self.params = uniq
        order = list(reversed(uniq))
        import math
        world = int(shard_world or fabric.world)
        self.shard_world = world
        seg_align = 4 * world * _ALIGN // math.gcd(4 * world, _ALIGN)

        # ---- layout: offsets (in elements) and segments --------------------------------
        self.offset = {}
        self.segments = []
        cap = int(first_bucket_mb * 1024 * 1024) // 4
        cur = Segment(0, 0)
        pos = 0
        for p in order:
            n = p.numel()
            if cur.params and (pos - cur.start) + n > cap:
                pos = _round_up(pos, seg_align)
                cur.end = pos
                self.segments.append(cur)
                cur = Segment(len(self.segments), pos)
                cap = int(bucket_cap_mb * 1024 * 1024) // 4
            self.offset[id(p)] = pos
            cur.params.append(p)
            pos = _round_up(pos + n, _ALIGN)
        pos = _round_up(pos, seg_align)
        cur.end = pos
        self.segments.append(cur)
        self.total = pos
        self.grads_zeroed = False
        self.weights_version = 0      # bumped by every writer of the weights; cached re-layouts compare against it
        self.segment_of = {id(p): s for s in self.segments for p in s.params}

        # ---- storage ---------------------------------------------------------------------
        self.param_buf = fabric.alloc(self.total * 4)
        self.grad_buf = fabric.alloc(self.total * 4)
        self.flat_params = self.param_buf.view(0, self.total, torch.float32)
        self.flat_grads = self.grad_buf.view(0, self.total, torch.float32)
        self.momentum = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        self.shadow_buf = None
        self.flat_shadow = None
        if shadow_dtype is not None:
            self.shadow_buf = fabric.alloc(self.total * 2)
            self.flat_shadow = self.shadow_buf.view(0, self.total, shadow_dtype)

        # ---- move the parameters in ----------------------------------------------------------
        with torch.no_grad():
            for p in order:
                view = self._view(self.flat_params, p)
                view.copy_(p.data)
                p.data = view
                if p.grad is not None:      # built after a backward (lazy optimizer binding): keep it
                    gv = self._view(self.flat_grads, p)
                    gv.copy_(p.grad)
                    p.grad = gv
                p._pdt_arena = self
        self.refresh_shadow()

    # -- views ---------------------------------------------------------------------------
    def _view(self, flat, p):
        off, n = self.offset[id(p)], p.numel()
        t = flat[off:off + n]
        if p.dim() == 4:   # OIHW logical, KRSC physical
            o, i, h, w = p.shape
            return t.view(o, h, w, i).permute(0, 3, 1, 2)
        return t.view(p.shape)

    def grad_view(self, p):
        return self._view(self.flat_grads, p)

    def momentum_view(self, p):
        return self._view(self.momentum, p)

    def shadow_view(self, p):
        return None if self.flat_shadow is None else self._view(self.flat_shadow, p)

    def weights_changed(self):
        """Every writer of the parameter arena calls this (optimizer kernels, load_state_dict, broadcasts): layouts
        derived from the weights (the data-gradient filters cached by ops.conv_native) are stale until re-made."""
        self.weights_version += 1

    def refresh_shadow(self):
        """Recompute the low-precision shadow from the fp32 master (init / load_state_dict)."""
        if self.flat_shadow is not None:
            self.flat_shadow.copy_(self.flat_params)
        self.weights_changed()

    def watch_state_dict(self, module):
        """``module.load_state_dict`` writes the fp32 master weights in place; keep the shadow in step."""
        module.register_load_state_dict_post_hook(lambda m, incompatible: self.refresh_shadow())

    def gather_grads(self, drop=False):
        """Copy every ``p.grad`` that is not already an arena view into the gradient arena."""
        src, dst = [], []
        for p in self.params:
            v = self.grad_view(p)
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                src.append(p.grad)
                dst.append(v)
            if drop:
                p.grad = None
        if src:
            torch._foreach_copy_(dst, src)

    def zero_grads(self):
        """One memset for the whole gradient arena.  Afterwards (until the next optimizer step) the native
        weight-gradient kernels accumulate straight into the arena views, so no per-parameter zero-fill and
        no gradient -> bucket copy is ever launched."""
        self.flat_grads.zero_()
        self.grads_zeroed = True

    def direct_grad(self, p):
        """A fresh view of ``p``'s slot in the gradient arena if a backward kernel may write (accumulate)
        into it directly, else None."""
        if not self.grads_zeroed or p.grad is not None or id(p) not in self.offset or not _DIRECT:
            return None
        return self.grad_view(p)

    def describe(self):
        return {"total_elems": self.total,
                "segments": [(s.start * 4, s.length * 4, len(s.params)) for s in self.segments]}
