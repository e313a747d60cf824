"""Optimizer and LR schedule (SURVEY layer L6, N12).

``SGD`` has ``torch.optim.SGD``'s constructor and maths
(torch/optim/sgd.py:343-380 with dampening 0, no nesterov: ``g += wd*p;
buf = mu*buf + g; p -= lr*buf``; weight decay applies to every parameter
including BN affine terms, /root/reference/distributed.py:148) but operates on
the flat arenas:

* under our DDP wrapper, ``step()`` enqueues ``pdt_sgd_allgather`` per segment:
  inf consensus, SGD on the shard this rank owns, then the new fp32 weights and
  their low-precision shadow are stored into every peer's arena over NVLink —
  the second half of the gradient all-reduce never happens;
* stand-alone on one GPU it is one ``pdt_sgd_flat`` launch instead of the
  reference's 3-4 foreach launch groups over 62 tensors;
* on CPU it is the same arithmetic with torch ops (the oracle).

``MultiStepLR`` is the closed form the reference's ``lr_scheduler.step(epoch)``
reaches: epochs 0-2 -> 0.1, 3 -> 0.01, 4 -> 0.001 with the defaults.
"""
import bisect

import torch
import torch.distributed as dist

from .ops import _native
from .parallel.arena import ParamArena
from .parallel.fabric import Fabric


class SGD(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
        if dampening != 0.0 or nesterov:
            raise ValueError("fused SGD implements dampening=0, nesterov=False (what the reference uses)")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening,
                                      weight_decay=weight_decay, nesterov=nesterov))
        if len(self.param_groups) != 1:
            raise ValueError("fused SGD keeps one parameter group (distributed.py:148)")
        self.arena = None
        self.owner = None       # DistributedDataParallel / DataParallel that owns the arena, if any
        self.scaler = None
        self._hyper = None
        self._hyper_host = None
        self._unit_state = None
        # Under DDP the arena already exists and the gradient kernels must know about the
        # fused optimizer before the first backward.  A DataParallel wrapper builds its
        # replicas at the first forward, so there binding waits for the first step().
        if getattr(self.param_groups[0]["params"][0], "_pdt_arena", None) is not None:
            self._bind()

    @property
    def ddp(self):
        return self.owner if hasattr(self.owner, "_launch_segment") else None

    def _bind(self):
        plist = self.param_groups[0]["params"]
        arena = getattr(plist[0], "_pdt_arena", None)
        if arena is None:
            arena = ParamArena(plist, Fabric(plist[0].device, group="solo"), shadow_dtype=None)
        self.arena = arena
        self.owner = getattr(arena, "owner", None)
        if self.ddp is not None:
            self.owner.optimizer = self
        self.device = arena.device
        self.native = self.device.type == "cuda" and _native.available()
        self._hyper = torch.zeros(4, dtype=torch.float32, device=self.device)
        self.sync_hyper()
        for p in plist:
            if id(p) in arena.offset:
                self.state[p]["momentum_buffer"] = arena.momentum_view(p)
        if self.scaler is not None and self.ddp is not None:
            self.owner.amp_state = self.scaler.state

    # -- hyper-parameters live on the device so a captured graph sees LR changes ---------
    def sync_hyper(self):
        if self._hyper is None:
            return
        g = self.param_groups[0]
        amp = 1.0 if (self.scaler is not None and self.scaler.is_enabled()) else 0.0
        host = (float(g["lr"]), float(g["momentum"]), float(g["weight_decay"]), amp)
        if host != self._hyper_host:
            self._hyper.copy_(torch.tensor(host, dtype=torch.float32))
            self._hyper_host = host

    def attach_scaler(self, scaler):
        self.scaler = scaler
        if self.ddp is not None:
            self.owner.amp_state = scaler.state
        self.sync_hyper()

    def _amp_state(self):
        if self.scaler is not None and self.scaler.state is not None:
            return self.scaler.state
        if self._unit_state is None:
            self._unit_state = torch.zeros(8, dtype=torch.float32, device=self.device)
            self._unit_state[0] = 1.0
        return self._unit_state

    def zero_grad(self, set_to_none=True):
        # Gradients are views into the arena: drop the references and clear the arena with ONE memset;
        # backward kernels then write their results straight into it (ParamArena.direct_grad).
        for p in self.param_groups[0]["params"]:
            p.grad = None
        if self.arena is not None and self.device.type == "cuda":
            self.arena.zero_grads()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self.arena is None:
            self._bind()
        capturing = self.device.type == "cuda" and torch.cuda.is_current_stream_capturing()
        if not capturing:
            self.sync_hyper()
        a = self.arena
        a.grads_zeroed = False          # the arena now holds this step's gradients
        a.weights_changed()             # cached data-gradient filter layouts are stale after this step
        if self.owner is not None and not hasattr(self.owner, "_launch_segment"):
            self.owner.fused_step(self)          # DataParallel: per-device reduce + SGD + push
        elif self.native:
            lib = _native.lib()
            st = self._amp_state()
            sh_dtype = _native.dtype_code(a.shadow_dtype) if a.shadow_dtype is not None else _native.BF16
            if self.ddp is not None:
                ddp = self.ddp
                stream = ddp._comm_stream
                stream.wait_stream(torch.cuda.current_stream(self.device))
                nvls = ddp.fabric.use_nvls
                with torch.cuda.stream(stream):
                    # every segment in ONE launch (one inf consensus, one pair of cross-GPU barriers); nothing else
                    # runs at this point of the step, so the kernel may take most of the GPU
                    for lo in range(0, len(a.segments), _native.MAX_SEGMENTS):
                        segs = _native.make_segments((s.start, s.length) for s in a.segments[lo:lo + _native.MAX_SEGMENTS])
                        _native.check(lib.pdt_sgd_allgather(
                            ddp.fabric.comm, a.grad_buf.peers(), a.param_buf.peers(),
                            a.shadow_buf.peers() if a.shadow_buf is not None else None,
                            a.param_buf.mc(nvls), a.shadow_buf.mc(nvls) if a.shadow_buf is not None else None, sh_dtype,
                            a.momentum.data_ptr(), segs, self._hyper.data_ptr(), st.data_ptr(),
                            1 if ddp.world > 1 else 0, ddp.step_blocks, stream.cuda_stream), "pdt_sgd_allgather")
                ddp.join_comm()
            else:
                a.gather_grads()
                s = _native.stream_ptr(self.device)
                if self.scaler is not None and self.scaler.is_enabled():
                    _native.check(lib.pdt_reduce_scatter_f32(
                        a.fabric.comm, a.grad_buf.peers(), None, 0, a.total, st.data_ptr(), 0, 148, s),
                        "pdt_reduce_scatter_f32(unscale)")
                _native.check(lib.pdt_sgd_flat(
                    a.flat_params.data_ptr(), a.flat_grads.data_ptr(), a.momentum.data_ptr(),
                    a.flat_shadow.data_ptr() if a.flat_shadow is not None else None, sh_dtype, a.total,
                    self._hyper.data_ptr(), st.data_ptr(), s), "pdt_sgd_flat")
        else:
            self._step_torch()
        return loss

    def _step_torch(self):
        """Oracle / CPU path: identical arithmetic on the flat tensors."""
        a = self.arena
        g = self.param_groups[0]
        lr, mu, wd = g["lr"], g["momentum"], g["weight_decay"]
        if self.ddp is None:
            a.gather_grads()
        grads = a.flat_grads
        if self.scaler is not None and self.scaler.is_enabled():
            st = self.scaler.state
            grads.mul_(1.0 / st[0])
            bad = (~torch.isfinite(grads)).any().float()
            if dist.is_initialized() and dist.get_world_size() > 1:
                dist.all_reduce(bad, op=dist.ReduceOp.MAX)
            skip = bool(bad.item() > 0)
            st.view(torch.int32)[4] = int(skip)
            if skip:
                return
        d = grads.add(a.flat_params, alpha=wd) if wd != 0 else grads
        if mu != 0:
            a.momentum.mul_(mu).add_(d)
            d = a.momentum
        a.flat_params.add_(d, alpha=-lr)
        a.refresh_shadow()

    def consolidate_state(self):
        """Momentum is sharded across ranks under fused DDP; gather it before ``state_dict()``."""
        if self.ddp is None or self.ddp.world == 1 or not self.native:
            return
        a, W, r = self.arena, self.ddp.world, self.ddp.rank
        full = torch.zeros_like(a.momentum)
        for seg in a.segments:
            sh = seg.length // W
            lo = seg.start + r * sh
            full[lo:lo + sh] = a.momentum[lo:lo + sh]
        dist.all_reduce(full, group=self.ddp.process_group)
        a.momentum.copy_(full)


class MultiStepLR:
    """``lr = base_lr * gamma ** (#milestones <= epoch)``; ``step(epoch)`` like distributed.py:192."""

    def __init__(self, optimizer, milestones, gamma=0.1, last_epoch=-1):
        self.optimizer = optimizer
        self.milestones = sorted(int(m) for m in milestones)
        self.gamma = gamma
        self.base_lrs = [g["lr"] for g in optimizer.param_groups]
        self.last_epoch = last_epoch
        self.step(0 if last_epoch < 0 else last_epoch)

    def lr_at(self, epoch):
        k = bisect.bisect_right(self.milestones, epoch)
        return [b * self.gamma ** k for b in self.base_lrs]

    def step(self, epoch=None):
        self.last_epoch = self.last_epoch + 1 if epoch is None else int(epoch)
        for g, lr in zip(self.optimizer.param_groups, self.lr_at(self.last_epoch)):
            g["lr"] = lr
        if hasattr(self.optimizer, "sync_hyper"):
            self.optimizer.sync_hyper()

    def get_last_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def state_dict(self):
        return {"milestones": self.milestones, "gamma": self.gamma, "base_lrs": self.base_lrs,
                "last_epoch": self.last_epoch}

    def load_state_dict(self, sd):
        self.milestones, self.gamma = sd["milestones"], sd["gamma"]
        self.base_lrs, self.last_epoch = sd["base_lrs"], sd["last_epoch"]
