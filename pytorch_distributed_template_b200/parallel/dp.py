"""Single-process multi-GPU DataParallel.

User contract of ``nn.DataParallel`` as the reference uses it
(``nn.DataParallel(model).cuda()``, /root/reference/dataparallel.py:119): the
whole batch arrives on GPU 0, ``forward`` returns logits for the whole batch on
GPU 0, ``.module`` is the real network whose parameters the optimizer owns.

What is different underneath:
* replicas are PERSISTENT (built once) instead of re-broadcast every forward
  (D3: 46.8 MB x 7 per step in the reference);
* scatter (D2), gather (D5), the gradient-of-gather (D7a) and the BN-buffer
  broadcast are destination-side pull kernels over NVLink peer mappings
  (``pdt_peer_copy``), not NCCL / cudaMemcpyPeer calls;
* the gradient reduce-to-GPU-0 + SGD-on-GPU-0 + next-forward broadcast
  (D7, D8, D3) is ONE pair of kernels per device: device d sums shard d of
  every replica's gradient arena through peer loads, applies momentum SGD to
  that shard and stores the new weights (and bf16/fp16 shadow) into every
  replica's arena through peer stores — the optimizer work is spread over all
  GPUs and no parameter ever travels as a separate broadcast.
* replica forwards are driven from one persistent thread per device; backward
  uses autograd's own per-device threads (N4).

Without CUDA the wrapper is transparent (like torch's with no GPUs).
"""
import copy
import threading

import torch
import torch.nn as nn

from ..ops import _native
from .arena import ParamArena
from .ddp import BufferArena
from .fabric import Fabric


def _peer_copy(src, dst):
    """dst (on the current device) <- src (possibly on a peer); both contiguous, same nbytes."""
    nbytes = dst.numel() * dst.element_size()
    if _native.available() and src.data_ptr() % 16 == 0 and dst.data_ptr() % 16 == 0:
        _native.check(_native.lib().pdt_peer_copy(src.data_ptr(), dst.data_ptr(), nbytes,
                                                  _native.stream_ptr(dst.device)), "pdt_peer_copy")
    else:
        dst.copy_(src, non_blocking=True)


class _Scatter(torch.autograd.Function):
    """Chunk a GPU-0 tensor along dim 0 onto the devices; each device pulls its slice."""

    @staticmethod
    def forward(ctx, dp, x):
        chunks = x.chunk(len(dp.device_ids), 0)
        outs = []
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(x.device))
        for dev, c in zip(dp.devices, chunks):
            if dev == x.device:
                outs.append(c)
                continue
            with torch.cuda.device(dev), torch.cuda.stream(dp.stream_of(dev)):
                torch.cuda.current_stream(dev).wait_event(ev)
                o = torch.empty(c.shape, dtype=c.dtype, device=dev)
                _peer_copy(c.contiguous(), o)
                outs.append(o)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *grads):
        return None, None


class _Gather(torch.autograd.Function):
    """Concatenate per-device outputs on GPU 0 (pull), and scatter the gradient back (pull)."""

    @staticmethod
    def forward(ctx, dp, *outs):
        ctx.dp = dp
        ctx.sizes = [o.shape[0] for o in outs]
        dev0 = dp.devices[0]
        full = torch.empty((sum(ctx.sizes),) + tuple(outs[0].shape[1:]), dtype=outs[0].dtype, device=dev0)
        pos = 0
        with torch.cuda.device(dev0):
            s0 = torch.cuda.current_stream(dev0)
            for o in outs:
                if o.device != dev0:
                    ev = torch.cuda.Event()
                    ev.record(dp.stream_of(o.device))
                    s0.wait_event(ev)
                _peer_copy(o.contiguous(), full[pos:pos + o.shape[0]])
                pos += o.shape[0]
        return full

    @staticmethod
    def backward(ctx, g):
        dp = ctx.dp
        g = g.contiguous()
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(g.device))
        outs, pos = [], 0
        for dev, n in zip(dp.devices, ctx.sizes):
            piece = g[pos:pos + n]
            pos += n
            if dev == g.device:
                outs.append(piece)
                continue
            with torch.cuda.device(dev), torch.cuda.stream(dp.stream_of(dev)):
                torch.cuda.current_stream(dev).wait_event(ev)
                o = torch.empty(piece.shape, dtype=piece.dtype, device=dev)
                _peer_copy(piece, o)
                outs.append(o)
        return (None,) + tuple(outs)


class _Worker(threading.Thread):
    """One persistent launch thread per device (the reference spawns W threads per forward)."""

    def __init__(self, device, stream=None):
        super().__init__(daemon=True)
        self.device = device
        self.stream = stream
        self.job = None
        self.result = None
        self.cv = threading.Condition()
        self.start()

    def run(self):
        torch.cuda.set_device(self.device)
        if self.stream is not None:      # the device's one work stream (see DataParallel._build)
            torch.cuda.set_stream(self.stream)
        while True:
            with self.cv:
                while self.job is None:
                    self.cv.wait()
                fn, args = self.job
            try:
                res = (True, fn(*args))
            except BaseException as e:  # propagate to the caller
                res = (False, e)
            with self.cv:
                self.job, self.result = None, res
                self.cv.notify_all()

    def submit(self, fn, *args):
        with self.cv:
            self.result = None
            self.job = (fn, args)
            self.cv.notify_all()

    def wait(self):
        with self.cv:
            while self.result is None:
                self.cv.wait()
            ok, val = self.result
        if not ok:
            raise val
        return val


class GraphStep:
    """One DataParallel training step as 3 W + 1 CUDA graphs replayed from ONE thread.

    The reference drives W replicas with W Python threads per forward and autograd's device threads per
    backward, all serialised by the GIL (torch/nn/parallel/parallel_apply.py:123-136); at 150 images per GPU
    the launches, not the GPUs, set the pace.  Here the step is cut where data crosses devices and each piece
    is captured once on its device's stream:

      F[d]  (device d)  pull rank 0's BN buffers and slice d of the input batch over NVLink, replica forward
      L     (device 0)  pull every replica's logits, loss + top-1 + meters, gradient of the loss w.r.t. the logits
      B[d]  (device d)  pull slice d of that gradient, clear the gradient arena, replica backward
      S[d]  (device d)  sum shard d of all W gradient arenas, momentum SGD on it, push the new weights to all

    Between the pieces only events order the devices (F[*] -> L -> B[*] -> S[*] -> next F[*]); a replay is
    3 W + 1 graph launches and the same number of event records from the calling thread, no worker threads,
    no host synchronisation.  ``images`` / ``target`` are static device-0 tensors the caller refills."""

    def __init__(self, dp, images, target, loss_fn, metrics_fn, amp_dtype=None):
        from .. import amp
        from ..engine import CAPTURE_MODE
        self.dp, self.images, self.target = dp, images, target
        W = len(dp.devices)
        devs, streams = dp.devices, dp._streams
        n = images.shape[0]
        bounds = [0]
        for c in images.chunk(W, 0):
            bounds.append(bounds[-1] + c.shape[0])
        self.bounds = bounds
        self.gF, self.gB, self.gS = [None] * W, [None] * W, [None] * W
        self.evF = [torch.cuda.Event() for _ in range(W)]
        self.evB = [torch.cuda.Event() for _ in range(W)]
        self.evS = [torch.cuda.Event() for _ in range(W)]
        self.evL = torch.cuda.Event()
        self.evIn = torch.cuda.Event()
        outs = [None] * W
        sig = tuple(m.training for m in dp.module.modules())
        for rep in dp.replicas[1:]:
            for m, flag in zip(rep.modules(), sig):
                m.training = flag
        dp._train_sig = sig
        n0 = _native.launch_count
        # ---- F[d]
        for d, dev in enumerate(devs):
            with torch.cuda.device(dev):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=streams[d], capture_error_mode=CAPTURE_MODE):
                    if d > 0 and dp.buffer_arenas[0].used:
                        _peer_copy(dp.buffer_arenas[0].flat, dp.buffer_arenas[d].flat)
                    xs = images[bounds[d]:bounds[d + 1]]
                    if d > 0:
                        xd = torch.empty(xs.shape, dtype=xs.dtype, device=dev)
                        _peer_copy(xs, xd)
                    else:
                        xd = xs
                    with torch.enable_grad(), amp.autocast(enabled=amp_dtype is not None, dtype=amp_dtype):
                        outs[d] = dp.replicas[d](xd)
                self.gF[d] = g
        # ---- L
        dev0 = devs[0]
        with torch.cuda.device(dev0):
            full = torch.empty((n,) + tuple(outs[0].shape[1:]), dtype=outs[0].dtype, device=dev0)
            self.gL = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.gL, stream=streams[0], pool=self.gF[0].pool(), capture_error_mode=CAPTURE_MODE):
                for d in range(W):
                    _peer_copy(outs[d].detach(), full[bounds[d]:bounds[d + 1]])
                leaf = full.detach().requires_grad_(True)
                with torch.enable_grad():
                    loss, pair = loss_fn(leaf, target)
                metrics_fn(pair, n)                 # also copies the reduced (loss, acc1) to the runner's pinned host slot
                (dfull,) = torch.autograd.grad(loss, leaf)
        self.full, self.dfull = full, dfull
        # ---- B[d]
        for d, dev in enumerate(devs):
            with torch.cuda.device(dev):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=streams[d], pool=self.gF[d].pool(), capture_error_mode=CAPTURE_MODE):
                    piece = dfull[bounds[d]:bounds[d + 1]]
                    if d > 0:
                        dl = torch.empty(piece.shape, dtype=piece.dtype, device=dev)
                        _peer_copy(piece, dl)
                    else:
                        dl = piece
                    for p in dp.replicas[d].parameters():
                        p.grad = None
                    dp.arenas[d].zero_grads()
                    torch.autograd.backward([outs[d]], [dl.to(outs[d].dtype)])
                    dp.arenas[d].gather_grads(drop=False)
                    dp.arenas[d].grads_zeroed = False
                self.gB[d] = g
        outs = None
        # ---- S[d]
        lib = _native.lib()
        sh_dtype = _native.dtype_code(dp.shadow_dtype) if dp.shadow_dtype is not None else _native.BF16
        for d, dev in enumerate(devs):
            a = dp.arenas[d]
            with torch.cuda.device(dev):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=streams[d], pool=self.gF[d].pool(), capture_error_mode=CAPTURE_MODE):
                    s_ = streams[d].cuda_stream
                    _native.check(lib.pdt_reduce_scatter_f32(dp._comms[d], dp._grads, None, 0, a.total, dp._state[d].data_ptr(), 0,
                                                            128, s_), "pdt_reduce_scatter_f32(dp)")
                    _native.check(lib.pdt_sgd_allgather(dp._comms[d], dp._grads, dp._params, dp._shadow, None, None, sh_dtype,
                                                       a.momentum.data_ptr(), dp._whole, dp._hypers[d].data_ptr(),
                                                       dp._state[d].data_ptr(), 0, 128, s_), "pdt_sgd_allgather(dp)")
                self.gS[d] = g
        self.launches = _native.launch_count - n0

    def replay(self, opt):
        dp = self.dp
        devs, streams = dp.devices, dp._streams
        W = len(devs)
        dp.sync_hyper(opt)
        self.evIn.record(streams[0])                      # the caller filled the input buffers on device 0's stream
        for d in range(W):
            s = streams[d]
            if d > 0:
                s.wait_event(self.evIn)
            for e in self.evS:                            # last step's weight pushes (and gradient reads) are done
                s.wait_event(e)
            with torch.cuda.device(devs[d]), torch.cuda.stream(s):
                self.gF[d].replay()
            self.evF[d].record(s)
        s0 = streams[0]
        for e in self.evF[1:]:
            s0.wait_event(e)
        with torch.cuda.device(devs[0]), torch.cuda.stream(s0):
            self.gL.replay()
        self.evL.record(s0)
        for d in range(W):
            s = streams[d]
            if d > 0:
                s.wait_event(self.evL)
            with torch.cuda.device(devs[d]), torch.cuda.stream(s):
                self.gB[d].replay()
            self.evB[d].record(s)
        for d in range(W):
            s = streams[d]
            for e in self.evB:
                s.wait_event(e)
            with torch.cuda.device(devs[d]), torch.cuda.stream(s):
                self.gS[d].replay()
            self.evS[d].record(s)
        for e in self.evS[1:]:                            # device 0's stream ends the step: it has seen every push
            s0.wait_event(e)
        for a in dp.arenas:
            a.weights_changed()


class DataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, dim=0, shadow_dtype=None):
        super().__init__()
        if dim != 0:
            raise ValueError("only dim=0 is supported (what the reference uses)")
        self.module = module
        self.dim = 0
        if device_ids is None:
            device_ids = list(range(torch.cuda.device_count())) if torch.cuda.is_available() else []
        self.device_ids = [d if isinstance(d, int) else torch.device(d).index for d in device_ids]
        if len(self.device_ids) > _native.MAX_RANKS:
            raise ValueError("DataParallel drives at most {} devices (got {})".format(_native.MAX_RANKS, len(self.device_ids)))
        self.devices = [torch.device("cuda", d) for d in self.device_ids]
        self.output_device = self.device_ids[0] if self.device_ids else None
        self.shadow_dtype = shadow_dtype
        self.optimizer = None
        self._built = False
        self._train_sig = None
        self.replicas, self.arenas, self.buffer_arenas, self._workers = [], [], [], []
        self._state = []

    # ------------------------------------------------------------------ lazy construction
    def _build(self):
        """Called at the first forward: by then ``.cuda()`` has moved the module to GPU 0."""
        dev0 = self.devices[0]
        if next(self.module.parameters()).device != dev0:
            self.module.to(dev0)
        for a in self.devices:       # map every pair once (torch enables peer access on first copy)
            for b in self.devices:
                if a != b:
                    if not torch.cuda.can_device_access_peer(a.index, b.index):
                        raise RuntimeError("DataParallel needs peer access between GPU {} and {}".format(a, b))
                    torch.zeros(1, device=a).to(b)
        self.replicas = [self.module] + [copy.deepcopy(self.module).to(d) for d in self.devices[1:]]
        for rep, dev in zip(self.replicas, self.devices):
            with torch.cuda.device(dev):
                fab = Fabric(dev, group="solo")
                arena = ParamArena(rep.parameters(), fab, shadow_dtype=self.shadow_dtype, shard_world=len(self.devices))
                arena.owner = self
                self.arenas.append(arena)
                self.buffer_arenas.append(BufferArena(rep, fab))
                st = torch.zeros(8, dtype=torch.float32, device=dev)
                st[0] = 1.0 / len(self.devices)      # k_reduce_scatter multiplies by 1/(W*scale): DP grads are sums
                self._state.append(st)
        # ONE work stream per device for everything this wrapper does there (forward in the launch threads,
        # backward in autograd's device threads, which follow the forward's stream, the optimizer kernels, and
        # the captured graphs): autograd remembers the stream of the first forward, and a graph capture must not
        # have to synchronise with the legacy default stream.
        self._streams = [torch.cuda.current_stream(dev0)] + [torch.cuda.Stream(d) for d in self.devices[1:]]
        keep_dev = torch.cuda.current_device()
        for st_ in self._streams[1:]:
            torch.cuda.set_stream(st_)           # per-device current stream of THIS thread (also switches device)
        torch.cuda.set_device(keep_dev)
        self._workers = [None] + [_Worker(d, st_) for d, st_ in zip(self.devices[1:], self._streams[1:])]
        W = len(self.devices)
        self._grads = _native.make_peers([a.grad_buf.ptrs[0] for a in self.arenas])
        self._params = _native.make_peers([a.param_buf.ptrs[0] for a in self.arenas])
        self._shadow = _native.make_peers([a.shadow_buf.ptrs[0] for a in self.arenas]) if self.shadow_dtype else None
        self._comms = [_native.make_comm(r, W, []) for r in range(W)]
        self._hypers = [torch.zeros(4, dtype=torch.float32, device=d) for d in self.devices]
        self._hyper_host = None
        self._whole = _native.make_segments([(0, self.arenas[0].total)])
        for d in self.devices:
            torch.cuda.synchronize(d)
        self._built = True
        self.module.register_load_state_dict_post_hook(lambda m, incompatible: self._push_master())

    @torch.no_grad()
    def _push_master(self):
        """After ``load_state_dict`` on the real module: replicas and every shadow follow device 0's master weights."""
        a0 = self.arenas[0]
        torch.cuda.synchronize(self.devices[0])
        for dev, a in zip(self.devices, self.arenas):
            with torch.cuda.device(dev):
                if a is not a0:
                    a.flat_params.copy_(a0.flat_params)
                a.refresh_shadow()
        for d in self.devices:
            torch.cuda.synchronize(d)

    @property
    def arena(self):
        return self.arenas[0] if self.arenas else None

    def _bcast_buffers(self):
        ba0 = self.buffer_arenas[0]
        if ba0.used == 0:
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.devices[0]))
        for dev, ba in zip(self.devices[1:], self.buffer_arenas[1:]):
            with torch.cuda.device(dev), torch.cuda.stream(self.stream_of(dev)):
                torch.cuda.current_stream(dev).wait_event(ev)
                _peer_copy(ba0.flat, ba.flat)

    # ------------------------------------------------------------------ forward
    def forward(self, *inputs, **kwargs):
        if len(self.device_ids) <= 1 or not torch.cuda.is_available():
            return self.module(*inputs, **kwargs)
        if not self._built:
            self._build()
        x = inputs[0]
        if x.shape[0] < len(self.devices):
            return self.module(*inputs, **kwargs)
        # replicas mirror the training flag of EVERY sub-module (a BatchNorm the user froze with .eval() stays
        # frozen on all devices, like torch's replicate()); 68 attribute reads, applied only on change
        sig = tuple(m.training for m in self.module.modules())
        if sig != self._train_sig:
            for rep in self.replicas[1:]:
                for m, flag in zip(rep.modules(), sig):
                    m.training = flag
            self._train_sig = sig
        self._bcast_buffers()
        xs = _Scatter.apply(self, x)
        grad_on = torch.is_grad_enabled()
        from .. import amp
        lp = amp.active_dtype()

        def run(rep, xi):
            with torch.set_grad_enabled(grad_on), amp.autocast(enabled=lp is not None, dtype=lp):
                return rep(xi)

        for w, rep, xi in zip(self._workers[1:], self.replicas[1:], xs[1:]):
            w.submit(run, rep, xi)
        outs = [run(self.replicas[0], xs[0])] + [w.wait() for w in self._workers[1:]]
        return _Gather.apply(self, *outs)

    def stream_of(self, device):
        """The work stream of one of this wrapper's devices (see _build)."""
        return self._streams[self.devices.index(torch.device(device))]

    def sync_hyper(self, opt):
        """lr / momentum / weight decay live in device memory the optimizer kernels read (also from replayed
        graphs); copies happen only when the schedule changes them."""
        g = opt.param_groups[0]
        hyper = (float(g["lr"]), float(g["momentum"]), float(g["weight_decay"]), 0.0)
        if hyper != self._hyper_host:
            host = torch.tensor(hyper, dtype=torch.float32).pin_memory()
            for r, dev in enumerate(self.devices):
                with torch.cuda.device(dev):
                    self._hypers[r].copy_(host, non_blocking=True)
            for dev in self.devices:
                torch.cuda.synchronize(dev)          # `host` may go away; happens once per LR change
            self._hyper_host = hyper

    def graph_step(self, images, target, loss_fn, metrics_fn, amp_dtype=None):
        """Capture this wrapper's training step for the static device-0 batch (images, target): see GraphStep."""
        if not self._built:
            raise RuntimeError("run at least one eager step before capturing (replicas are built at the first forward)")
        return GraphStep(self, images, target, loss_fn, metrics_fn, amp_dtype)

    # ------------------------------------------------------------------ fused optimizer step
    def fused_step(self, opt):
        """D7 + D8 + D3 in one pair of kernels per device (see module docstring)."""
        W = len(self.devices)
        lib = _native.lib()
        self.sync_hyper(opt)
        # phase 1: every device moves its replica's gradients into its arena
        done = []
        for r, dev in enumerate(self.devices):
            with torch.cuda.device(dev), torch.cuda.stream(self._streams[r]):
                self.arenas[r].gather_grads(drop=(r != 0))
                self.arenas[r].grads_zeroed = False
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(dev))
                done.append(ev)
        sh_dtype = _native.dtype_code(self.shadow_dtype) if self.shadow_dtype is not None else _native.BF16
        after = []
        # phase 2: once ALL arenas are complete, device r reduces shard r over every arena,
        # applies SGD to it and stores the new weights into every replica
        for r, dev in enumerate(self.devices):
            a = self.arenas[r]
            with torch.cuda.device(dev), torch.cuda.stream(self._streams[r]):
                s = self._streams[r]
                for ev in done:
                    s.wait_event(ev)
                _native.check(lib.pdt_reduce_scatter_f32(self._comms[r], self._grads, None, 0, a.total,
                                                        self._state[r].data_ptr(), 0, 128, s.cuda_stream),
                              "pdt_reduce_scatter_f32(dp)")
                _native.check(lib.pdt_sgd_allgather(self._comms[r], self._grads, self._params, self._shadow, None, None,
                                                   sh_dtype, a.momentum.data_ptr(), self._whole, self._hypers[r].data_ptr(),
                                                   self._state[r].data_ptr(), 0, 128, s.cuda_stream),
                              "pdt_sgd_allgather(dp)")
                ev = torch.cuda.Event()
                ev.record(s)
                after.append(ev)
        for a in self.arenas:
            a.weights_changed()
        for r, dev in enumerate(self.devices):      # new weights were written by peers: everybody waits for everybody
            s = self._streams[r]
            for ev in after:
                s.wait_event(ev)
            if r > 0:      # replicas' gradients were dropped above: clear their arenas for direct writes
                with torch.cuda.device(dev), torch.cuda.stream(s):
                    self.arenas[r].zero_grads()

    def zero_replica_grads(self):
        for rep in self.replicas[1:]:
            for p in rep.parameters():
                p.grad = None
