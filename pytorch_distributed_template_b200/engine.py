"""The training step as a product feature: ``StepRunner`` + ``DeviceFeed``.

The reference's hot loop (/root/reference/distributed_syncBN_amp.py:250-288; distributed.py:237-273;
dataparallel.py:203-232) is ~10^3 eager kernel launches and 2-23 host synchronisations per iteration.
Here one iteration — forward, loss + top-1, the cross-rank metric reduction with its device-side
``AverageMeter`` update, ``zero_grad``, backward with the in-kernel gradient reduce-scatter, the fused
SGD / weight all-gather, the ``GradScaler`` update — is ONE ``StepRunner.step`` call that

* runs eagerly for the first few iterations (lazy initialisation: replicas, scratch buffers, layout plans),
* is then captured ONCE per input staging buffer into a CUDA graph (cross-GPU kernels included: nothing in
  the step touches the host) and afterwards replayed — one launch per iteration,
* never waits for the CPU: the loss / accuracy averages live in a device meter read only at print time,
  the learning rate and the loss scale live in device memory the captured kernels read, so ``lr_scheduler.step``
  and scaler back-offs need no re-capture.

``DeviceFeed`` is the matching input side: batches go from (pinned) host memory straight into the runner's
static staging buffers on a copy stream, double-buffered, and each buffer has its own graph, so there is no
device-to-device copy and the H2D transfer of batch i+1 overlaps the compute of batch i.

Both ``trainer.train`` (the three entry scripts) and ``bench.py`` drive training through this module; what the
benchmark times is what a user of ``start.sh`` runs.  Environment: ``PDT_GRAPH=0`` keeps every step eager,
``PDT_GRAPH_WARMUP`` sets the number of eager iterations before capture (default 3), ``PDT_NVTX=0`` drops the
NVTX ranges (forward / backward / optimizer / metrics) that profilers pick up.
"""
import contextlib
import os
import sys

import torch
import torch.distributed as dist

from . import amp
from .ops import _native

_streams = {}

# Graph captures run in "thread_local" error mode: a training process has other threads that legitimately call the
# CUDA runtime while a capture is open (the DataLoader's pin-memory thread allocates pinned buffers), which the
# default "global" mode turns into an invalidated capture.  Kernels that autograd's device threads launch into the
# capturing stream are captured in every mode.
CAPTURE_MODE = "thread_local"


def work_stream(device):
    """The one non-default stream everything of a training process runs on, made current on first use.

    Autograd's AccumulateGrad nodes remember the stream of the first forward; a capture that has to synchronise
    with the legacy default stream is invalid.  So construction, warm-up, capture, replay and validation all use
    this stream — call it before the model's first forward (``trainer.setup`` and ``StepRunner`` do)."""
    device = torch.device(device)
    if device.type != "cuda":
        return None
    key = device.index if device.index is not None else torch.cuda.current_device()
    s = _streams.get(key)
    if s is None:
        s = _streams[key] = torch.cuda.Stream(device)
    if torch.cuda.current_stream(device) != s:
        torch.cuda.set_stream(s)
    return s


_NVTX = os.environ.get("PDT_NVTX", "1") == "1"


@contextlib.contextmanager
def nvtx_range(name):
    """NVTX range around a phase of the step (visible in nsys / ncu --nvtx; free without a profiler)."""
    on = _NVTX and torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class Meter:
    """Device-resident AverageMeters for (loss, top-1): float32[8] = {sum_loss, sum_acc, count, val_loss, val_acc}.
    Updated by ``pdt_metrics_update`` inside the step; ``read()`` is the only host synchronisation."""

    def __init__(self, device):
        self.t = torch.zeros(8, dtype=torch.float32, device=device)

    def reset(self):
        self.t.zero_()

    def read(self):
        sl, sa, c, vl, va = self.t.tolist()[:5]
        c = c if c else 1.0
        return {"loss": sl / c, "acc1": sa / c, "count": c, "val_loss": vl, "val_acc1": va}


class StepRunner:
    def __init__(self, model, criterion, optimizer, scaler=None, device=None, use_amp=False, amp_dtype=None,
                 graph=None, eager_steps=None, nbuf=None):
        self.model, self.criterion, self.optimizer, self.scaler = model, criterion, optimizer, scaler
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.use_amp, self.amp_dtype = bool(use_amp), amp_dtype
        self.fabric = getattr(model, "fabric", None)
        self.is_dp = hasattr(model, "fused_step")
        cuda = self.device.type == "cuda"
        self.native = cuda and _native.available()
        if graph is None:
            graph = os.environ.get("PDT_GRAPH", "1") == "1"
        # DataParallel drives several devices from one process: its step is captured as per-device pieces
        # (parallel.dp.GraphStep) when no loss scaling is involved; with one device it is the wrapped module's step.
        dp_ok = (not self.is_dp) or (os.environ.get("PDT_DP_GRAPH", "1") == "1" and len(getattr(model, "device_ids", [])) > 1
                                     and (scaler is None or not scaler.is_enabled()))
        self.graph_enabled = bool(graph) and self.native and dp_ok
        self.eager_left = int(os.environ.get("PDT_GRAPH_WARMUP", "3")) if eager_steps is None else int(eager_steps)
        self.stream = work_stream(self.device) if cuda else None
        self.nbuf = nbuf = int(os.environ.get("PDT_FEED_BUFFERS", "2")) if nbuf is None else int(nbuf)
        self._static = None            # [(images, target)] * nbuf, owned here, written by DeviceFeed
        self._graphs = {}
        self._pool = None
        self.launches_per_step = None  # own kernels per captured step (bench.py reports it)
        self.meter = Meter(self.device)
        self.result = torch.zeros(2, dtype=torch.float32, device=self.device)   # cross-rank mean (loss, acc1), last step
        self.consumed = [torch.cuda.Event() for _ in range(nbuf)] if cuda else None
        self.capture_error = None
        # (loss, acc1) of the last step in pinned HOST memory, written by a device-to-host copy that is part of
        # the step itself (also of the captured graph): reading ``result_host`` after a stream / event
        # synchronisation costs the training loop no extra launch between two replays.
        self.result_host = torch.zeros(2, dtype=torch.float32).pin_memory() if cuda else torch.zeros(2)

    # ------------------------------------------------------------------ static input buffers
    def ensure_static(self, images, target):
        """Staging buffers shaped like this batch (allocated once); None if the batch has another shape
        (the last, smaller batch of an epoch runs eagerly from its own tensors)."""
        if self._static is None:
            self._static = [(torch.empty(images.shape, dtype=images.dtype, device=self.device),
                             torch.empty(target.shape, dtype=target.dtype, device=self.device)) for _ in range(self.nbuf)]
        img0, tgt0 = self._static[0]
        if img0.shape != images.shape or img0.dtype != images.dtype or tgt0.shape != target.shape or tgt0.dtype != target.dtype:
            return None
        return self._static

    def _slot_of(self, images):
        if self._static is not None:
            for i, (img, _) in enumerate(self._static):
                if img.data_ptr() == images.data_ptr():
                    return i
        return None

    # ------------------------------------------------------------------ the step itself
    def update_metrics(self, pair, n, meter=None):
        """reduce_mean(loss), reduce_mean(acc1), the barrier before them and the two AverageMeter updates
        (distributed.py:253-258) as one stream-ordered launch; DataParallel and single-GPU runs skip the exchange."""
        meter = self.meter if meter is None else meter
        fab = self.fabric if not self.is_dp else None
        world = fab.world if fab is not None else 1
        if self.native and pair.is_cuda:
            comm = fab.comm if (fab is not None and fab.native and world > 1) else None
            if comm is not None:
                fab.check_stream(1)
            with _native.guard(self.device):
                _native.check(_native.lib().pdt_metrics_update(comm, pair.data_ptr(), self.result.data_ptr(), 2, 1.0 / world, 1,
                                                               meter.t.data_ptr(), float(n), _native.stream_ptr(self.device)),
                              "pdt_metrics_update")
            return self.result
        red = pair.detach().float().clone()
        if world > 1 and dist.is_initialized():
            dist.all_reduce(red, group=fab.group)
            red /= world
        self.result.copy_(red)
        meter.t[0:2] += red * n
        meter.t[2] += n
        meter.t[3:5] = red
        return self.result

    def _forward(self, images, target, training):
        with amp.autocast(enabled=self.use_amp and training, dtype=self.amp_dtype, device_type=self.device.type):
            out = self.model(images)
            loss = self.criterion(out, target)
        pair = getattr(self.criterion, "last_pair", None)
        if pair is None:
            with torch.no_grad():
                acc = (out.argmax(1) == target).float().mean()
            pair = torch.stack([loss.detach().float(), acc])
        return loss, pair

    def _body(self, images, target):
        n = images.shape[0]
        with nvtx_range("forward"):
            loss, pair = self._forward(images, target, True)
        with nvtx_range("metrics"):
            self.update_metrics(pair, n)
        self.optimizer.zero_grad()
        with nvtx_range("backward"):
            (self.scaler.scale(loss) if self.scaler is not None else loss).backward()
        with nvtx_range("optimizer"):
            if self.scaler is not None:
                self.scaler.step(self.optimizer)
                self.scaler.update()
            else:
                self.optimizer.step()
        self.result_host.copy_(self.result, non_blocking=True)
        return self.result

    def _capture_dp(self, slot):
        """DataParallel: 3 W + 1 graphs (forward / loss / backward / optimizer pieces per device)."""
        img, tgt = self._static[slot]

        def loss_fn(logits, target):
            loss = self.criterion(logits, target)
            pair = getattr(self.criterion, "last_pair", None)
            if pair is None:
                with torch.no_grad():
                    pair = torch.stack([loss.detach().float(), (logits.argmax(1) == target).float().mean()])
            return loss, pair

        try:
            for d in self.model.devices:
                torch.cuda.synchronize(d)
            def metrics_fn(pair, n):
                self.update_metrics(pair, n)
                self.result_host.copy_(self.result, non_blocking=True)

            g = self.model.graph_step(img, tgt, loss_fn, metrics_fn, self.amp_dtype if self.use_amp else None)
        except Exception as e:
            self.capture_error = repr(e)
            self.graph_enabled = False
            for d in self.model.devices:
                try:
                    torch.cuda.synchronize(d)
                except Exception:
                    pass
            print("[pdt] CUDA-graph capture of the DataParallel step failed, continuing eagerly: {!r}".format(e), file=sys.stderr)
            return None
        self.launches_per_step = g.launches
        self._graphs[slot] = g
        return g

    def _capture(self, slot):
        if self.is_dp:
            return self._capture_dp(slot)
        img, tgt = self._static[slot]
        g = torch.cuda.CUDAGraph()
        n0 = _native.launch_count
        try:
            with torch.cuda.graph(g, stream=self.stream, pool=self._pool, capture_error_mode=CAPTURE_MODE):
                self._body(img, tgt)
        except Exception as e:        # keep training: report once, continue eagerly
            self.capture_error = repr(e)
            self.graph_enabled = False
            torch.cuda.synchronize(self.device)
            print("[pdt] CUDA-graph capture of the training step failed, continuing eagerly: {!r}".format(e), file=sys.stderr)
            return None
        self.launches_per_step = _native.launch_count - n0
        if self._pool is None:
            self._pool = g.pool()
        self._graphs[slot] = g
        return g

    def step(self, images, target):
        """One optimisation step.  ``images`` / ``target`` are device tensors — ideally the runner's own staging
        buffers as handed out by ``DeviceFeed`` (zero-copy graph replay); any other tensor of the same shape is
        copied into buffer 0 first; other shapes run eagerly.  Returns the device tensor ``result`` =
        cross-rank mean (loss, top-1) of this step; nothing synchronises with the host."""
        training = self.model.training and torch.is_grad_enabled()
        if not (self.graph_enabled and training):
            return self._body(images, target)
        slot = self._slot_of(images)
        if slot is None:
            if self.ensure_static(images, target) is None:
                return self._body(images, target)
            slot = 0
            self._static[0][0].copy_(images, non_blocking=True)
            self._static[0][1].copy_(target, non_blocking=True)
        img, tgt = self._static[slot]
        if self.eager_left > 0:
            self.eager_left -= 1
            self._body(img, tgt)
        else:
            g = self._graphs.get(slot) or self._capture(slot)
            if g is None:
                self._body(img, tgt)
            else:
                with nvtx_range("step(graph)"):
                    if self.is_dp:
                        g.replay(self.optimizer)
                    else:
                        g.replay()
                self._weights_changed()      # the replayed optimizer kernels rewrote the weights behind Python's back
        self.consumed[slot].record(torch.cuda.current_stream(self.device))
        return self.result

    def _weights_changed(self):
        arenas = getattr(self.model, "arenas", None) or [getattr(self.model, "arena", None)]
        for a in arenas:
            if a is not None:
                a.weights_changed()

    @torch.no_grad()
    def eval_step(self, images, target, meter):
        """Validation iteration (never under autocast): forward, loss, top-1, reduced into ``meter``."""
        _, pair = self._forward(images, target, False)
        return self.update_metrics(pair, images.shape[0], meter)

    @property
    def graphs(self):
        return len(self._graphs)


class DeviceFeed:
    """Iterate ``(images, target)`` on the device, staged for a ``StepRunner``.

    Batches i+1 .. i+nbuf-1 are copied host -> device on a side stream into the runner's staging buffers while
    batch i computes; a copy waits for the step that last read that buffer (``runner.consumed``), the step
    waits for the copy.  Pinned host batches (``DataLoader(pin_memory=True)``) make the copy asynchronous.
    Batches of another shape than the first (the tail of an epoch) are moved with ordinary tensors."""

    def __init__(self, loader, runner):
        self.loader, self.runner = loader, runner
        self.device = runner.device
        self.stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._n = 0

    def __len__(self):
        return len(self.loader)

    def _stage(self, batch):
        images, target = batch
        target = torch.as_tensor(target)
        if self.stream is None:
            return images.to(self.device), target.to(self.device), None, None
        bufs = self.runner.ensure_static(images, target)
        with torch.cuda.stream(self.stream):
            if bufs is None:
                di = images.to(self.device, non_blocking=True)
                dt = target.to(self.device, non_blocking=True)
                slot = None
            else:
                slot = self._n % len(bufs)
                self._n += 1
                self.stream.wait_event(self.runner.consumed[slot])
                di, dt = bufs[slot]
                di.copy_(images, non_blocking=True)
                dt.copy_(target, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(self.stream)
        return di, dt, ready, (images, target)      # the host tensors stay referenced until the copy is consumed

    def __iter__(self):
        """Up to nbuf - 1 batches are staged ahead of the one being consumed (nbuf = 2: the classic double buffer;
        nbuf = 3 gives every copy two step times to land, which matters when eight ranks pull from the same host)."""
        import collections
        it = iter(self.loader)
        depth = max(1, self.runner.nbuf - 1)
        queue = collections.deque()

        def fill():
            while len(queue) < depth + 1:
                try:
                    queue.append(self._stage(next(it)))
                except StopIteration:
                    return

        fill()
        while queue:
            di, dt, ready, _keep = queue.popleft()
            fill()
            if ready is not None:
                main = torch.cuda.current_stream(self.device)
                main.wait_event(ready)
                if self.runner._slot_of(di) is None:
                    di.record_stream(main)
                    dt.record_stream(main)
            yield di, dt
