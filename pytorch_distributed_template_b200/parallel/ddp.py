"""DistributedDataParallel for one process per GPU.

Same user contract as ``torch.nn.parallel.DistributedDataParallel`` as the
reference uses it (``DDP(model, device_ids=[local_rank])``,
/root/reference/distributed.py:144): construction makes every rank's
parameters and buffers equal to rank 0's, ``forward`` re-broadcasts rank 0's
buffers (``broadcast_buffers=True``), ``backward`` leaves every rank with the
same averaged gradient (or, with the fused optimizer, the same updated
weights), the wrapped network is ``.module`` so ``state_dict()`` keys carry
the ``module.`` prefix while ``model.module.state_dict()`` does not.

What is different underneath:
* parameters, gradients and float buffers live in flat arenas in peer-mapped
  memory (``arena.py``); a "bucket" is a contiguous *segment* of the arena;
* per-parameter ``post_accumulate_grad`` hooks count a segment down; when it
  reaches zero an event is recorded and OUR kernel (``csrc/comm.cu``) is
  enqueued on a side stream, overlapping the rest of backward.  It reads the
  peers' gradient arenas directly over NVLink and folds in 1/world, the AMP
  unscale and the inf check;
* if a fused optimizer (``optim.SGD``) is bound, the second half of the
  all-reduce is replaced by "SGD on my shard, store new weights to every
  peer" at ``optimizer.step()``; otherwise the kernel is a plain in-place
  all-reduce and any torch optimizer works on the averaged ``p.grad``;
* nothing in the step touches the host, so it can be captured in a CUDA graph.

On CPU (gloo) the same bookkeeping runs with ``dist.all_reduce`` per segment —
that is the oracle the GPU path is tested against.
"""
import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import _native
from .arena import ParamArena
from .fabric import Fabric


class BufferArena:
    """All floating-point buffers (BN running stats) in one peer-mapped range."""

    def __init__(self, module, fabric):
        bufs = [b for b in module.buffers() if b.is_floating_point()]
        self.total = sum((b.numel() + 3) // 4 * 4 for b in bufs)
        self.total = max(4, (self.total + 3) // 4 * 4)
        self.buf = fabric.alloc(self.total * 4)
        self.flat = self.buf.view(0, self.total, torch.float32)
        pos = 0
        self.views = []
        with torch.no_grad():
            for b in bufs:
                if b.dtype != torch.float32:
                    continue
                v = self.flat[pos:pos + b.numel()].view(b.shape)
                v.copy_(b)
                b.data = v
                self.views.append(v)
                pos += (b.numel() + 3) // 4 * 4
        self.used = pos


class DistributedDataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, dim=0, broadcast_buffers=True,
                 process_group=None, bucket_cap_mb=25, find_unused_parameters=False, check_reduction=False,
                 gradient_as_bucket_view=True, static_graph=False, shadow_dtype=None, comm_blocks=32, step_blocks=128):
        super().__init__()
        self.module = module
        p0 = next(module.parameters())
        self.device = p0.device
        if device_ids is not None and len(device_ids) == 1 and self.device.type == "cuda":
            want = device_ids[0] if isinstance(device_ids[0], int) else torch.device(device_ids[0]).index
            assert want == self.device.index, "module must already be on device_ids[0]"
        self.process_group = process_group if process_group is not None else (
            dist.group.WORLD if dist.is_initialized() else None)
        self.broadcast_buffers = broadcast_buffers
        self.fabric = Fabric(self.device, self.process_group)
        self.world, self.rank = self.fabric.world, self.fabric.rank
        self.native = self.fabric.native
        self.comm_blocks = comm_blocks      # reduce-scatter launches share the GPU with the rest of backward
        self.step_blocks = step_blocks      # the SGD / all-gather launch at step end has it to itself
        self.arena = ParamArena(module.parameters(), self.fabric, first_bucket_mb=1.0,
                                bucket_cap_mb=float(bucket_cap_mb), shadow_dtype=shadow_dtype)
        self.arena.owner = self
        self.arena.watch_state_dict(module)
        from .syncbn import SyncBatchNorm
        for m in module.modules():
            if isinstance(m, SyncBatchNorm):
                m._pdt_fabric = self.fabric
        self.buffers_arena = BufferArena(module, self.fabric)
        self._int_buffers = [b for b in module.buffers() if not b.is_floating_point()]
        self.optimizer = None        # set by optim.SGD.bind()
        self.amp_state = None        # device tensor, set by amp.GradScaler.bind()
        self._unit_state = None
        self._comm_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._works = []
        self._fired = set()
        self._callback_queued = False
        self._launched = set()
        self.require_backward_grad_sync = True
        self._sync_module_states()
        self._register_hooks()

    # ------------------------------------------------------------------ construction sync
    def _sync_module_states(self):
        """Rank 0's parameters and buffers become everybody's (torch DDP ctor, X1)."""
        if self.world == 1:
            return
        if self.native:
            lib, s = _native.lib(), _native.stream_ptr(self.device)
            self.fabric.barrier()
            _native.check(lib.pdt_bcast_pull(self.fabric.comm, self.arena.param_buf.peers(), 0, self.arena.total * 4, 0, 1,
                                            64, s), "pdt_bcast_pull(params)")
            self._bcast_buffers_native()
            self.arena.refresh_shadow()
        else:
            dist.broadcast(self.arena.flat_params, src=0, group=self.process_group)
            dist.broadcast(self.buffers_arena.flat, src=0, group=self.process_group)
        for b in self._int_buffers:
            dist.broadcast(b, src=0, group=self.process_group)

    def _bcast_buffers_native(self):
        ba = self.buffers_arena
        if ba.used == 0:
            return
        self.fabric.check_stream(1)
        _native.check(_native.lib().pdt_bcast_pull(self.fabric.comm, ba.buf.peers(), 0, ba.total * 4, 0, 1, 4,
                                                  _native.stream_ptr(self.device)), "pdt_bcast_pull(buffers)")

    def _sync_buffers(self):
        """broadcast_buffers=True: rank 0's BN statistics overwrite the others' before a
        training forward (torch/nn/parallel/distributed.py:2214-2257, X2)."""
        if self.world == 1 or not self.broadcast_buffers or self.buffers_arena.used == 0:
            return
        if self.native:
            self._bcast_buffers_native()
        else:
            dist.broadcast(self.buffers_arena.flat, src=0, group=self.process_group)

    # ------------------------------------------------------------------ hooks
    def _register_hooks(self):
        for seg in self.arena.segments:
            seg.pending = len(seg.params)
        for p in self.arena.params:
            p.register_post_accumulate_grad_hook(self._make_hook(p))

    def _make_hook(self, p):
        arena = self.arena
        seg = arena.segment_of[id(p)]
        view = arena.grad_view(p)
        vptr = view.data_ptr()

        def hook(param):
            if not self.require_backward_grad_sync:
                return
            g = param.grad
            if g is None:
                return
            if g.data_ptr() != vptr:
                view.copy_(g)
                param.grad = view
            if not self._callback_queued:
                self._callback_queued = True
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            if id(param) in self._fired:
                return
            self._fired.add(id(param))
            seg.pending -= 1
            if seg.pending == 0:
                self._launch_segment(seg)
        return hook

    def _amp_ptr(self):
        if self.amp_state is not None:
            return self.amp_state.data_ptr()
        if self._unit_state is None:   # scale 1.0, never reports inf to anybody
            self._unit_state = torch.zeros(8, dtype=torch.float32, device=self.device)
            self._unit_state[0] = 1.0
        return self._unit_state.data_ptr()

    def _launch_segment(self, seg):
        self._launched.add(seg.index)
        if self.native:
            main = torch.cuda.current_stream(self.device)
            self._comm_stream.wait_stream(main)
            with torch.cuda.stream(self._comm_stream):
                s = self._comm_stream.cuda_stream
                lib = _native.lib()
                if self.optimizer is not None:
                    _native.check(lib.pdt_reduce_scatter_f32(
                        self.fabric.comm, self.arena.grad_buf.peers(), self.arena.grad_buf.mc(self.fabric.use_nvls),
                        seg.start, seg.length, self._amp_ptr(), 1 if self.world > 1 else 0, self.comm_blocks, s),
                        "pdt_reduce_scatter_f32")
                elif self.world > 1:
                    _native.check(lib.pdt_allreduce_f32(
                        self.fabric.comm, self.arena.grad_buf.peers(), seg.start, seg.length, 1.0 / self.world,
                        self.comm_blocks, s), "pdt_allreduce_f32")
        elif self.world > 1:
            flat = self.arena.flat_grads[seg.start:seg.end]
            self._works.append((dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.process_group, async_op=True), flat))

    def _finalize_backward(self):
        self._callback_queued = False
        # parameters that produced no gradient this iteration contribute zeros
        for seg in self.arena.segments:
            if seg.index in self._launched:
                continue
            for p in seg.params:
                if id(p) not in self._fired:
                    self.arena.grad_view(p).zero_()
                    p.grad = self.arena.grad_view(p)
            self._launch_segment(seg)
        for work, flat in self._works:
            work.wait()
            flat.div_(self.world)
        self._works = []
        if self.native and self.optimizer is None:
            torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
        for seg in self.arena.segments:
            seg.pending = len(seg.params)
        self._fired.clear()
        self._launched.clear()

    def join_comm(self):
        """Make the current stream wait for everything queued on the communication stream."""
        if self._comm_stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)

    # ------------------------------------------------------------------ forward
    def forward(self, *inputs, **kwargs):
        if self.training and torch.is_grad_enabled():
            self._sync_buffers()
        return self.module(*inputs, **kwargs)

    def no_sync(self):
        ddp = self

        class _NoSync:
            def __enter__(self_inner):
                self_inner.old = ddp.require_backward_grad_sync
                ddp.require_backward_grad_sync = False

            def __exit__(self_inner, *a):
                ddp.require_backward_grad_sync = self_inner.old
        return _NoSync()
