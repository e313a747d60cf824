"""NVLink fabric bring-up: peer-mapped arenas, flag pads, pointer tables.

One ``Fabric`` per process group.  ``alloc(nbytes)`` returns a byte tensor that
every rank has mapped into its address space (CUDA VMM allocation exchanged
through ``torch.distributed._symmetric_memory``; PyTorch stays the allocator
and rendezvous layer, the kernels that use the pointers are ours).  With one
rank, or on CPU, the same API hands out ordinary tensors so the rest of the
stack has a single code path.

Replaces ProcessGroupNCCL's role for the hot path; NCCL remains
only for process-group bootstrap and as the baseline.
"""
import os

import torch
import torch.distributed as dist

from ..ops import _native


class SymBuffer:
    """A byte range present on every rank; ``ptrs[r]`` is rank r's copy as seen from here."""

    def __init__(self, local, ptrs, multicast_ptr=0, handle=None):
        self.local = local            # uint8 tensor (this rank's copy)
        self.ptrs = list(ptrs)        # int device pointers, len == world
        self.multicast_ptr = multicast_ptr
        self._handle = handle         # keeps the mapping alive

    def view(self, offset, numel, dtype):
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        return self.local[offset:offset + nbytes].view(dtype)

    def peers(self, offset=0):
        return _native.make_peers([p + offset for p in self.ptrs])

    def mc(self, use=True):
        """The buffer through its NVLS multicast mapping (an int device pointer), or None."""
        return self.multicast_ptr if (use and self.multicast_ptr) else None


class Fabric:
    def __init__(self, device, group=None):
        self.device = torch.device(device)
        solo = group == "solo"
        self.group = None if solo else (group if group is not None else (dist.group.WORLD if dist.is_initialized() else None))
        self.rank = dist.get_rank(self.group) if self.group is not None else 0
        self.world = dist.get_world_size(self.group) if self.group is not None else 1
        self.native = self.device.type == "cuda" and _native.available()
        self.symmetric = False
        self._buffers = []
        self.pad = None
        self.comm = None
        # NVLS data plane (multimem.ld_reduce / multimem.st on the multicast mapping) when the allocation got one;
        # PDT_NVLS=0 forces the peer load/store plane.
        self.use_nvls = os.environ.get("PDT_NVLS", "1") == "1"
        self._channel_stream = {}
        if self.native:
            if self.world > _native.MAX_RANKS:
                raise RuntimeError("native fabric supports world sizes 1..{} (got {})".format(_native.MAX_RANKS, self.world))
            self.pad = self.alloc(_native.lib().pdt_comm_pad_bytes(), zero=True)
            self.comm = _native.make_comm(self.rank, self.world, self.pad.ptrs, self.pad.mc(self.use_nvls))

    # -- allocation ---------------------------------------------------------------------
    def alloc(self, nbytes, zero=True):
        nbytes = (int(nbytes) + 255) // 256 * 256
        if self.device.type != "cuda" or self.world == 1:
            t = torch.zeros(nbytes, dtype=torch.uint8, device=self.device) if zero else \
                torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            buf = SymBuffer(t, [t.data_ptr()] if self.device.type == "cuda" else [0])
        else:
            buf = self._alloc_symmetric(nbytes, zero)
        self._buffers.append(buf)
        return buf

    def _alloc_symmetric(self, nbytes, zero):
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(nbytes, dtype=torch.uint8, device=self.device)
        if zero:
            t.zero_()
        torch.cuda.synchronize(self.device)
        handle = symm.rendezvous(t, group=self.group)
        ptrs = [int(p) for p in handle.buffer_ptrs]
        mc = int(getattr(handle, "multicast_ptr", 0) or 0)
        self.symmetric = True
        # every rank must have zeroed before anybody signals into the buffer
        dist.barrier(group=self.group)
        return SymBuffer(t, ptrs, mc, handle)

    # -- collectives used outside the gradient path ---------------------------------------
    def check_stream(self, channel, stream_ptr=None):
        """The flag protocol is self-resetting: every user of one channel must enqueue on ONE stream, in the same
        order on every rank (SyncBN exchanges, metric reductions and buffer broadcasts share channel 1).  A second
        stream on the same channel could interleave two exchanges on the same flag row and deadlock until the
        spin budget traps; refuse it here instead."""
        if self.world == 1:
            return
        s = _native.stream_ptr(self.device) if stream_ptr is None else stream_ptr
        first = self._channel_stream.setdefault(channel, s)
        if first != s:
            raise RuntimeError("fabric channel {} is bound to stream {:#x} but was used from stream {:#x}: all users of a "
                               "channel must share one stream (use another channel for a side stream)".format(channel, first, s))

    def barrier(self, channel=1):
        self.check_stream(channel)
        if self.native and self.world > 1:
            _native.check(_native.lib().pdt_barrier(self.comm, channel, _native.stream_ptr(self.device)), "pdt_barrier")
        elif self.world > 1:
            dist.barrier(group=self.group)

    def allreduce_small_(self, t, mean=False, channel=1):
        """In-place sum (or mean) of a small fp32 tensor across ranks, stream-ordered, no host sync."""
        if self.world == 1:
            return t
        if self.native and t.is_cuda and t.dtype == torch.float32 and t.numel() <= 8192 and t.is_contiguous():
            self.check_stream(channel)
            scale = 1.0 / self.world if mean else 1.0
            _native.check(_native.lib().pdt_allreduce_small_f32(
                self.comm, t.data_ptr(), t.data_ptr(), t.numel(), scale, channel, _native.stream_ptr(self.device)),
                "pdt_allreduce_small_f32")
            return t
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        if mean:
            t /= self.world
        return t

    def describe(self):
        mc = bool(self.pad is not None and self.pad.multicast_ptr)
        return {"rank": self.rank, "world": self.world, "native": self.native, "symmetric": self.symmetric,
                "multicast": mc, "data_plane": "nvls" if (mc and self.use_nvls and self.world > 1) else "p2p"}
