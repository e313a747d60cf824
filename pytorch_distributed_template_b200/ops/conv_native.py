"""tcgen05 implicit-GEMM convolution (``csrc/conv.cu``) behind ``ops.conv2d`` / ``ops.linear``.

Native on the training path:
  * forward of every conv with C_in % 64 == 0 (IM2COL-TMA fprop kernel);
  * data gradient: stride 1 = the fprop kernel on dy with the flipped/transposed filter;
    stride 2 = one fprop launch per output parity class (sub-filters of <= 2x2 taps), each
    writing its own interleaved positions of dx;
  * weight gradient: split-K kernel over the pixel dimension, fp32 result;
  * the 7x7/2 stem (C_in = 3) through a 2x2 space-to-depth transform (``stem_*`` below);
  * nn.Linear as a 1x1 convolution with a bias epilogue.
``PDT_CONV=torch`` sends everything through ``aten::convolution*`` instead (A/B runs, debugging).
"""
import os

import torch
import torch.nn.functional as F

from . import _native
from . import conv as _conv

_vp, _i, _ll = _native._vp, _native._i, _native.ctypes.c_longlong
_native.register("pdt_conv_fprop", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp])
_native.register("pdt_conv3x3_wgrad_halo_supported", [_i, _i, _i, _i, _i, _i])
_native.register("pdt_conv_wgrad", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp])
_native.register("pdt_conv_fprop_ex", [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _ll, _ll, _ll, _i, _i, _i, _i,
                                       _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp])
_native.register("pdt_conv_fprop_ex_bnstats", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _ll, _ll, _ll, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i,
                                               _i, _vp, _native._PC, _vp, _vp, _native._f, _native._f, _vp, _vp, _vp])
_native.register("pdt_conv_fprop_add", [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp])
_native.register("pdt_conv3x3_halo_add", [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _native._PC, _vp, _vp, _native._f, _native._f, _vp, _vp, _vp])
_native.register("pdt_conv_wgrad_ex", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _ll, _ll, _ll, _i, _i, _i, _i,
                                       _i, _i, _i, _i, _i, _i, _i, _vp])
_native.register("pdt_stem_s2d", [_vp, _vp, _i, _i, _i, _i, _i, _vp])
_native.register("pdt_stem_s2d_u8", [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp])
_native.register("pdt_stem_conv_supported", [_i, _i, _i, _i, _i])
_native.register("pdt_stem_weight_s2d", [_vp, _vp, _i, _i, _ll, _ll, _ll, _ll, _vp])
_native.register("pdt_stem_weight_grad_unpack", [_vp, _vp, _i, _ll, _ll, _ll, _ll, _vp])
_native.register("pdt_stem_wgrad_supported", [_i, _i, _i, _i, _i])
_native.register("pdt_stem_wgrad", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp])
_native.register("pdt_stem_conv", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _native._PC, _vp, _vp, _native._f, _native._f, _vp, _vp, _vp])
_native.register("pdt_dgrad_weight", [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp])
_native.register("pdt_dgrad_weight_batched", [_vp, _i, _i, _vp])
_native.register("pdt_dgrad_tile", [])
_native.register("pdt_conv_fprop_bnstats", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _native._PC, _vp, _vp,
                                            _native._f, _native._f, _vp, _vp, _vp])

_native.register("pdt_conv3x3_halo_supported", [_i, _i, _i, _i, _i, _i])
_native.register("pdt_conv3x3_halo", [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _native._f, _native._f, _vp, _vp, _vp])

CL = torch.channels_last


def _halo_ok(x, w_lp, stride, pad):
    """Large-map 3x3/1/1 layers go to the patch-reuse kernel (csrc/conv_halo.cu); PDT_HALO=0 disables it."""
    K, C, R, S = w_lp.shape
    if R != 3 or S != 3 or stride != 1 or pad != 1 or os.environ.get("PDT_HALO", "1") != "1":
        return False
    if x.dtype == torch.float32 and os.environ.get("PDT_HALO_TF32", "1") != "1":      # fp32 runs as kind::tf32, streamed filter taps
        return False
    N, _, H, W = x.shape
    return bool(_native.lib().pdt_conv3x3_halo_supported(_native.dtype_code(x.dtype), N, H, W, C, K))


def _halo(x, w_lp, y, scratch=None, rm=None, rv=None, momentum=0.1, eps=1e-5, mean=None, invstd=None, addend=None, comm=None):
    N, C, H, W = x.shape
    K = w_lp.shape[0]
    _native.check(_native.lib().pdt_conv3x3_halo_add(
        x.data_ptr(), w_lp.data_ptr(), y.data_ptr(), _native.ptr(addend), _native.dtype_code(x.dtype), N, H, W, C, K,
        _native.ptr(scratch), comm, _native.ptr(rm), _native.ptr(rv), momentum, eps, _native.ptr(mean), _native.ptr(invstd),
        _native.stream_ptr(x.device)), "pdt_conv3x3_halo")
    return y


def enabled():
    return os.environ.get("PDT_CONV", "native") == "native"


def tf32_enabled():
    """fp32 convolutions run on tcgen05 kind::tf32 exactly when torch would let cuDNN use TF32."""
    return torch.backends.cudnn.allow_tf32 and os.environ.get("PDT_TF32", "native") == "native"


def _row(dtype):
    return 32 if dtype == torch.float32 else 64     # channels per 128-byte K row


def supported(x, weight, stride, padding):
    K, C, R, S = weight.shape
    ok_dtype = x.dtype in (torch.bfloat16, torch.float16) or (x.dtype == torch.float32 and tf32_enabled())
    return (x.is_cuda and ok_dtype and C % _row(x.dtype) == 0 and K % 8 == 0
            and stride[0] == stride[1] and padding[0] == padding[1] and R == S and enabled())


# ------------------------------------------------------------------------------------------ kernels
def fprop(x, w_lp, stride, pad, addend=None):
    """x: [N,C,H,W] channels_last, w_lp: [K,C,R,S] channels_last -> y [N,K,P,Q] channels_last.
    ``addend`` (shaped and laid out like y) is summed into the result in the kernel's epilogue."""
    N, C, H, W = x.shape
    K, _, R, S = w_lp.shape
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    y = torch.empty((N, K, P, Q), dtype=x.dtype, device=x.device, memory_format=CL)
    if _halo_ok(x, w_lp, stride, pad):
        return _halo(x, w_lp, y, addend=addend)
    if addend is not None:
        _native.check(_native.lib().pdt_conv_fprop_add(x.data_ptr(), w_lp.data_ptr(), y.data_ptr(), addend.data_ptr(),
                                                       _native.dtype_code(x.dtype), N, H, W, C, K, R, S, stride, pad,
                                                       _native.stream_ptr(x.device)), "pdt_conv_fprop_add")
        return y
    _native.check(_native.lib().pdt_conv_fprop(x.data_ptr(), w_lp.data_ptr(), y.data_ptr(), _native.dtype_code(x.dtype),
                                               N, H, W, C, K, R, S, stride, pad, _native.stream_ptr(x.device)),
                  "pdt_conv_fprop")
    return y


def fprop_bnstats(x, w_lp, stride, pad, running_mean, running_var, momentum, eps, comm=None):
    """Convolution whose epilogue also yields the BatchNorm batch statistics of its output; with ``comm`` (a native
    CommDesc, world > 1) the statistics are SyncBatchNorm's: exchanged with the peers inside the kernel."""
    from .bn import _bn_scratch
    N, C, H, W = x.shape
    K, _, R, S = w_lp.shape
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    y = torch.empty((N, K, P, Q), dtype=x.dtype, device=x.device, memory_format=CL)
    mean = torch.empty(K, dtype=torch.float32, device=x.device)
    invstd = torch.empty_like(mean)
    if _halo_ok(x, w_lp, stride, pad):
        _halo(x, w_lp, y, _bn_scratch(x.device, K), running_mean, running_var, momentum, eps, mean, invstd, comm=comm)
        return y, mean, invstd
    _native.check(_native.lib().pdt_conv_fprop_bnstats(
        x.data_ptr(), w_lp.data_ptr(), y.data_ptr(), _native.dtype_code(x.dtype), N, H, W, C, K, R, S, stride, pad,
        _bn_scratch(x.device, K).data_ptr(), comm, _native.ptr(running_mean), _native.ptr(running_var), momentum, eps,
        mean.data_ptr(), invstd.data_ptr(), _native.stream_ptr(x.device)), "pdt_conv_fprop_bnstats")
    return y, mean, invstd


def fprop_ex(x_ptr, w, y, bias, dtype, N, H, W, C, Cx, sw, sh, sn, K, R, S, stride, pads, P, Q, OH, OW, os_, oa, ob, device,
             dil_w=1, addend=None):
    _native.check(_native.lib().pdt_conv_fprop_ex(
        x_ptr, w.data_ptr(), y.data_ptr(), _native.ptr(bias), _native.dtype_code(dtype), N, H, W, C, Cx, sw, sh, sn,
        K, R, S, stride, pads[0], pads[1], pads[2], pads[3], P, Q, OH, OW, os_, oa, ob, dil_w, _native.ptr(addend),
        _native.stream_ptr(device)), "pdt_conv_fprop_ex")
    return y


def direct_grad(param):
    """Slot of ``param`` in its gradient arena when backward may write there directly (see ParamArena)."""
    arena = getattr(param, "_pdt_arena", None)
    return None if arena is None else arena.direct_grad(param)


def wgrad(x, dy, weight_shape, stride, pad, out=None):
    """fp32 dW [K,C,R,S] (channels_last, i.e. KRSC) from x [N,C,H,W] and dy [N,K,P,Q], both channels_last.
    ``out``: an already-zeroed destination the kernel accumulates into (the gradient arena)."""
    N, C, H, W = x.shape
    K, _, R, S = weight_shape
    _native.ensure_deterministic(x.device)
    if out is None:
        out = torch.empty((K, C, R, S), dtype=torch.float32, device=x.device, memory_format=CL)
        out.zero_()
    _native.check(_native.lib().pdt_conv_wgrad(x.data_ptr(), dy.data_ptr(), out.data_ptr(), _native.dtype_code(x.dtype),
                                               N, H, W, C, K, R, S, stride, pad, _native.stream_ptr(x.device)),
                  "pdt_conv_wgrad")
    return out


def wgrad_supported(C, K, dtype=torch.bfloat16):
    # fp32: both operands are MN-major kind::tf32 tiles (32-byte-atom swizzle); PDT_WGRAD_TF32=0 routes
    # fp32 weight gradients through aten instead.
    if dtype == torch.float32 and os.environ.get("PDT_WGRAD_TF32", "1") != "1":
        return False
    return C % _row(dtype) == 0 and K % 8 == 0 and os.environ.get("PDT_WGRAD", "native") == "native"


def _parity_taps(size_in, ksize, pad, a):
    """Taps of the stride-2 data gradient that reach input rows h = 2i + a.
    Returns (filter rows in increasing dy-offset order, lowest dy offset)."""
    rs = [r for r in range(ksize) if (a + pad - r) % 2 == 0]
    if not rs:
        return [], 0
    rs = sorted(rs, reverse=True)                 # decreasing r == increasing offset into dy
    return rs, (a + pad - rs[0]) // 2


_index_cache = {}
_tap_cache = {}


def _index_tensor(values, device):
    """Small index tensors are created once (an H2D copy is not allowed inside a graph capture)."""
    key = (tuple(values), device)
    t = _index_cache.get(key)
    if t is None:
        t = torch.tensor(list(values), device=device)
        _index_cache[key] = t
    return t


class _DgradJob(_native.ctypes.Structure):
    _fields_ = [("src", _native.ctypes.c_void_p), ("dst", _native.ctypes.c_void_p),
                ("K", _i), ("C", _i), ("R", _i), ("S", _i), ("Th", _i), ("Tw", _i),
                ("rs", _i * 8), ("ss", _i * 8), ("tile_begin", _i), ("pad_", _i)]


class _DgradPlan:
    """Per device: every (filter, tap subset) the data-gradient kernels asked for, with a persistent output
    buffer each.  ``prepare()`` (called at the start of a training forward, i.e. after the optimizer has
    written the step's weights) re-lays all of them out in ONE launch; until the next ``prepare`` the
    backward pass picks the buffers up without launching anything."""

    def __init__(self, device):
        self.device = device
        self.jobs = {}            # (src_ptr, K, C, R, S, rs, ss) -> output tensor
        self.order = []
        self.table = None         # device copy of the descriptor array
        self.table_jobs = 0
        self.total_tiles = 0
        self.fresh = False
        self.version = -1          # ParamArena.weights_version the buffers were made from

    def lookup(self, w_lp, rs, ss, version=None):
        K, C, R, S = w_lp.shape
        key = (w_lp.data_ptr(), K, C, R, S, tuple(rs), tuple(ss))
        out = self.jobs.get(key)
        if out is None:
            out = torch.empty((C, K, len(rs), len(ss)), dtype=w_lp.dtype, device=w_lp.device, memory_format=CL)
            self.jobs[key] = out
            self.order.append(key)
            return out, False
        return out, self.fresh and self.table_jobs == len(self.order) and (version is None or version == self.version)

    def prepare(self, version=None):
        self.fresh = False
        self.version = version
        if not self.order or os.environ.get("PDT_DGRAD_BATCH", "1") != "1":
            return
        if self.table is None or self.table_jobs != len(self.order):
            if torch.cuda.is_current_stream_capturing():
                return                                   # the descriptor upload is a host-to-device copy
            arr = (_DgradJob * len(self.order))()
            tiles = 0
            te = int(_native.lib().pdt_dgrad_tile())        # tile edge of the batched re-layout kernel
            for n, key in enumerate(self.order):
                ptr, K, C, R, S, rs, ss = key
                j = arr[n]
                j.src, j.dst = ptr, self.jobs[key].data_ptr()
                j.K, j.C, j.R, j.S, j.Th, j.Tw = K, C, R, S, len(rs), len(ss)
                for i, v in enumerate(rs):
                    j.rs[i] = v
                for i, v in enumerate(ss):
                    j.ss[i] = v
                j.tile_begin = tiles
                tiles += ((K + te - 1) // te) * ((C + te - 1) // te) * len(rs) * len(ss)
            host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            self.table = host.to(self.device)
            self.table_jobs = len(self.order)
            self.total_tiles = tiles
        with _native.guard(self.device):
            _native.check(_native.lib().pdt_dgrad_weight_batched(self.table.data_ptr(), self.table_jobs, self.total_tiles,
                                                                 _native.stream_ptr(self.device)), "pdt_dgrad_weight_batched")
        self.fresh = True


def prepare_dgrad_weights(arena):
    """Call once per training forward (after the weights of the step are final).  The plan lives on the
    parameter arena whose shadow weights it reads, and dies with it."""
    plan = getattr(arena, "_dgrad_plan", None)
    if plan is not None:
        plan.prepare(getattr(arena, "weights_version", None))


def dgrad_weight(w_lp, rs, ss, stable=None):
    """[C, K, Th, Tw] channels_last (memory [C][Th][Tw][K]) filter of the data-gradient convolution:
    out[c, k, i, j] = w_lp[k, c, rs[i], ss[j]] in one launch (w_lp: [K, C, R, S] channels_last, 16-bit)."""
    K, C, R, S = w_lp.shape
    Th, Tw = len(rs), len(ss)
    if w_lp.element_size() != 2 or not w_lp.is_cuda or Th > 8 or Tw > 8:
        ir, is_ = _index_tensor(rs, w_lp.device), _index_tensor(ss, w_lp.device)
        return w_lp.index_select(2, ir).index_select(3, is_).transpose(0, 1).contiguous(memory_format=CL)
    if stable is not None and stable is not False:   # an arena: `w_lp` keeps its address across steps, cache the layout
        plan = getattr(stable, "_dgrad_plan", None)
        if plan is None:
            plan = stable._dgrad_plan = _DgradPlan(w_lp.device)
        # a backward whose forward did not call prepare() after the last weight update (eval-mode forward with
        # gradients, a retained graph) sees a version mismatch and re-makes the layout below
        out, ready = plan.lookup(w_lp, rs, ss, getattr(stable, "weights_version", None))
        if ready:
            return out
    else:
        out = torch.empty((C, K, Th, Tw), dtype=w_lp.dtype, device=w_lp.device, memory_format=CL)
    key = (tuple(rs), tuple(ss))
    taps = _tap_cache.get(key)
    if taps is None:
        taps = ((_native.ctypes.c_int * Th)(*rs), (_native.ctypes.c_int * Tw)(*ss))
        _tap_cache[key] = taps
    _native.check(_native.lib().pdt_dgrad_weight(w_lp.data_ptr(), out.data_ptr(), K, C, R, S, Th, Tw,
                                                 _native.ctypes.addressof(taps[0]), _native.ctypes.addressof(taps[1]),
                                                 _native.stream_ptr(w_lp.device)), "pdt_dgrad_weight")
    return out


def dgrad_stride2(dy, w_lp, x_shape, pad, stable_w=None, addend=None):
    """dx for a stride-2 convolution: one fprop launch per parity class of the input grid.  ``addend`` (like dx)
    is summed in by the launches' epilogues when the four classes cover every pixel, else by one add afterwards."""
    N, C, H, W = x_shape
    K, _, R, S = w_lp.shape
    _, _, P, Q = dy.shape
    plan = []
    for a in (0, 1):
        rs, dh = _parity_taps(H, R, pad, a)
        for b in (0, 1):
            ss, dw_ = _parity_taps(W, S, pad, b)
            if rs and ss and (dh > 0 or dw_ > 0):
                return None                        # would need negative padding: not produced by ResNet shapes
            plan.append((a, b, rs, ss, dh, dw_))
    full = all(p[2] and p[3] for p in plan)
    dx = torch.empty((N, C, H, W), dtype=dy.dtype, device=dy.device, memory_format=CL)
    if not full:
        dx.zero_()
    es = dy.element_size()
    sw, sh, sn = K * es, Q * K * es, P * Q * K * es
    for a, b, rs, ss, dh, dw_ in plan:
        if not rs or not ss:
            continue
        Hs, Ws = (H - a + 1) // 2, (W - b + 1) // 2
        if Hs <= 0 or Ws <= 0:
            continue
        w_sub = dgrad_weight(w_lp, rs, ss, stable_w)
        Th, Tw = len(rs), len(ss)
        pads = (-dh, -dw_, Hs + dh + Th - 1 - P, Ws + dw_ + Tw - 1 - Q)
        if pads[2] < 0 or pads[3] < 0:
            return None
        fprop_ex(dy.data_ptr(), w_sub, dx, None, dy.dtype, N, P, Q, K, K, sw, sh, sn, C, Th, Tw, 1, pads, Hs, Ws,
                 H, W, 2, a, b, dy.device, addend=addend if full else None)
    if addend is not None and not full:
        dx += addend
    return dx


# ------------------------------------------------------------------------------------------ autograd
class GradLink:
    """Hands the shortcut gradient of a residual block to the data-gradient kernel of the block's first convolution.

    In ``out = relu(bn2(conv2(relu(bn1(conv1(x))))) + shortcut(x))`` the gradient of x is the sum of conv1's data
    gradient and the shortcut's gradient; autograd would add the two with an elementwise kernel (three passes over
    the block's largest tensor, eight times per ResNet-18 backward).  Instead the PRODUCER of the shortcut gradient
    (the block's last BatchNorm when the shortcut is the identity, the projection convolution otherwise) deposits
    its result here and reports "no gradient" for x, and the CONSUMER (conv1's backward) passes the deposit to its
    kernel as an epilogue addend.  Ordering: with an identity shortcut the producer precedes the consumer by data
    dependency; a projection convolution additionally takes the consumer's ``token`` output as an input, which
    makes the autograd engine run it first.

    Measured on B200 (ResNet-18, fp16, 1200 images): the eight ATen adds cost 0.92 ms, but the epilogue addend —
    each lane pulling its own 64-byte row pieces next to the accumulator read-out — slows the data-gradient kernels
    by 2.2 ms (k_conv3x3_halo<64> +0.70, k_conv_igemm<64> parity launches +0.65, <128> +0.34, <256> +0.31,
    halo<128> +0.14): step 25.1 -> 25.9 ms.  The link is therefore OFF unless PDT_RESADD_FUSION=1; making it pay
    needs the addend tile staged through shared memory by TMA like the output tile is."""
    __slots__ = ("grad", "claimed", "token")

    def __init__(self):
        self.grad, self.claimed, self.token = None, False, None


def link_enabled(x):
    return (x.is_cuda and x.requires_grad and torch.is_grad_enabled() and enabled()
            and os.environ.get("PDT_RESADD_FUSION", "0") == "1")


class _ConvNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, w_lp, stride, padding, bnpack=None, consume=None, produce=None, token=None):
        with _native.guard(x.device):
            outs = _ConvNative._forward_impl(ctx, x, weight, w_lp, stride, padding, bnpack)
        ctx.consume, ctx.produce = consume, produce
        if consume is None:
            return outs
        tok = x.new_empty(0)           # differentiable dummy output: a producer that takes it runs before this node
        return (outs + (tok,)) if isinstance(outs, tuple) else (outs, tok)

    @staticmethod
    def _forward_impl(ctx, x, weight, w_lp, stride, padding, bnpack=None):
        ctx.save_for_backward(x, w_lp)
        ctx.geom = (stride, padding)
        ctx.weight = weight
        ctx.set_materialize_grads(False)      # no zero tensors for the unused (mean, invstd) output gradients
        # w_lp is the arena's shadow of `weight` (fixed address, rewritten in place by the optimizer kernel): its
        # data-gradient layouts may be cached in the per-device plan.  Anything else is a temporary.
        arena = getattr(weight, "_pdt_arena", None)
        sv = arena.shadow_view(weight) if arena is not None and getattr(arena, "flat_shadow", None) is not None else None
        ctx.stable_w = arena if (sv is not None and sv.data_ptr() == w_lp.data_ptr()) else None
        if bnpack is None:
            return fprop(x, w_lp, stride[0], padding[0])
        y, mean, invstd = fprop_bnstats(x, w_lp, stride[0], padding[0], *bnpack)      # bnpack may carry a 5th item: comm
        ctx.mark_non_differentiable(mean, invstd)
        return y, mean, invstd

    @staticmethod
    def backward(ctx, dy, *unused_stat_grads):
        with _native.guard(dy.device):
            return _ConvNative._backward_impl(ctx, dy) + (None, None, None)

    @staticmethod
    def _backward_impl(ctx, dy):
        x, w_lp = ctx.saved_tensors
        stride, padding = ctx.geom
        dy = dy.contiguous(memory_format=CL)
        K, C, R, S = w_lp.shape
        need_dx = ctx.needs_input_grad[0]
        dx = None
        add = None
        if ctx.consume is not None:        # the shortcut gradient deposited by this block's producer (GradLink)
            add, ctx.consume.grad = ctx.consume.grad, None
            if add is not None and (not need_dx or add.shape != x.shape or add.dtype != dy.dtype):
                raise RuntimeError("GradLink: deposited shortcut gradient does not match the consumer's input")
            if add is not None and not add.is_contiguous(memory_format=CL):
                add = add.contiguous(memory_format=CL)
        if need_dx and K % 8 == 0 and C % 8 == 0:
            if stride[0] == 1 and R - 1 - padding[0] >= 0:
                # dx = conv(dy, flip(w)^T, pad' = R-1-pad): the same kernel with the roles of C and K swapped
                w_t = dgrad_weight(w_lp, list(range(R - 1, -1, -1)), list(range(S - 1, -1, -1)), ctx.stable_w)
                dx = fprop(dy, w_t, 1, R - 1 - padding[0], addend=add)
                add = None
            elif stride[0] == 2:
                dx = dgrad_stride2(dy, w_lp, x.shape, padding[0], ctx.stable_w, addend=add)
                if dx is not None:
                    add = None
        native_dx = dx is not None
        native_dw = wgrad_supported(C, K, x.dtype)
        dw = wgrad(x, dy, w_lp.shape, stride[0], padding[0], out=direct_grad(ctx.weight)) if native_dw else None
        if (need_dx and not native_dx) or not native_dw:
            dx2, dw2, _ = torch.ops.aten.convolution_backward(
                dy, x, w_lp, None, stride, padding, (1, 1), False, (0, 0), 1,
                (need_dx and not native_dx, not native_dw, False))
            if not native_dx:
                dx = dx2
            if not native_dw:
                dw = dw2.float()
        if add is not None:                 # a data-gradient path without the epilogue addend: add here
            dx = dx + add
        if ctx.produce is not None and dx is not None:
            ctx.produce.grad, dx = dx, None     # the consumer (same input x) returns the sum
        return dx, dw, None, None, None, None


def conv_native(x, weight, w_lp, stride, padding, bnpack=None, consume=None, produce=None):
    """bnpack = (running_mean, running_var, momentum, eps): also return the batch statistics of the
    output, computed in the convolution's epilogue -> (y, mean, invstd).
    consume / produce: ``GradLink`` roles of this convolution in a residual block (see GradLink)."""
    if not supported(x, weight, stride, padding):
        return None
    if not w_lp.is_contiguous(memory_format=CL):
        w_lp = w_lp.contiguous(memory_format=CL)
    if bnpack is not None and (weight.shape[0] > 512 or os.environ.get("PDT_BNSTATS_FUSION", "1") != "1"):
        bnpack = None
    token = None
    if produce is not None:
        token = produce.token if (produce.claimed and produce.token is not None and x.requires_grad) else None
        if token is None:
            produce = None               # no consumer behind this link: behave like an ordinary convolution
    if consume is not None and not x.requires_grad:
        consume = None
    out = _ConvNative.apply(x, weight, w_lp, tuple(stride), tuple(padding), bnpack, consume, produce, token)
    if consume is not None:
        consume.claimed, consume.token = True, out[-1]
        out = out[:-1]
        if len(out) == 1:
            out = out[0]
    return out      # y or (y, mean, invstd)


# ------------------------------------------------------------------------------------------ stem (7x7/2, C_in = 3)
def _stem_weight_native(t):
    return (t.is_cuda and _native.available() and t.dtype in (torch.float16, torch.bfloat16, torch.float32)
            and os.environ.get("PDT_STEM_WEIGHT", "native") == "native")


def stem_weight_s2d(w_lp):
    """[K,3,7,7] -> [K,4,4,16] (k, r', s', (dy,dx,c) padded to 16): filter of the equivalent 4x4/1 conv."""
    K = w_lp.shape[0]
    if _stem_weight_native(w_lp) and tuple(w_lp.shape[1:]) == (3, 7, 7):       # one launch (csrc/misc.cu: k_stem_weight_s2d)
        out = torch.empty((K, 4, 4, 16), dtype=w_lp.dtype, device=w_lp.device)
        sk, sc, sr, ss = w_lp.stride()
        _native.check(_native.lib().pdt_stem_weight_s2d(w_lp.data_ptr(), out.data_ptr(), _native.dtype_code(w_lp.dtype), K, sk, sc, sr, ss,
                                                        _native.stream_ptr(w_lp.device)), "pdt_stem_weight_s2d")
        return out
    wp = F.pad(w_lp, (0, 1, 0, 1))                                   # 8x8, zero last row/col
    wp = wp.reshape(K, 3, 4, 2, 4, 2).permute(0, 2, 4, 3, 5, 1)       # k, r', s', dy, dx, c
    wp = wp.reshape(K, 4, 4, 12)
    return F.pad(wp, (0, 4)).contiguous()


def stem_weight_grad_from_s2d(dwp, K):
    """inverse of stem_weight_s2d for the fp32 gradient: [K,4,4,16] -> [K,3,7,7]."""
    if _stem_weight_native(dwp) and dwp.dtype == torch.float32 and dwp.is_contiguous() and dwp.numel() == K * 256:
        dw = torch.empty((K, 3, 7, 7), dtype=torch.float32, device=dwp.device, memory_format=CL)
        sk, sc, sr, ss = dw.stride()
        _native.check(_native.lib().pdt_stem_weight_grad_unpack(dwp.data_ptr(), dw.data_ptr(), K, sk, sc, sr, ss,
                                                                _native.stream_ptr(dwp.device)), "pdt_stem_weight_grad_unpack")
        return dw
    g = dwp.reshape(K, 4, 4, 16)[..., :12].reshape(K, 4, 4, 2, 2, 3).permute(0, 5, 1, 3, 2, 4)   # k,c,r',dy,s',dx
    return g.reshape(K, 3, 8, 8)[:, :, :7, :7]


def stem_supported(x, conv):
    return (enabled() and x.is_cuda and x.dtype in (torch.float32, torch.uint8) and x.dim() == 4 and x.shape[1] == 3
            and conv.kernel_size == (7, 7) and conv.stride == (2, 2) and conv.padding == (3, 3)
            and conv.out_channels % 8 == 0 and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0
            and os.environ.get("PDT_STEM", "native") == "native")


class _StemNative(torch.autograd.Function):
    """conv1 of torchvision ResNet on fp32 NCHW images: space-to-depth + 4x4/1 implicit GEMM.

    The transformed image has 16 channels per pixel.  In 16-bit types 4 horizontal neighbours are one
    128-byte K row (taps 4 x 1); in fp32/TF32 two neighbours are (taps 4 x 2, horizontal tap distance
    2 pixels).  Either way the filter is the same [K, 4, 4, 16] tensor."""

    @staticmethod
    def _geom(dtype, HP, WP, Q):
        es = 4 if dtype == torch.float32 else 2
        px = 16 * es                                   # bytes per transformed pixel
        if es == 2:
            return dict(C=64, S=1, dil=1, Wt=Q, sw=px, sh=WP * px, sn=HP * WP * px)
        return dict(C=32, S=2, dil=2, Wt=Q + 2, sw=px, sh=WP * px, sn=HP * WP * px)

    @staticmethod
    def forward(ctx, x, weight, w_lp, bnpack=None):
        with _native.guard(x.device):
            return _StemNative._forward_impl(ctx, x, weight, w_lp, bnpack)

    @staticmethod
    def _forward_impl(ctx, x, weight, w_lp, bnpack=None):
        ctx.set_materialize_grads(False)
        N, _, H, W = x.shape
        K = weight.shape[0]
        lp = w_lp.dtype
        HP, WP = (H + 6) // 2, (W + 6) // 2
        P, Q = HP - 3, WP - 3
        xs = torch.empty((N, HP, WP, 16), dtype=lp, device=x.device)
        x = x.contiguous()
        if x.dtype == torch.uint8:      # raw pixels: ToTensor + Normalize are fused into the layout kernel
            from . import input_norm
            sc, bs = input_norm.scale_bias_ctypes()
            _native.check(_native.lib().pdt_stem_s2d_u8(x.data_ptr(), xs.data_ptr(), _native.dtype_code(lp), N, H, W, 3,
                                                        _native.ctypes.addressof(sc), _native.ctypes.addressof(bs),
                                                        _native.stream_ptr(x.device)), "pdt_stem_s2d_u8")
        else:
            _native.check(_native.lib().pdt_stem_s2d(x.data_ptr(), xs.data_ptr(), _native.dtype_code(lp), N, H, W, 3,
                                                     _native.stream_ptr(x.device)), "pdt_stem_s2d")
        wp = stem_weight_s2d(w_lp)
        y = torch.empty((N, K, P, Q), dtype=lp, device=x.device, memory_format=CL)
        g = _StemNative._geom(lp, HP, WP, Q)
        lib = _native.lib()
        if (os.environ.get("PDT_STEM_CONV", "patch") == "patch"
                and lib.pdt_stem_conv_supported(_native.dtype_code(lp), N, HP, WP, K) == 1):
            # one linear TMA load per tile, sixteen row-shifted 32-byte-swizzled views (conv_halo.cu: k_stem_conv);
            # with bnpack = (running_mean, running_var, momentum, eps) its epilogue also produces the batch statistics
            stats = None
            if bnpack is not None:
                from .bn import _bn_scratch
                rm, rv, momentum, eps = bnpack[:4]
                comm = bnpack[4] if len(bnpack) > 4 else None
                mean = torch.empty(K, dtype=torch.float32, device=x.device)
                invstd = torch.empty_like(mean)
                stats = (mean, invstd)
                _native.check(lib.pdt_stem_conv(xs.data_ptr(), wp.data_ptr(), y.data_ptr(), _native.dtype_code(lp), N, HP, WP, K,
                                                _bn_scratch(x.device, K).data_ptr(), comm, _native.ptr(rm), _native.ptr(rv),
                                                momentum, eps, mean.data_ptr(), invstd.data_ptr(),
                                                _native.stream_ptr(x.device)), "pdt_stem_conv")
            else:
                _native.check(lib.pdt_stem_conv(xs.data_ptr(), wp.data_ptr(), y.data_ptr(), _native.dtype_code(lp), N, HP, WP, K,
                                                None, None, None, None, 0.1, 1e-5, None, None, _native.stream_ptr(x.device)),
                              "pdt_stem_conv")
            if stats is not None:
                ctx.save_for_backward(xs)
                ctx.geom = (N, HP, WP, P, Q, K)
                ctx.mark_non_differentiable(*stats)
                return y, stats[0], stats[1]
        elif bnpack is not None and K <= 512 and os.environ.get("PDT_BNSTATS_FUSION", "1") == "1":
            # generic im2col kernel (fp32 / shapes the patch kernel declines) with the statistics epilogue
            from .bn import _bn_scratch
            rm, rv, momentum, eps = bnpack[:4]
            comm = bnpack[4] if len(bnpack) > 4 else None
            mean = torch.empty(K, dtype=torch.float32, device=x.device)
            invstd = torch.empty_like(mean)
            _native.check(lib.pdt_conv_fprop_ex_bnstats(
                xs.data_ptr(), wp.data_ptr(), y.data_ptr(), _native.dtype_code(lp), N, HP, g["Wt"], g["C"], g["C"], g["sw"], g["sh"],
                g["sn"], K, 4, g["S"], 1, 0, 0, 0, 0, P, Q, g["dil"], _bn_scratch(x.device, K).data_ptr(), comm, _native.ptr(rm),
                _native.ptr(rv), momentum, eps, mean.data_ptr(), invstd.data_ptr(), _native.stream_ptr(x.device)),
                "pdt_conv_fprop_ex_bnstats")
            ctx.save_for_backward(xs)
            ctx.geom = (N, HP, WP, P, Q, K)
            ctx.mark_non_differentiable(mean, invstd)
            return y, mean, invstd
        else:
            fprop_ex(xs.data_ptr(), wp, y, None, lp, N, HP, g["Wt"], g["C"], g["C"], g["sw"], g["sh"], g["sn"], K, 4, g["S"], 1,
                     (0, 0, 0, 0), P, Q, P, Q, 1, 0, 0, x.device, dil_w=g["dil"])
        ctx.save_for_backward(xs)
        ctx.geom = (N, HP, WP, P, Q, K)
        return y

    @staticmethod
    def backward(ctx, dy, *unused_stat_grads):
        with _native.guard(dy.device):
            return _StemNative._backward_impl(ctx, dy) + (None,)

    @staticmethod
    def _backward_impl(ctx, dy):
        (xs,) = ctx.saved_tensors
        N, HP, WP, P, Q, K = ctx.geom
        dy = dy.contiguous(memory_format=CL)
        if xs.dtype == torch.float32 and os.environ.get("PDT_WGRAD_TF32", "1") != "1":
            # same reason as wgrad_supported(): 4x4/1 weight gradient over the transformed image through aten
            xs_nchw = xs.permute(0, 3, 1, 2)
            wshape = (K, 16, 4, 4)
            dwp4 = torch.ops.aten.convolution_backward(dy, xs_nchw, torch.empty(wshape, device=dy.device), None,
                                                       (1, 1), (0, 0), (1, 1), False, (0, 0), 1, (False, True, False))[1]
            return None, stem_weight_grad_from_s2d(dwp4.permute(0, 2, 3, 1).reshape(K, 256), K), None
        dwp = torch.zeros((K, 256), dtype=torch.float32, device=dy.device)
        _native.ensure_deterministic(dy.device)
        lib = _native.lib()
        if (os.environ.get("PDT_STEM_WGRAD", "patch") == "patch" and not _native.deterministic()
                and lib.pdt_stem_wgrad_supported(_native.dtype_code(dy.dtype), N, HP, WP, K) == 1):
            # the image rows of a tile arrive once; every tap is a shifted MN-major view of them (conv_halo.cu: k_stem_wgrad)
            _native.check(lib.pdt_stem_wgrad(xs.data_ptr(), dy.data_ptr(), dwp.data_ptr(), _native.dtype_code(dy.dtype), N, HP, WP, K,
                                             _native.stream_ptr(dy.device)), "pdt_stem_wgrad")
            return None, stem_weight_grad_from_s2d(dwp, K), None
        g = _StemNative._geom(xs.dtype, HP, WP, Q)
        _native.check(_native.lib().pdt_conv_wgrad_ex(
            xs.data_ptr(), dy.data_ptr(), dwp.data_ptr(), _native.dtype_code(dy.dtype), N, HP, g["Wt"], g["C"], g["C"],
            g["sw"], g["sh"], g["sn"], K, 4, g["S"], 1, 0, 0, 0, 0, P, Q, g["dil"], _native.stream_ptr(dy.device)),
            "pdt_conv_wgrad_ex")
        return None, stem_weight_grad_from_s2d(dwp, K), None


def stem_native(x, conv, w_lp, bnpack=None):
    """bnpack as in conv_native(): returns (y, mean, invstd) when the patch kernel could fuse the statistics."""
    return _StemNative.apply(x, conv.weight, w_lp, bnpack)


# ------------------------------------------------------------------------------------------ linear
class _LinearNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, w_lp):
        with _native.guard(x.device):
            return _LinearNative._forward_impl(ctx, x, weight, bias, w_lp)

    @staticmethod
    def _forward_impl(ctx, x, weight, bias, w_lp):
        B, C = x.shape
        K = w_lp.shape[0]
        x = x.contiguous()
        y = torch.empty((B, K), dtype=x.dtype, device=x.device)
        bias32 = bias.detach().float().contiguous() if bias is not None else None
        es = x.element_size()
        fprop_ex(x.data_ptr(), w_lp, y, bias32, x.dtype, B, 1, 1, C, C, C * es, C * es, C * es, K, 1, 1, 1, (0, 0, 0, 0),
                 1, 1, 1, 1, 1, 0, 0, x.device)
        ctx.save_for_backward(x, w_lp)
        ctx.has_bias = bias is not None
        ctx.weight = weight
        return y

    @staticmethod
    def backward(ctx, dy):
        with _native.guard(dy.device):
            return _LinearNative._backward_impl(ctx, dy)

    @staticmethod
    def _backward_impl(ctx, dy):
        x, w_lp = ctx.saved_tensors
        B, C = x.shape
        K = w_lp.shape[0]
        dy = dy.contiguous()
        dx = None
        if ctx.needs_input_grad[0]:
            w_t = w_lp.t().contiguous()                      # [C][K]: "filter" of the transposed product
            dx = torch.empty((B, C), dtype=dy.dtype, device=dy.device)
            es = dy.element_size()
            fprop_ex(dy.data_ptr(), w_t, dx, None, dy.dtype, B, 1, 1, K, K, K * es, K * es, K * es, C, 1, 1, 1, (0, 0, 0, 0),
                     1, 1, 1, 1, 1, 0, 0, dy.device)
        if wgrad_supported(C, K, dy.dtype):
            dw = direct_grad(ctx.weight)
            if dw is None:
                dw = torch.zeros((K, C), dtype=torch.float32, device=dy.device)
            _native.ensure_deterministic(dy.device)
            _native.check(_native.lib().pdt_conv_wgrad(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), _native.dtype_code(dy.dtype),
                                                       B, 1, 1, C, K, 1, 1, 1, 0, _native.stream_ptr(dy.device)),
                          "pdt_conv_wgrad(fc)")
        else:
            dw = (dy.t() @ x).float()
        db = dy.float().sum(0) if ctx.has_bias else None
        return dx, dw, db, None


def linear_supported(x, fc):
    # fp32 training: the classifier runs on tcgen05 kind::tf32 like the convolutions whenever TF32 convolutions
    # are allowed (torch itself would keep cuBLAS fp32 unless matmul.allow_tf32 is also set; PDT_FC_TF32=0
    # restores exactly that rule) — no cuBLAS call on the training path.
    ok_dtype = x.dtype in (torch.bfloat16, torch.float16) or (
        x.dtype == torch.float32 and tf32_enabled()
        and (torch.backends.cuda.matmul.allow_tf32 or os.environ.get("PDT_FC_TF32", "1") == "1"))
    return (enabled() and x.is_cuda and x.dim() == 2 and ok_dtype
            and fc.in_features % 64 == 0 and fc.out_features % 8 == 0)


def linear_native(x, fc, w_lp):
    return _LinearNative.apply(x, fc.weight, fc.bias, w_lp.contiguous())


_conv._conv_native = conv_native
_conv._stem_native = (stem_supported, stem_native)
_conv._linear_native = (linear_supported, linear_native)
