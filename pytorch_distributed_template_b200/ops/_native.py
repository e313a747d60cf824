"""ctypes binding of ``lib/libpdt_b200.so`` (C ABI in ``include/pdt_b200.h``).

The library is plain CUDA with no torch dependency; tensors cross the boundary
as raw device pointers plus the current torch stream.  On a machine with a GPU
a missing library is an error, not a reason to fall back silently: call
``require()`` from every native code path.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libpdt_b200.so")

MAX_RANKS = 8
BF16, FP16, FP32 = 0, 1, 2
_DTYPE_CODE = {torch.bfloat16: BF16, torch.float16: FP16, torch.float32: FP32}

_lib = None
_load_error = None
launch_count = 0   # kernels launched through this binding (bench.py reports it)


class Peers(ctypes.Structure):
    _fields_ = [("p", ctypes.c_void_p * MAX_RANKS)]


class CommDesc(ctypes.Structure):
    _fields_ = [("rank", ctypes.c_int), ("world", ctypes.c_int), ("pad", ctypes.c_void_p * MAX_RANKS),
                ("pad_mc", ctypes.c_void_p)]


MAX_SEGMENTS = 8


class Segments(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int), ("off", ctypes.c_size_t * MAX_SEGMENTS), ("len", ctypes.c_size_t * MAX_SEGMENTS)]


class EmuRanks(ctypes.Structure):
    _fields_ = [("pad", ctypes.c_void_p * MAX_RANKS), ("amp_state", ctypes.c_void_p * MAX_RANKS),
                ("momentum", ctypes.c_void_p * MAX_RANKS), ("io", ctypes.c_void_p * MAX_RANKS)]


def make_segments(pairs):
    """[(offset, length), ...] in elements -> PdtSegments (at most MAX_SEGMENTS)."""
    pairs = list(pairs)
    if not 1 <= len(pairs) <= MAX_SEGMENTS:
        raise ValueError("1..{} segments per launch (got {})".format(MAX_SEGMENTS, len(pairs)))
    s = Segments()
    s.n = len(pairs)
    for i, (o, l) in enumerate(pairs):
        s.off[i], s.len[i] = int(o), int(l)
    return s


def dtype_code(dtype):
    return _DTYPE_CODE[dtype]


def make_peers(ptrs):
    s = Peers()
    for i in range(MAX_RANKS):
        s.p[i] = int(ptrs[i]) if i < len(ptrs) and ptrs[i] else None
    return s


def make_comm(rank, world, pad_ptrs, pad_mc=0):
    d = CommDesc()
    d.rank, d.world = rank, world
    for i in range(MAX_RANKS):
        d.pad[i] = int(pad_ptrs[i]) if i < len(pad_ptrs) and pad_ptrs[i] else None
    d.pad_mc = int(pad_mc) if pad_mc else None
    return d


_vp, _i, _l, _f, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_float, ctypes.c_size_t
_PC, _PP = ctypes.POINTER(CommDesc), ctypes.POINTER(Peers)
_PS, _PE = ctypes.POINTER(Segments), ctypes.POINTER(EmuRanks)

_SIGNATURES = {
    "pdt_comm_pad_bytes": ([], _sz),
    "pdt_reduce_scatter_f32": ([_PC, _PP, _vp, _sz, _sz, _vp, _i, _i, _vp], _i),
    "pdt_sgd_allgather": ([_PC, _PP, _PP, _PP, _vp, _vp, _i, _vp, _PS, _vp, _vp, _i, _i, _vp], _i),
    "pdt_set_wgrad_workspace": ([_vp, _sz], _i),
    "pdt_metrics_update": ([_PC, _vp, _vp, _i, _f, _i, _vp, _f, _vp], _i),
    "pdt_emu_step": ([_PE, _i, _PP, _PP, _PP, _PS, _vp, _i, _vp], _i),
    "pdt_emu_small": ([_PE, _i, _i, _f, _PP, _sz, _i, _vp], _i),
    "pdt_allreduce_f32": ([_PC, _PP, _sz, _sz, _f, _i, _vp], _i),
    "pdt_allreduce_small_f32": ([_PC, _vp, _vp, _i, _f, _i, _vp], _i),
    "pdt_bcast_pull": ([_PC, _PP, _sz, _sz, _i, _i, _i, _vp], _i),
    "pdt_barrier": ([_PC, _i, _vp], _i),
    "pdt_scale_update": ([_vp, _f, _f, _i, _vp], _i),
    "pdt_sgd_flat": ([_vp, _vp, _vp, _vp, _i, _sz, _vp, _vp, _vp], _i),
    "pdt_bn_scratch_bytes": ([_i], _sz),
    "pdt_bn_fwd_stats": ([_vp, _l, _i, _i, _vp, _PC, _vp, _vp, _f, _f, _vp, _vp, _vp], _i),
    "pdt_bn_fwd_apply": ([_vp, _vp, _vp, _l, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _f, _vp], _i),
    "pdt_bn_bwd_reduce": ([_vp, _vp, _vp, _l, _i, _i, _vp, _vp, _vp, _vp, _i, _vp, _PC, _vp, _vp, _vp, _vp], _i),
    "pdt_bn_bwd_apply": ([_vp, _vp, _vp, _vp, _vp, _l, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp], _i),
    "pdt_bn_pool_fwd": ([_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp], _i),
    "pdt_bn_pool_bwd_reduce": ([_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _PC, _vp, _vp, _vp, _vp], _i),
    "pdt_bn_pool_bwd_apply": ([_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp], _i),
    "pdt_ce_scratch_bytes": ([_i], _sz),
    "pdt_ce_fwd": ([_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp], _i),
    "pdt_scale_cast": ([_vp, _vp, _vp, _i, _sz, _vp], _i),
    "pdt_maxpool_fwd": ([_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp], _i),
    "pdt_maxpool_bwd": ([_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp], _i),
    "pdt_avgpool_fwd": ([_vp, _vp, _i, _i, _i, _i, _vp], _i),
    "pdt_avgpool_bwd": ([_vp, _vp, _i, _i, _i, _i, _vp], _i),
    "pdt_peer_copy": ([_vp, _vp, _sz, _vp], _i),
    "pdt_nchw_to_nhwc": ([_vp, _vp, _i, _i, _i, _i, _i, _i, _vp], _i),
}


def _header_abi():
    import re
    hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "include", "pdt_b200.h")
    try:
        m = re.search(r"#define\s+PDT_ABI_VERSION\s+(\d+)", open(hdr).read())
        return int(m.group(1)) if m else None
    except OSError:
        return None


def _rebuild_if_stale():
    """The library next to the package must have been built from the sources next to the package (content hash,
    ``build.source_id()`` vs the id recorded by the build).  If it was not — a fresh checkout, an edited kernel —
    build it now.  Ranks of one node serialise on a lock file; whoever comes second finds the library current.
    A build error is raised, never swallowed: running a stale kernel is worse than not running."""
    from .. import build as _build
    try:
        want = _build.source_id()
    except OSError:
        return None                      # sources not shipped (binary-only install): nothing to compare with
    if _build.built_id() == want:
        return want
    if os.environ.get("PDT_NO_AUTOBUILD", "0") == "1":
        raise OSError("{} was not built from the sources in {} (PDT_NO_AUTOBUILD=1): run "
                      "`python -m pytorch_distributed_template_b200.build`".format(LIB_PATH, _build.CSRC))
    import fcntl
    os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if _build.built_id() != want:
                _build.build()           # raises when nvcc is missing or fails
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return want


def _load():
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return _lib
    try:
        want_id = _rebuild_if_stale()
        lib = ctypes.CDLL(LIB_PATH)
        want = _header_abi()
        have = lib.pdt_abi_version() if hasattr(lib, "pdt_abi_version") else None
        if want is not None and have != want:
            raise OSError("{} reports ABI {} but include/pdt_b200.h is at {}: stale build, run "
                          "`python -m pytorch_distributed_template_b200.build`".format(LIB_PATH, have, want))
        if want_id is not None:
            lib.pdt_build_id.restype = ctypes.c_char_p
            have_id = lib.pdt_build_id().decode()
            if have_id != want_id:
                raise OSError("{} was built from other sources (id {} != {}): run "
                              "`python -m pytorch_distributed_template_b200.build --force`".format(LIB_PATH, have_id, want_id))
        for name, (argtypes, restype) in _SIGNATURES.items():
            fn = getattr(lib, name, None)
            if fn is None:
                continue
            fn.argtypes, fn.restype = argtypes, restype
        _lib = lib
    except (OSError, RuntimeError) as e:  # not built / stale / build failed: require() reports it
        _load_error = e
    return _lib


def register(name, argtypes, restype=_i):
    """Used by the other ops modules to add signatures (conv/gemm)."""
    _SIGNATURES[name] = (argtypes, restype)
    if _lib is not None and hasattr(_lib, name):
        fn = getattr(_lib, name)
        fn.argtypes, fn.restype = argtypes, restype


def available():
    """True when the library is built AND a CUDA device is usable."""
    return torch.cuda.is_available() and _load() is not None


def require():
    lib = _load()
    if lib is None:
        raise RuntimeError(
            "libpdt_b200.so is not built ({}); run `python -m pytorch_distributed_template_b200.build` "
            "or `python -c 'import __graft_entry__ as g; g.build()'`".format(_load_error))
    return lib


def lib():
    return require()


_det_ws = {}


def deterministic():
    """PDT_DETERMINISTIC=1 (set by the trainer when --seed is given, like cudnn.deterministic in the reference,
    distributed.py:91-99): weight gradients fold their split-K partials in a fixed order instead of using atomics."""
    return os.environ.get("PDT_DETERMINISTIC", "0") == "1"


def ensure_deterministic(device):
    """Register (once per device) the workspace that switches the weight-gradient kernels to their
    deterministic path; a no-op unless ``deterministic()``."""
    if not deterministic() or device.type != "cuda":
        return
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx in _det_ws:
        return
    if torch.cuda.is_current_stream_capturing():
        raise RuntimeError("deterministic mode must be set up before graph capture (run one eager step first)")
    ws = torch.empty(int(os.environ.get("PDT_DETERMINISTIC_WS_MB", "256")) << 20, dtype=torch.uint8, device=device)
    with torch.cuda.device(idx):
        code = lib().pdt_set_wgrad_workspace(ws.data_ptr(), ws.numel())
    if code != 0:
        raise RuntimeError("pdt_set_wgrad_workspace failed ({})".format(code))
    _det_ws[idx] = ws


class _NoGuard:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NO_GUARD = _NoGuard()


def guard(device):
    """Make ``device`` the current CUDA device while our kernels are enqueued.

    torch ops carry their own device guards; a raw launch from the C library goes to whatever
    device is current, so a tensor on cuda:1 touched from a thread whose current device is 0
    (DataParallel, user code) would otherwise be launched into the wrong context."""
    if device.type != "cuda" or device.index is None or torch.cuda.current_device() == device.index:
        return _NO_GUARD
    return torch.cuda.device(device)


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t):
    return None if t is None else t.data_ptr()


def check(code, what):
    global launch_count
    launch_count += 1
    if code != 0:
        if code > 0:
            raise RuntimeError("{}: CUDA error {} ({})".format(what, code, _cuda_error_name(code)))
        raise RuntimeError("{}: unsupported arguments (code {})".format(what, code))


def _cuda_error_name(code):
    try:
        return torch.cuda.cudart().cudaGetErrorString(code) if hasattr(torch.cuda.cudart(), "cudaGetErrorString") else str(code)
    except Exception:
        return str(code)
