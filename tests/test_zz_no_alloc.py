"""Allocation-free BatchNorm statistics entry point, called through the C ABI.

``pdt_bn_fwd_stats`` needs a caller-provided scratch buffer.  Callers that cannot allocate one at
the call site (no cudaMalloc allowed inside their loop, a foreign memory pool they must not grow, a
bare C program) use ``pdt_bn_fwd_stats_noalloc`` instead, whose scratch is a fixed device array inside
the library.  This test loads the shared library with plain ctypes (no package wrapper), runs both
entry points on the same input twice in a row and requires the same statistics.  It is collected last
so that a failure here cannot hide the other GPU tests in a ``-x`` run.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorch_distributed_template_b200", "lib", "libpdt_b200.so")
BF16 = 0
E_TOO_LARGE = -3


def _lib():
    if not os.path.exists(LIB):
        from pytorch_distributed_template_b200 import build
        build.build()
    lib = ctypes.CDLL(LIB)
    vp, i, l, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_float
    lib.pdt_bn_scratch_bytes.argtypes, lib.pdt_bn_scratch_bytes.restype = [i], ctypes.c_size_t
    lib.pdt_bn_fwd_stats.argtypes = [vp, l, i, i, vp, vp, vp, vp, f, f, vp, vp, vp]
    lib.pdt_bn_fwd_stats.restype = i
    lib.pdt_bn_fwd_stats_noalloc.argtypes = [vp, l, i, i, vp, vp, f, f, vp, vp, vp]
    lib.pdt_bn_fwd_stats_noalloc.restype = i
    return lib


@pytest.mark.parametrize("C,M", [(64, 6 * 56 * 56), (512, 9 * 7 * 7)])
def test_noalloc_stats_match_scratch_version_twice(C, M):
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device="cpu").manual_seed(C)
    scratch = torch.zeros(lib.pdt_bn_scratch_bytes(C), dtype=torch.uint8, device=dev)
    for call in range(2):                                   # second call reuses whatever the first left behind
        x = (torch.randn(M, C, generator=g) * (1.0 + call) + 0.25 * call).to(dev).to(torch.bfloat16)
        out = {}
        for name in ("scratch", "noalloc"):
            rm = torch.zeros(C, device=dev)
            rv = torch.ones(C, device=dev)
            mean = torch.empty(C, device=dev)
            invstd = torch.empty(C, device=dev)
            if name == "scratch":
                rc = lib.pdt_bn_fwd_stats(x.data_ptr(), M, C, BF16, scratch.data_ptr(), None, rm.data_ptr(), rv.data_ptr(),
                                          0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(), stream)
            else:
                # free bytes as the DRIVER reports them: a cudaMalloc inside the library would show up here, which
                # torch.cuda.memory_allocated (torch's own caching allocator only) would not notice.  The first
                # call on a device may load the module that holds the static array, so prime it once.
                if call == 0:
                    lib.pdt_bn_fwd_stats_noalloc(x.data_ptr(), M, C, BF16, rm.data_ptr(), rv.data_ptr(), 0.0, 1e-5,
                                                 mean.data_ptr(), invstd.data_ptr(), stream)
                    torch.cuda.synchronize(dev)
                    rm.zero_(); rv.fill_(1.0)
                    torch.cuda.synchronize(dev)
                free_before = torch.cuda.mem_get_info(dev)[0]
                torch_before = torch.cuda.memory_allocated(dev)
                rc = lib.pdt_bn_fwd_stats_noalloc(x.data_ptr(), M, C, BF16, rm.data_ptr(), rv.data_ptr(), 0.1, 1e-5,
                                                  mean.data_ptr(), invstd.data_ptr(), stream)
                torch.cuda.synchronize(dev)
                assert torch.cuda.mem_get_info(dev)[0] == free_before
                assert torch.cuda.memory_allocated(dev) == torch_before
            assert rc == 0, (name, rc)
            torch.cuda.synchronize(dev)
            out[name] = (mean, invstd, rm, rv)
        for a, b in zip(out["scratch"], out["noalloc"]):
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
        xf = x.float()                                      # and both agree with a plain PyTorch fp32 reference
        assert torch.allclose(out["noalloc"][0], xf.mean(0), rtol=1e-3, atol=1e-3)
        assert torch.allclose(out["noalloc"][1], torch.rsqrt(xf.var(0, unbiased=False) + 1e-5), rtol=1e-3, atol=1e-3)


def test_noalloc_rejects_sizes_beyond_the_fixed_limit():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    C, M = 1024, 64
    x = torch.zeros(M, C, dtype=torch.bfloat16, device=dev)
    buf = torch.zeros(4, C, device=dev)
    rc = lib.pdt_bn_fwd_stats_noalloc(x.data_ptr(), M, C, BF16, buf[0].data_ptr(), buf[1].data_ptr(), 0.1, 1e-5,
                                      buf[2].data_ptr(), buf[3].data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    assert rc == E_TOO_LARGE
