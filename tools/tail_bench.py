"""What the cross-CTA statistics tail costs at small per-GPU batches (diagnostic): convolution with and without the
BatchNorm-statistics epilogue, and the BatchNorm backward reduction, per launch, at the shapes of a 150-image step.
CUDA events around 50 back-to-back launches after 5 warm-ups."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, bn as bnops, conv_native  # noqa: E402

dev = torch.device("cuda", 0)
CL = torch.channels_last
_native.require()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 150


def timed(fn, iters=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


for (c, h, k, r, s, p) in ((64, 56, 64, 3, 1, 1), (128, 28, 128, 3, 1, 1), (256, 14, 256, 3, 1, 1), (512, 7, 512, 3, 1, 1),
                           (256, 14, 512, 3, 2, 1), (256, 14, 512, 1, 2, 0)):
    x = torch.randn(N, c, h, h, device=dev).to(torch.float16).contiguous(memory_format=CL)
    w = (torch.randn(k, c, r, r, device=dev) / (c * r * r) ** 0.5).to(torch.float16).contiguous(memory_format=CL)
    rm, rv = torch.zeros(k, device=dev), torch.ones(k, device=dev)
    t0 = timed(lambda: conv_native.fprop(x, w, s, p))
    t1 = timed(lambda: conv_native.fprop_bnstats(x, w, s, p, rm, rv, 0.1, 1e-5))
    print("conv N{} {}->{} {}x{} k{} s{}: plain {:.1f} us, with statistics {:.1f} us (+{:.1f})".format(N, c, k, h, h, r, s, t0, t1, t1 - t0),
          flush=True)

lib = _native.lib()
for (c, h) in ((64, 56), (128, 28), (256, 14), (512, 7)):
    x = torch.randn(N, c, h, h, device=dev).to(torch.float16).contiguous(memory_format=CL)
    dy = torch.randn_like(x)
    mean, invstd, g, b = torch.zeros(c, device=dev), torch.ones(c, device=dev), torch.ones(c, device=dev), torch.zeros(c, device=dev)
    dg, db, kk = torch.empty(c, device=dev), torch.empty(c, device=dev), torch.empty(2 * c, device=dev)
    scratch = bnops._bn_scratch(dev, c)
    st, dt, M = _native.stream_ptr(dev), _native.dtype_code(x.dtype), N * h * h

    def red():
        _native.check(lib.pdt_bn_bwd_reduce(dy.data_ptr(), None, x.data_ptr(), M, c, dt, mean.data_ptr(), invstd.data_ptr(), g.data_ptr(),
                                            b.data_ptr(), 1, scratch.data_ptr(), None, dg.data_ptr(), db.data_ptr(), kk.data_ptr(), st), "reduce")
    t = timed(red)
    ideal = 2 * x.numel() * 2 / 6.5e12 * 1e6
    print("bwd_reduce N{} C{} {}x{}: {:.1f} us (streaming floor {:.1f} us)".format(N, c, h, h, t, ideal), flush=True)
print("TAIL_BENCH_DONE")
