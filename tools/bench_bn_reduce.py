"""Time pdt_bn_bwd_reduce / apply on one shape with CUDA events (diagnostic)."""
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, bn as bnops  # noqa: E402


def main():
    N, C, H = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    dev = torch.device("cuda", 0)
    x = torch.randn(N, C, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dy = torch.randn_like(x)
    mean = torch.zeros(C, device=dev); invstd = torch.ones(C, device=dev)
    w = torch.ones(C, device=dev); b = torch.zeros(C, device=dev)
    dg = torch.empty(C, device=dev); db = torch.empty(C, device=dev); k = torch.empty(2 * C, device=dev)
    scratch = bnops._bn_scratch(dev, C)
    lib, s, dt = _native.lib(), _native.stream_ptr(dev), _native.dtype_code(x.dtype)
    M = N * H * H

    def run():
        _native.check(lib.pdt_bn_bwd_reduce(dy.data_ptr(), None, x.data_ptr(), M, C, dt, mean.data_ptr(), invstd.data_ptr(),
                                            w.data_ptr(), b.data_ptr(), 1, scratch.data_ptr(), None, dg.data_ptr(), db.data_ptr(),
                                            k.data_ptr(), s), "reduce")
    for _ in range(5):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        run()
    e1.record()
    torch.cuda.synchronize()
    print("N=%d C=%d H=%d stream=%s cap=%s: %.1f us/launch" % (N, C, H, os.environ.get("PDT_BN_STREAM", "1"),
                                                            os.environ.get("PDT_BN_GRID_CAP", "-"), e0.elapsed_time(e1) * 20))


if __name__ == "__main__":
    main()
