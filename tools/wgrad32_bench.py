"""fp32 (tf32) 64->64 3x3 weight gradient: time per launch and error vs aten.  PDT_WGRAD_HALO=0 selects the generic kernel."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, conv_native  # noqa: E402

dev = torch.device("cuda", 0)
CL = torch.channels_last
_native.require()
torch.backends.cudnn.allow_tf32 = False
torch.manual_seed(0)
print("PDT_WGRAD_HALO =", os.environ.get("PDT_WGRAD_HALO", "1"))
for n in (8, 150, 1200):
    x = torch.randn(n, 64, 56, 56, device=dev).contiguous(memory_format=CL)
    dy = (torch.randn(n, 64, 56, 56, device=dev) / 8).contiguous(memory_format=CL)
    nr = min(n, 32)
    dws = conv_native.wgrad(x[:nr].contiguous(memory_format=CL), dy[:nr].contiguous(memory_format=CL), (64, 64, 3, 3), 1, 1)
    ref = torch.ops.aten.convolution_backward(dy[:nr], x[:nr], torch.empty(64, 64, 3, 3, device=dev), None, (1, 1), (1, 1), (1, 1), False,
                                              (0, 0), 1, (False, True, False))[1]
    err = ((dws - ref).abs().max() / ref.abs().max()).item()
    out = torch.zeros(64, 64, 3, 3, device=dev).contiguous(memory_format=CL)
    for _ in range(3):
        conv_native.wgrad(x, dy, (64, 64, 3, 3), 1, 1, out=out)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        conv_native.wgrad(x, dy, (64, 64, 3, 3), 1, 1, out=out)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) / 20 * 1e3
    print("N{}: err {:.2e}  {:.1f} us  {:.0f} TFLOP/s".format(n, err, us, 2.0 * n * 56 * 56 * 64 * 64 * 9 / us / 1e6), flush=True)
print("WGRAD32_DONE")
