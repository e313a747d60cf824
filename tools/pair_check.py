"""k_conv_igemm with CTA pairs (PDT_CONV_PAIR=1, tcgen05.mma.cta_group::2) against F.conv2d, and its time per launch.
Run once with PDT_CONV_PAIR=0 and once with 1 (the switch is read once per process)."""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, conv_native  # noqa: E402

dev = torch.device("cuda", 0)
CL = torch.channels_last
_native.require()
torch.manual_seed(0)
print("PDT_CONV_PAIR =", os.environ.get("PDT_CONV_PAIR", "0"))
for (n, c, h, k, r, s, p) in ((100, 256, 14, 256, 3, 1, 1), (1200, 256, 14, 256, 3, 1, 1), (1200, 128, 28, 256, 3, 2, 1),
                              (1200, 512, 7, 512, 3, 1, 1), (1200, 128, 28, 128, 1, 1, 0), (600, 256, 14, 512, 1, 1, 0)):
    x = torch.randn(n, c, h, h, device=dev).to(torch.float16).contiguous(memory_format=CL)
    w = (torch.randn(k, c, r, r, device=dev) / (c * r * r) ** 0.5).to(torch.float16).contiguous(memory_format=CL)
    y = conv_native.fprop(x, w, s, p)
    torch.cuda.synchronize()
    ref = F.conv2d(x[:64].float(), w.float(), None, s, p)
    err = ((y[:64].float() - ref).abs().max() / ref.abs().max()).item()
    ref2 = F.conv2d(x[-32:].float(), w.float(), None, s, p)
    err2 = ((y[-32:].float() - ref2).abs().max() / ref2.abs().max()).item()
    for _ in range(3):
        conv_native.fprop(x, w, s, p)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        conv_native.fprop(x, w, s, p)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) / 20 * 1e3
    flops = 2.0 * y.numel() * c * r * r
    print("N{} {}->{} {}x{} k{} s{}: err {:.2e} / {:.2e}  {:.1f} us  {:.0f} TFLOP/s".format(n, c, k, h, h, r, s, err, err2, us, flops / us / 1e6),
          flush=True)
    assert err < 4e-3 and err2 < 4e-3
print("PAIR_CHECK_DONE")
