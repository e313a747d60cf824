"""Diagnostic (library built with -DPDT_TAIL_STAMPS): SM-clock stamps of the LAST CTA of k_conv_igemm's statistics tail."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, bn as bnops, conv_native  # noqa: E402

dev = torch.device("cuda", 0)
CL = torch.channels_last
_native.require()
N = 150
names = ["epilogue start", "tiles done", "row written", "fence+bar", "ticket taken", "fold done", "finalised"]
for (c, h, k, r, s, p) in ((256, 14, 256, 3, 1, 1), (512, 7, 512, 3, 1, 1), (256, 14, 512, 1, 2, 0)):
    x = torch.randn(N, c, h, h, device=dev).to(torch.float16).contiguous(memory_format=CL)
    w = (torch.randn(k, c, r, r, device=dev) / (c * r * r) ** 0.5).to(torch.float16).contiguous(memory_format=CL)
    rm, rv = torch.zeros(k, device=dev), torch.ones(k, device=dev)
    for rep in range(3):
        conv_native.fprop_bnstats(x, w, s, p, rm, rv, 0.1, 1e-5)
        torch.cuda.synchronize()
        st = bnops._bn_scratch(dev, k)[64:64 + 56].view(torch.int64).tolist()
        d = [(b - a) / 1.9e3 for a, b in zip(st[:-1], st[1:])]
        print("{}->{} {}x{} k{} s{}: ".format(c, k, h, h, r, s) + "  ".join("{} +{:.2f}us".format(n, v) for n, v in zip(names[1:], d)), flush=True)
