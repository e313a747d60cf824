"""Diagnostic: does k_conv3x3_halo<64> run faster when the padded row length Wp = W + 2 is a multiple of 8
(row-shifted tap views that start on an 8-row swizzle-atom boundary)?  Time per 128-slot tile for several W."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, conv_native  # noqa: E402

dev = torch.device("cuda", 0)
CL = torch.channels_last
_native.require()
torch.manual_seed(0)
N, H = 1184, 56
for W in (14, 22, 30, 38, 46, 52, 53, 54, 55, 56):
    x = torch.randn(N, 64, H, W, device=dev).to(torch.float16).contiguous(memory_format=CL)
    w = (torch.randn(64, 64, 3, 3, device=dev) / 24).to(torch.float16).contiguous(memory_format=CL)
    for _ in range(3):
        conv_native.fprop(x, w, 1, 1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        conv_native.fprop(x, w, 1, 1)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) / 10 * 1e3
    Wp = W + 2
    TR = min(128 // Wp, H)
    tiles = N * ((H + TR - 1) // TR)
    per_sm = (tiles + 147) // 148
    print("W {:2d} Wp {:2d} (mod 8 = {}) TR {} slots {:3d}: {:7.1f} us, {:6.0f} ns per tile per SM, {:5.0f} TFLOP/s".format(
        W, Wp, Wp % 8, TR, TR * Wp, us, us * 1e3 / per_sm, 2.0 * N * H * W * 64 * 64 * 9 / us / 1e6), flush=True)
