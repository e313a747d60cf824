"""k_conv_wgrad with CTA pairs (tcgen05.mma.cta_group::2, the default) against aten's weight gradient, and its time
per launch.  Run once with PDT_WGRAD_PAIR=0 and once without (the switch is read once per process)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, conv_native  # noqa: E402

dev = torch.device("cuda", 0)
CL = torch.channels_last
_native.require()
torch.manual_seed(0)
torch.backends.cudnn.allow_tf32 = False
print("PDT_WGRAD_PAIR =", os.environ.get("PDT_WGRAD_PAIR", "1"))
CASES = (
    # n, c, h, k, r, s, p, dtype
    (8, 64, 14, 128, 3, 1, 1, torch.float16),       # odd number of (r,s,c) tiles: the last pair is half empty
    (8, 64, 14, 200, 3, 1, 1, torch.bfloat16),      # short N tile
    (8, 128, 8, 512, 1, 1, 0, torch.float16),       # one (r,s,c) tile: single-CTA path
    (8, 64, 16, 256, 3, 2, 1, torch.float32),       # tf32, stride 2
    (150, 128, 28, 128, 3, 1, 1, torch.float16),
    (150, 128, 28, 256, 3, 2, 1, torch.float16),
    (150, 256, 14, 256, 3, 1, 1, torch.float16),
    (150, 256, 14, 512, 3, 2, 1, torch.float16),
    (150, 512, 7, 512, 3, 1, 1, torch.float16),
    (150, 256, 14, 512, 1, 2, 0, torch.float16),
    (1200, 128, 28, 128, 3, 1, 1, torch.float16),
    (1200, 256, 14, 256, 3, 1, 1, torch.float16),
    (1200, 512, 7, 512, 3, 1, 1, torch.float16),
    (1200, 256, 14, 256, 3, 1, 1, torch.float32),
)
for (n, c, h, k, r, s, p, dt) in CASES:
    x = torch.randn(n, c, h, h, device=dev).to(dt).contiguous(memory_format=CL)
    P = (h + 2 * p - r) // s + 1
    dy = (torch.randn(n, k, P, P, device=dev) / 8).to(dt).contiguous(memory_format=CL)
    dw = conv_native.wgrad(x, dy, (k, c, r, r), s, p)
    torch.cuda.synchronize()
    nref = min(n, 64)      # the oracle on a slice of the batch (fp32 aten), compared with our result on the same slice
    dws = conv_native.wgrad(x[:nref].contiguous(memory_format=CL), dy[:nref].contiguous(memory_format=CL), (k, c, r, r), s, p)
    ref = torch.ops.aten.convolution_backward(dy[:nref].float(), x[:nref].float(), torch.empty(k, c, r, r, device=dev), None, (s, s), (p, p),
                                              (1, 1), False, (0, 0), 1, (False, True, False))[1]
    err = ((dws - ref).abs().max() / ref.abs().max()).item()
    again = conv_native.wgrad(x, dy, (k, c, r, r), s, p)
    drift = ((again - dw).abs().max() / dw.abs().max()).item()       # atomics reorder: small, never large
    out = torch.zeros_like(dw)
    for _ in range(3):
        conv_native.wgrad(x, dy, (k, c, r, r), s, p, out=out)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        conv_native.wgrad(x, dy, (k, c, r, r), s, p, out=out)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) / 20 * 1e3
    flops = 2.0 * n * P * P * k * c * r * r
    print("N{} {}->{} {}x{} k{} s{} {}: err {:.2e} drift {:.1e}  {:.1f} us  {:.0f} TFLOP/s".format(
        n, c, k, h, h, r, s, str(dt)[6:], err, drift, us, flops / us / 1e6), flush=True)
    assert err < (4e-3 if dt != torch.float32 else 2e-3) and drift < 1e-3, (err, drift)
print("WGRAD_PAIR_CHECK_DONE")
