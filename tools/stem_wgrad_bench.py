"""Stem weight gradient per launch: k_stem_wgrad (patch reuse) vs the generic split-K kernel (PDT_STEM_WGRAD=generic)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, conv_native  # noqa: E402

dev = torch.device("cuda", 0)
_native.require()
torch.manual_seed(0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1200
conv = torch.nn.Conv2d(3, 64, 7, 2, 3, bias=False).to(dev)
x = torch.randint(0, 256, (N, 3, 224, 224), dtype=torch.uint8, device=dev)
w_lp = conv.weight.detach().half()
for mode in ("patch", "generic"):
    os.environ["PDT_STEM_WGRAD"] = mode
    y = conv_native.stem_native(x, conv, w_lp)
    dy = torch.randn_like(y)
    for _ in range(3):
        y.backward(dy, retain_graph=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        y.backward(dy, retain_graph=True)
    b.record()
    torch.cuda.synchronize()
    print("{}: {:.1f} us per backward (weight gradient + zero fill + re-layout)".format(mode, a.elapsed_time(b) / 10 * 1e3), flush=True)
