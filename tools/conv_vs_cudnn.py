"""Per ResNet-18 convolution shape: our kernels against cuDNN (torch's F.conv2d / aten.convolution_backward on
channels_last tensors, cudnn.benchmark on) — forward, data gradient, weight gradient, per launch.

    python tools/conv_vs_cudnn.py [--batch 1200] [--dtype fp16|bf16|fp32] [--md profiles/conv_vs_cudnn.md]

CUDA events around 20 back-to-back launches after 5 warm-ups (inputs >> L2 at batch 1200).  Ours goes through the
same Python entry points the autograd Functions use (conv_native.fprop / dgrad_stride2 / wgrad), so patch-reuse and
CTA-pair variants are picked exactly as in training; the dgrad filter re-layout is outside the timed region for
both (ours caches it per optimizer step, cuDNN does it internally)."""
import argparse
import os
import sys
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native, conv_native  # noqa: E402

CL = torch.channels_last
# (name, C_in, C_out, H_in, kernel, stride, pad) — every distinct convolution of torchvision's ResNet-18 at 224 x 224
# except the 3-channel stem (which has its own space-to-depth kernel and no cuDNN counterpart of the same shape)
SHAPES = [
    ("layer1 3x3", 64, 64, 56, 3, 1, 1),
    ("layer2.0 3x3/2", 64, 128, 56, 3, 2, 1),
    ("layer2.0 1x1/2", 64, 128, 56, 1, 2, 0),
    ("layer2 3x3", 128, 128, 28, 3, 1, 1),
    ("layer3.0 3x3/2", 128, 256, 28, 3, 2, 1),
    ("layer3.0 1x1/2", 128, 256, 28, 1, 2, 0),
    ("layer3 3x3", 256, 256, 14, 3, 1, 1),
    ("layer4.0 3x3/2", 256, 512, 14, 3, 2, 1),
    ("layer4.0 1x1/2", 256, 512, 14, 1, 2, 0),
    ("layer4 3x3", 512, 512, 7, 3, 1, 1),
]


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1200)
    ap.add_argument("--dtype", default="fp16")
    ap.add_argument("--md", default=None)
    a = ap.parse_args()
    dt = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[a.dtype]
    dev = torch.device("cuda", 0)
    _native.require()
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = True
    torch.manual_seed(0)
    rows = []
    for (name, c, k, h, r, s, p) in SHAPES:
        n = a.batch
        x = torch.randn(n, c, h, h, device=dev).to(dt).contiguous(memory_format=CL)
        w = (torch.randn(k, c, r, r, device=dev) / (c * r * r) ** 0.5).to(dt).contiguous(memory_format=CL)
        P = (h + 2 * p - r) // s + 1
        dy = (torch.randn(n, k, P, P, device=dev) / 8).to(dt).contiguous(memory_format=CL)
        flops = 2.0 * n * P * P * k * c * r * r
        out = {"name": name, "gflop": flops / 1e9}
        # ---- forward
        out["fwd_ours"] = timed(lambda: conv_native.fprop(x, w, s, p))
        out["fwd_cudnn"] = timed(lambda: torch.nn.functional.conv2d(x, w, None, s, p))
        # ---- data gradient
        if s == 1:
            w_t = conv_native.dgrad_weight(w, list(range(r - 1, -1, -1)), list(range(r - 1, -1, -1)))
            out["dgrad_ours"] = timed(lambda: conv_native.fprop(dy, w_t, 1, r - 1 - p))
        else:
            st = types.SimpleNamespace(weights_version=0)         # stands for the parameter arena: the tap sub-filters are
            conv_native.dgrad_stride2(dy, w, x.shape, p, st)         # laid out once per optimizer step, not per backward
            conv_native.prepare_dgrad_weights(st)
            out["dgrad_ours"] = timed(lambda: conv_native.dgrad_stride2(dy, w, x.shape, p, st))
        out["dgrad_cudnn"] = timed(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, (s, s), (p, p), (1, 1), False, (0, 0), 1,
                                                                                  (True, False, False)))
        # ---- weight gradient (ours: fp32 accumulation into a zeroed fp32 tensor, as the gradient arena receives it)
        dw = torch.zeros(k, c, r, r, device=dev).contiguous(memory_format=CL)
        out["wgrad_ours"] = timed(lambda: conv_native.wgrad(x, dy, (k, c, r, r), s, p, out=dw))
        out["wgrad_cudnn"] = timed(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, (s, s), (p, p), (1, 1), False, (0, 0), 1,
                                                                                  (False, True, False)))
        rows.append(out)
        print("{:16s} {:7.1f} GF | fwd {:7.1f} / {:7.1f} us | dgrad {:7.1f} / {:7.1f} us | wgrad {:7.1f} / {:7.1f} us   (ours / cuDNN)".format(
            name, out["gflop"], out["fwd_ours"], out["fwd_cudnn"], out["dgrad_ours"], out["dgrad_cudnn"], out["wgrad_ours"], out["wgrad_cudnn"]),
            flush=True)
    tot = {key: sum(r_[key] for r_ in rows) for key in rows[0] if key not in ("name",)}
    print("sum              fwd {:.0f} / {:.0f} us | dgrad {:.0f} / {:.0f} us | wgrad {:.0f} / {:.0f} us".format(
        tot["fwd_ours"], tot["fwd_cudnn"], tot["dgrad_ours"], tot["dgrad_cudnn"], tot["wgrad_ours"], tot["wgrad_cudnn"]))
    if a.md:
        with open(a.md, "a") as f:
            f.write("\n### batch {}, {} (cuDNN {}, cudnn.benchmark on, channels_last)\n\n".format(a.batch, a.dtype, torch.backends.cudnn.version()))
            f.write("| convolution | GFLOP | fwd ours (us) | fwd cuDNN | dgrad ours | dgrad cuDNN | wgrad ours | wgrad cuDNN | ours TFLOP/s fwd / dgrad / wgrad |\n")
            f.write("|---|---|---|---|---|---|---|---|---|\n")
            for r_ in rows:
                f.write("| {} | {:.1f} | {:.1f} | {:.1f} | {:.1f} | {:.1f} | {:.1f} | {:.1f} | {:.0f} / {:.0f} / {:.0f} |\n".format(
                    r_["name"], r_["gflop"], r_["fwd_ours"], r_["fwd_cudnn"], r_["dgrad_ours"], r_["dgrad_cudnn"], r_["wgrad_ours"],
                    r_["wgrad_cudnn"], r_["gflop"] / r_["fwd_ours"] * 1e3, r_["gflop"] / r_["dgrad_ours"] * 1e3,
                    r_["gflop"] / r_["wgrad_ours"] * 1e3))
            f.write("| **sum** | | {:.0f} | {:.0f} | {:.0f} | {:.0f} | {:.0f} | {:.0f} | |\n".format(
                tot["fwd_ours"], tot["fwd_cudnn"], tot["dgrad_ours"], tot["dgrad_cudnn"], tot["wgrad_ours"], tot["wgrad_cudnn"]))
    print("CONV_VS_CUDNN_DONE")


if __name__ == "__main__":
    main()
