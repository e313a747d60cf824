"""Smallest case of every hand-written kernel family, for compute-sanitizer:

    compute-sanitizer --tool memcheck  python tools/sanitize_cases.py     # one tool per gpurun call
    compute-sanitizer --tool racecheck python tools/sanitize_cases.py
    compute-sanitizer --tool synccheck python tools/sanitize_cases.py

Each case also checks its result against torch, so a run without the sanitizer doubles as a smoke test.
Covered: k_conv_igemm (fprop, stride-1 and stride-2 data gradient), k_conv_wgrad, k_conv3x3_halo<64/128>,
k_wgrad3x3_halo, k_stem_conv (+ k_stem_s2d), k_stem_wgrad, sp::k_fwd_stats / k_fwd_apply / k_bwd_reduce / k_bwd_apply,
k_bn_pool_*_321 (16-bit and, in the fp32 step of model_case, the _f32 pair together with k_conv3x3_halo<.., float> and
k_wgrad3x3_halo_tf32), k_maxpool / k_avgpool, k_ce_fwd / k_scale_cast, k_sgd_flat / k_reduce_scatter<1>,
the rank-emulation launch (k_reduce_scatter / k_sgd_allgather / k_allreduce_small / k_bcast_pull bodies + flag protocol).
(The CTA-pair forms need >= 2 tiles per SM and are exercised by tests/test_conv_gpu.py instead.)
"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200 import ops, optim  # noqa: E402
from pytorch_distributed_template_b200.ops import _native, conv_native  # noqa: E402

dev = torch.device("cuda", 0)
CL = torch.channels_last


def nhwc(*s, dtype=torch.bfloat16):
    return torch.randn(*s, device=dev).to(dtype).contiguous(memory_format=CL)


def close(a, b, tol, what):
    err = (a.float() - b.float()).abs().max().item() / max(1.0, b.float().abs().max().item())
    assert err <= tol, (what, err)
    print("ok {:28s} err {:.2e}".format(what, err), flush=True)


def conv_cases():
    for (n, c, h, k, r, s, p) in ((2, 64, 14, 256, 3, 1, 1), (2, 64, 16, 128, 3, 2, 1), (2, 128, 8, 64, 1, 1, 0)):
        x = nhwc(n, c, h, h)
        w = (torch.randn(k, c, r, r, device=dev) / (c * r * r) ** 0.5).to(torch.bfloat16).contiguous(memory_format=CL)
        y = conv_native.fprop(x, w, s, p)
        close(y, F.conv2d(x.float(), w.float(), None, s, p), 2e-2, "igemm fprop {}x{} s{}".format(c, k, s))
        dy = torch.randn_like(y)
        ref = torch.ops.aten.convolution_backward(dy.float(), x.float(), w.float(), None, (s, s), (p, p), (1, 1), False, (0, 0), 1,
                                                  (True, True, False))
        dw = conv_native.wgrad(x, dy, w.shape, s, p)
        close(dw, ref[1], 2e-2, "wgrad {}x{} s{}".format(c, k, s))
        if s == 2:
            dx = conv_native.dgrad_stride2(dy, w, x.shape, p)
            close(dx, ref[0], 2e-2, "dgrad stride 2")
    for (n, c, h, k) in ((2, 64, 20, 64), (2, 128, 12, 128)):
        x = nhwc(n, c, h, h)
        w = (torch.randn(k, c, 3, 3, device=dev) / (c * 9) ** 0.5).to(torch.bfloat16).contiguous(memory_format=CL)
        add = nhwc(n, k, h, h)
        y = conv_native.fprop(x, w, 1, 1, addend=add)
        close(y, F.conv2d(x.float(), w.float(), None, 1, 1) + add.float(), 2e-2, "halo<{}> + addend".format(k))
    x = nhwc(2, 64, 20, 20)
    dy = nhwc(2, 64, 20, 20)
    dw = conv_native.wgrad(x, dy, (64, 64, 3, 3), 1, 1)
    ref = torch.ops.aten.convolution_backward(dy.float(), x.float(), torch.empty(64, 64, 3, 3, device=dev), None, (1, 1), (1, 1),
                                              (1, 1), False, (0, 0), 1, (False, True, False))[1]
    close(dw, ref, 2e-2, "wgrad halo 64x64")


def model_case():
    """Tiny ResNet-18 step under amp: stem (s2d + patch conv), BN / pool kernels, CE, fused SGD."""
    from pytorch_distributed_template_b200 import amp
    from pytorch_distributed_template_b200.models import create_model
    net = create_model("resnet18", num_classes=16).to(dev)
    opt = optim.SGD(net.parameters(), 0.05, momentum=0.9, weight_decay=1e-4)
    crit = ops.CrossEntropyLoss()
    x = torch.randint(0, 256, (4, 3, 64, 64), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 16, (4,), device=dev)
    net.train()
    for dt in (torch.float16, None):
        with amp.autocast(dt is not None, dt or torch.float16):
            loss = crit(net(x), y)
        opt.zero_grad()
        loss.backward()
        opt.step()
    torch.cuda.synchronize()
    assert torch.isfinite(loss).item()
    print("ok resnet18 step (fp16 + fp32)      loss {:.3f}".format(loss.item()), flush=True)


def emu_case():
    lib = _native.lib()
    W, n = 4, 4 * 4 * 16 * 4
    pads = [torch.zeros(lib.pdt_comm_pad_bytes(), dtype=torch.uint8, device=dev) for _ in range(W)]
    grads = [torch.randn(n, device=dev) for _ in range(W)]
    params = [torch.ones(n, device=dev) for _ in range(W)]
    shadow = [torch.zeros(n, dtype=torch.bfloat16, device=dev) for _ in range(W)]
    mom = [torch.zeros(n, device=dev) for _ in range(W)]
    amp_s = []
    for _ in range(W):
        t = torch.zeros(8, device=dev)
        t[0] = 1.0
        amp_s.append(t)
    io = [torch.arange(16, dtype=torch.float32, device=dev) for _ in range(W)]
    e = _native.EmuRanks()
    for k in range(W):
        e.pad[k], e.amp_state[k], e.momentum[k], e.io[k] = pads[k].data_ptr(), amp_s[k].data_ptr(), mom[k].data_ptr(), io[k].data_ptr()
    want = 1.0 - 0.1 * (sum(grads) / W + 1e-4)
    hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device=dev)
    s = torch.cuda.current_stream().cuda_stream
    rc = lib.pdt_emu_step(e, W, _native.make_peers([t.data_ptr() for t in grads]), _native.make_peers([t.data_ptr() for t in params]),
                          _native.make_peers([t.data_ptr() for t in shadow]), _native.make_segments([(0, n)]), hyper.data_ptr(), 2, s)
    assert rc == 0, rc
    bufs = [torch.full((64,), float(k), device=dev) for k in range(W)]
    rc = lib.pdt_emu_small(e, W, 16, 1.0, _native.make_peers([b.data_ptr() for b in bufs]), 256, 1, s)
    assert rc == 0, rc
    torch.cuda.synchronize()
    close(params[2], want, 1e-5, "emulated 4-rank step")
    close(io[3], torch.arange(16, dtype=torch.float32, device=dev) * W, 1e-6, "emulated small all-reduce")


if __name__ == "__main__":
    _native.require()
    torch.manual_seed(0)
    conv_cases()
    model_case()
    emu_case()
    print("SANITIZE_CASES_DONE", flush=True)
