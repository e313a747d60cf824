"""T2: tcgen05 implicit-GEMM convolution vs F.conv2d (fp32, TF32 off) on every ResNet-18 conv shape
the kernel serves, forward and stride-1 data gradient."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# (Cin, Cout, H, k, stride, pad) — SURVEY §2.4(a)
SHAPES = [(64, 64, 56, 3, 1, 1), (64, 128, 56, 3, 2, 1), (128, 128, 28, 3, 1, 1), (64, 128, 56, 1, 2, 0),
          (128, 256, 28, 3, 2, 1), (256, 256, 14, 3, 1, 1), (128, 256, 28, 1, 2, 0), (256, 512, 14, 3, 2, 1),
          (512, 512, 7, 3, 1, 1), (256, 512, 14, 1, 2, 0)]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("cin,cout,hw,k,stride,pad", SHAPES)
@pytest.mark.parametrize("batch", [3, 10])
def test_conv_fprop_matches_conv2d(native_lib, dtype, cin, cout, hw, k, stride, pad, batch):
    from pytorch_distributed_template_b200.ops import conv_native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    dev = torch.device("cuda", 0)
    x = torch.randn(batch, cin, hw, hw, device=dev).to(dtype).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    y = conv_native.fprop(x, w, stride, pad)
    ref = F.conv2d(x.float(), w.float(), None, stride, pad)
    assert y.shape == ref.shape and y.is_contiguous(memory_format=torch.channels_last)
    tol = 2e-2 if dtype == torch.bfloat16 else 3e-3
    err = (y.float() - ref).abs().max().item()
    assert err <= tol * max(1.0, ref.abs().max().item()), (err, ref.abs().max().item())


@pytest.mark.parametrize("cin,cout,hw,k,stride,pad", [s for s in SHAPES if s[4] == 1] + [(64, 64, 56, 3, 2, 1)])
def test_conv_autograd_matches_torch(native_lib, cin, cout, hw, k, stride, pad):
    from pytorch_distributed_template_b200.ops import conv_native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(1)
    dev = torch.device("cuda", 0)
    dtype = torch.bfloat16
    x = torch.randn(4, cin, hw, hw, device=dev).to(dtype).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    wm = (torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    w_lp = wm.detach().to(dtype)
    y = conv_native.conv_native(x, wm, w_lp, (stride, stride), (pad, pad))
    assert y is not None
    dy = torch.randn_like(y)
    y.backward(dy)
    xf = x.detach().float().requires_grad_(True)
    wf = w_lp.float().requires_grad_(True)
    yf = F.conv2d(xf, wf, None, stride, pad)
    yf.backward(dy.float())
    assert (x.grad.float() - xf.grad).abs().max().item() <= 2e-2 * max(1.0, xf.grad.abs().max().item())
    assert wm.grad.dtype == torch.float32
    assert (wm.grad - wf.grad).abs().max().item() <= 3e-2 * max(1.0, wf.grad.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("cin,cout,hw,k,stride,pad", SHAPES)
@pytest.mark.parametrize("batch", [2, 9])
def test_conv_wgrad_matches_autograd(native_lib, dtype, cin, cout, hw, k, stride, pad, batch):
    from pytorch_distributed_template_b200.ops import conv_native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(2)
    dev = torch.device("cuda", 0)
    x = torch.randn(batch, cin, hw, hw, device=dev).to(dtype).contiguous(memory_format=torch.channels_last)
    P = (hw + 2 * pad - k) // stride + 1
    dy = (torch.randn(batch, cout, P, P, device=dev) / 8).to(dtype).contiguous(memory_format=torch.channels_last)
    dw = conv_native.wgrad(x, dy, (cout, cin, k, k), stride, pad)
    wf = torch.zeros(cout, cin, k, k, device=dev, requires_grad=True)
    F.conv2d(x.float(), wf, None, stride, pad).backward(dy.float())
    assert dw.dtype == torch.float32 and dw.shape == wf.grad.shape
    err = (dw - wf.grad).abs().max().item()
    assert err <= 2e-3 * max(1.0, wf.grad.abs().max().item()), (err, wf.grad.abs().max().item())


@pytest.mark.parametrize("cin,cout,hw,k,stride,pad", [s for s in SHAPES if s[4] == 2])
def test_strided_dgrad_native(native_lib, cin, cout, hw, k, stride, pad):
    from pytorch_distributed_template_b200.ops import conv_native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(4)
    dev = torch.device("cuda", 0)
    dtype = torch.bfloat16
    w = (torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    P = (hw + 2 * pad - k) // stride + 1
    dy = torch.randn(3, cout, P, P, device=dev).to(dtype).contiguous(memory_format=torch.channels_last)
    dx = conv_native.dgrad_stride2(dy, w, (3, cin, hw, hw), pad)
    assert dx is not None
    xf = torch.zeros(3, cin, hw, hw, device=dev, requires_grad=True)
    F.conv2d(xf, w.float(), None, stride, pad).backward(dy.float())
    assert (dx.float() - xf.grad).abs().max().item() <= 2e-2 * max(1.0, xf.grad.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("batch,hw", [(4, 224), (3, 64)])
def test_stem_space_to_depth(native_lib, dtype, batch, hw):
    from pytorch_distributed_template_b200.ops import conv_native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(5)
    dev = torch.device("cuda", 0)
    conv = torch.nn.Conv2d(3, 64, 7, 2, 3, bias=False).to(dev)
    x = torch.randn(batch, 3, hw, hw, device=dev)
    assert conv_native.stem_supported(x, conv)
    w_lp = conv.weight.detach().to(dtype)
    y = conv_native.stem_native(x, conv, w_lp)
    ref = F.conv2d(x.to(dtype).float(), w_lp.float(), None, 2, 3)
    tol = 2e-2 if dtype == torch.bfloat16 else 3e-3
    assert y.shape == ref.shape
    assert (y.float() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())
    dy = (torch.randn_like(ref) / 16).to(dtype).contiguous(memory_format=torch.channels_last)
    y.backward(dy)
    wf = w_lp.float().requires_grad_(True)
    F.conv2d(x.to(dtype).float(), wf, None, 2, 3).backward(dy.float())
    assert conv.weight.grad.shape == (64, 3, 7, 7)
    assert (conv.weight.grad - wf.grad).abs().max().item() <= 3e-3 * max(1.0, wf.grad.abs().max().item())


@pytest.mark.parametrize("shape", [(5, 68, 64), (2, 224, 160), (150, 32, 32)])
def test_stem_weight_gradient_patch_kernel(native_lib, shape):
    """k_stem_wgrad (shifted MN-major views of the space-to-depth rows) vs autograd; 68 rows: a partial last tile."""
    from pytorch_distributed_template_b200.ops import conv_native, _native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(15)
    dev = torch.device("cuda", 0)
    n, h, w = shape
    assert _native.lib().pdt_stem_wgrad_supported(_native.dtype_code(torch.float16), n, (h + 6) // 2, (w + 6) // 2, 64) == 1
    conv = torch.nn.Conv2d(3, 64, 7, 2, 3, bias=False).to(dev)
    x = torch.randn(n, 3, h, w, device=dev)
    w_lp = conv.weight.detach().to(torch.float16)
    y = conv_native.stem_native(x, conv, w_lp)
    dy = (torch.randn_like(y.float()) / 16).to(torch.float16).contiguous(memory_format=torch.channels_last)
    y.backward(dy)
    wf = w_lp.float().requires_grad_(True)
    F.conv2d(x.half().float(), wf, None, 2, 3).backward(dy.float())
    assert (conv.weight.grad - wf.grad).abs().max().item() <= 3e-3 * max(1.0, wf.grad.abs().max().item())


@pytest.mark.parametrize("B", [150, 37])
def test_linear_native(native_lib, B):
    from pytorch_distributed_template_b200.ops import conv_native
    torch.manual_seed(6)
    dev = torch.device("cuda", 0)
    dtype = torch.bfloat16
    fc = torch.nn.Linear(512, 1000).to(dev)
    x = torch.randn(B, 512, device=dev).to(dtype).requires_grad_(True)
    assert conv_native.linear_supported(x, fc)
    w_lp = fc.weight.detach().to(dtype)
    y = conv_native.linear_native(x, fc, w_lp)
    xf = x.detach().float().requires_grad_(True)
    wf = w_lp.float().requires_grad_(True)
    bf = fc.bias.detach().clone().requires_grad_(True)
    ref = F.linear(xf, wf, bf)
    assert (y.float() - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    dy = (torch.randn(B, 1000, device=dev) / 8).to(dtype)
    y.backward(dy)
    ref.backward(dy.float())
    assert (x.grad.float() - xf.grad).abs().max().item() <= 2e-2 * max(1.0, xf.grad.abs().max().item())
    assert (fc.weight.grad - wf.grad).abs().max().item() <= 3e-3 * max(1.0, wf.grad.abs().max().item())
    assert torch.allclose(fc.bias.grad, bf.grad, rtol=1e-3, atol=1e-3)


# ---- fp32 activations on tcgen05 kind::tf32 (the DataParallel / plain DDP modes) -------------------------
TF32_TOL = 4e-3      # TF32 keeps 10 mantissa bits; cuDNN with allow_tf32 has the same error


@pytest.mark.parametrize("cin,cout,hw,k,stride,pad", SHAPES)
def test_tf32_conv_fprop_and_wgrad(native_lib, cin, cout, hw, k, stride, pad):
    """fprop (K-major tiles) and wgrad (MN-major tiles, 32-byte-atom swizzle) on kind::tf32."""
    from pytorch_distributed_template_b200.ops import conv_native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(10)
    dev = torch.device("cuda", 0)
    x = torch.randn(5, cin, hw, hw, device=dev).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5).contiguous(memory_format=torch.channels_last)
    y = conv_native.fprop(x, w, stride, pad)
    ref = F.conv2d(x, w, None, stride, pad)
    assert y.dtype == torch.float32 and y.shape == ref.shape
    assert (y - ref).abs().max().item() <= TF32_TOL * max(1.0, ref.abs().max().item())
    dy = (torch.randn_like(ref) / 8).contiguous(memory_format=torch.channels_last)
    wf = w.clone().requires_grad_(True)
    F.conv2d(x, wf, None, stride, pad).backward(dy)
    dw = conv_native.wgrad(x, dy, (cout, cin, k, k), stride, pad)
    err = (dw - wf.grad).abs().max().item() / max(1.0, wf.grad.abs().max().item())
    assert err <= TF32_TOL, err


def test_tf32_model_paths(native_lib):
    """fp32 ResNet-18 step (stem, strided dgrad, linear) through the kind::tf32 kernels against torch fp32 (TF32
    off) on a well-conditioned batch (64 images).  Calibrated bound: cuDNN's own TF32 gradients are measured
    against the same fp32 truth (measured on B200: 10.4 % in the 2-norm after 20 BatchNorm layers — TF32 itself,
    not an implementation, sets that scale) and ours may be at most 1.3x as far away."""
    import copy
    import torchvision
    from pytorch_distributed_template_b200.models import create_model
    from pytorch_distributed_template_b200.ops import _native
    torch.manual_seed(12)
    dev = torch.device("cuda", 0)
    net = create_model("resnet18", num_classes=16).to(dev)
    tv = torchvision.models.resnet18(num_classes=16).to(dev)
    tv.load_state_dict(net.state_dict())
    tv_tf32 = copy.deepcopy(tv)
    x = torch.randn(64, 3, 128, 128, device=dev)
    y = torch.randint(0, 16, (64,), device=dev)
    net.train(); tv.train(); tv_tf32.train()
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    n0 = _native.launch_count
    lo = F.cross_entropy(net(x), y)
    lo.backward()
    assert _native.launch_count - n0 > 60, "tf32 native path not taken"
    lt = F.cross_entropy(tv_tf32(x), y)
    lt.backward()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    lr = F.cross_entropy(tv(x), y)
    lr.backward()
    assert lo.item() == pytest.approx(lr.item(), rel=2e-3)
    flat = lambda m: torch.cat([p.grad.reshape(-1) for p in m.parameters()])
    g_ours, g_torch, g_true = flat(net), flat(tv_tf32), flat(tv)
    assert torch.isfinite(g_ours).all()
    e_ours = ((g_ours - g_true).norm() / g_true.norm()).item()
    e_torch = ((g_torch - g_true).norm() / g_true.norm()).item()
    assert e_ours <= 1.3 * e_torch + 1e-3, (e_ours, e_torch)
    assert e_ours <= 0.2, (e_ours, e_torch)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("cin,cout,hw,batch", [(64, 64, 56, 5), (128, 128, 28, 7), (64, 64, 56, 1), (128, 128, 28, 1),
                                               (64, 128, 28, 3), (128, 64, 56, 2), (64, 64, 20, 3),
                                               (128, 128, 28, 43), (64, 128, 28, 44),    # >= 2 tiles per SM: CTA-pair multicast
                                               (64, 64, 56, 11)])                        # the same for the fp32 N = 64 form
def test_conv3x3_halo_kernel(native_lib, dtype, cin, cout, hw, batch):
    """Patch-reuse kernel (row-shifted smem descriptors) vs F.conv2d, with and without fused BN statistics
    (fp32: kind::tf32 over 32-channel chunks, streamed filter taps)."""
    from pytorch_distributed_template_b200.ops import conv_native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(20)
    dev = torch.device("cuda", 0)
    x = torch.randn(batch, cin, hw, hw, device=dev).to(dtype).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, 3, 3, device=dev) / (cin * 9) ** 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    assert conv_native._halo_ok(x, w, 1, 1)
    y = conv_native.fprop(x, w, 1, 1)
    ref = F.conv2d(x.float(), w.float(), None, 1, 1)
    tol = 2e-2 if dtype == torch.bfloat16 else 3e-3          # (tf32 keeps 10 mantissa bits, like fp16)
    assert (y.float() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())
    rm, rv = torch.zeros(cout, device=dev), torch.ones(cout, device=dev)
    y2, mean, invstd = conv_native.fprop_bnstats(x, w, 1, 1, rm, rv, 0.1, 1e-5)
    assert torch.equal(y2, y)
    assert torch.allclose(mean, ref.mean((0, 2, 3)), rtol=1e-2, atol=2e-3)
    assert torch.allclose(invstd, torch.rsqrt(ref.var((0, 2, 3), unbiased=False) + 1e-5), rtol=1e-2, atol=1e-3)
    assert torch.allclose(rm, 0.1 * ref.mean((0, 2, 3)), rtol=1e-2, atol=2e-3)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("hw,batch", [(56, 5), (20, 3), (31, 2), (8, 150), (55, 1)])
def test_conv3x3_wgrad_halo_kernel(native_lib, dtype, hw, batch):
    """64->64 3x3 weight gradient with patch reuse (row-shifted MN-major views, tap pairs through the
    descriptor's leading-dimension offset; fp32: kind::tf32, four views per instruction over two channel planes)
    vs autograd; includes partial last tiles (31, 55 rows)."""
    from pytorch_distributed_template_b200 import ops
    from pytorch_distributed_template_b200.ops import conv_native, _native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(21)
    dev = torch.device("cuda", 0)
    assert _native.lib().pdt_conv3x3_wgrad_halo_supported(_native.dtype_code(dtype), batch, hw, hw, 64, 64) == 1
    x = torch.randn(batch, 64, hw, hw, device=dev).to(dtype).contiguous(memory_format=torch.channels_last)
    dy = (torch.randn(batch, 64, hw, hw, device=dev) / 8).to(dtype).contiguous(memory_format=torch.channels_last)
    dw = conv_native.wgrad(x, dy, (64, 64, 3, 3), 1, 1)
    wf = torch.zeros(64, 64, 3, 3, device=dev, requires_grad=True)
    F.conv2d(x.float(), wf, None, 1, 1).backward(dy.float())
    err = (dw - wf.grad).abs().max().item()
    tol = 3e-3 if dtype == torch.float32 else 2e-3          # tf32 keeps 10 mantissa bits of fp32 inputs
    assert err <= tol * max(1.0, wf.grad.abs().max().item()), (err, wf.grad.abs().max().item())
    # accumulates into its destination
    dw2 = conv_native.wgrad(x, dy, (64, 64, 3, 3), 1, 1, out=dw.clone())
    assert (dw2 - 2 * wf.grad).abs().max().item() <= 2 * tol * max(1.0, wf.grad.abs().max().item())


@pytest.mark.parametrize("cin,cout,hw,k,stride,pad,batch", [(256, 256, 14, 3, 1, 1, 100), (512, 512, 7, 3, 1, 1, 400),
                                                             (128, 256, 28, 3, 2, 1, 101), (256, 512, 14, 1, 2, 0, 397)])
def test_conv_cluster_multicast(native_lib, cin, cout, hw, k, stride, pad, batch):
    """>= 148 M tiles: the 128/256-wide forward kernel runs as clusters of 4 CTAs that share one multicast
    filter tile per k-block (includes M-tile counts that are not a multiple of 4: phantom tiles)."""
    from pytorch_distributed_template_b200.ops import conv_native
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(31)
    dev = torch.device("cuda", 0)
    dtype = torch.bfloat16
    x = torch.randn(batch, cin, hw, hw, device=dev).to(dtype).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    P = (hw + 2 * pad - k) // stride + 1
    assert (batch * P * P + 127) // 128 >= 148
    rm, rv = torch.zeros(cout, device=dev), torch.ones(cout, device=dev)
    y, mean, invstd = conv_native.fprop_bnstats(x, w, stride, pad, rm, rv, 0.1, 1e-5)
    ref = F.conv2d(x.float(), w.float(), None, stride, pad)
    assert (y.float() - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    assert torch.allclose(mean, ref.mean((0, 2, 3)), rtol=1e-2, atol=2e-3)
    assert torch.allclose(invstd, torch.rsqrt(ref.var((0, 2, 3), unbiased=False) + 1e-5), rtol=1e-2, atol=1e-3)
    y2 = conv_native.fprop(x, w, stride, pad)
    assert torch.equal(y2, y)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_stem_uint8_input_fused_normalisation(native_lib, dtype):
    """Raw uint8 images through the native stem (normalisation fused into the layout kernel) vs the same
    stem on host-style normalised fp32 images; forward and weight gradient."""
    from pytorch_distributed_template_b200 import amp, ops
    from pytorch_distributed_template_b200.ops import input_norm
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(33)
    dev = torch.device("cuda", 0)
    conv = torch.nn.Conv2d(3, 64, 7, 2, 3, bias=False).to(dev)
    x8 = torch.randint(0, 256, (6, 3, 64, 64), dtype=torch.uint8, device=dev)
    xf = input_norm.normalize_u8(x8)
    mean, std = input_norm.get_input_norm()
    ref = (x8.float() / 255.0 - torch.tensor(mean, device=dev).view(1, 3, 1, 1)) / torch.tensor(std, device=dev).view(1, 3, 1, 1)
    assert torch.allclose(xf, ref, rtol=1e-5, atol=1e-5)
    with amp.autocast(dtype=dtype):
        y8 = ops.conv2d(x8, conv)
        y8.float().square().mean().backward()
        g8 = conv.weight.grad.clone()
        conv.weight.grad = None
        yf = ops.conv2d(xf, conv)
        yf.float().square().mean().backward()
        gf = conv.weight.grad.clone()
    assert y8.dtype == dtype and y8.shape == yf.shape
    tol = 2e-2 if dtype == torch.bfloat16 else 3e-3
    assert (y8.float() - yf.float()).abs().max().item() <= tol * max(1.0, yf.float().abs().max().item())
    assert (g8 - gf).abs().max().item() <= tol * max(1e-3, gf.abs().max().item())
    oracle = F.conv2d(ref, conv.weight.detach(), None, 2, 3)
    assert (y8.float() - oracle).abs().max().item() <= 2 * tol * max(1.0, oracle.abs().max().item())


def _dev():
    return torch.device("cuda", 0)


def _nhwc(*shape, dtype=torch.bfloat16):
    return torch.randn(*shape, device=_dev()).to(dtype).contiguous(memory_format=torch.channels_last)


TOL = {torch.bfloat16: 2e-2, torch.float16: 4e-3}


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("N,C,H,K,R,pad", [(6, 64, 56, 64, 3, 1),      # patch-reuse kernel, TMA-store epilogue
                                           (5, 128, 28, 128, 3, 1),    # patch-reuse kernel, direct-store epilogue
                                           (44, 128, 28, 128, 3, 1),   # ... with CTA pairs
                                           (7, 256, 14, 256, 3, 1),    # implicit-GEMM kernel
                                           (9, 512, 7, 512, 3, 1), (4, 256, 14, 64, 1, 0)])
def test_conv_epilogue_addend(native_lib, dtype, N, C, H, K, R, pad):
    """y = conv(x, w) + addend in the kernels' epilogues (the residual-gradient sum of a block's backward) vs
    F.conv2d + add in fp32."""
    from pytorch_distributed_template_b200.ops import conv_native
    torch.manual_seed(3)
    x = _nhwc(N, C, H, H, dtype=dtype)
    w = (torch.randn(K, C, R, R, device=_dev()) / (C * R * R) ** 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    add = _nhwc(N, K, H, H, dtype=dtype)
    y = conv_native.fprop(x, w, 1, pad, addend=add)
    y0 = conv_native.fprop(x, w, 1, pad)
    ref = F.conv2d(x.float(), w.float(), padding=pad) + add.float()
    tol = TOL[dtype]
    assert (y.float() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())
    # and it is the plain result plus the addend up to one rounding of the sum
    assert (y.float() - (y0.float() + add.float())).abs().max().item() <= tol * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_strided_dgrad_with_addend(native_lib, dtype):
    from pytorch_distributed_template_b200.ops import conv_native
    torch.manual_seed(5)
    N, C, H, K = 6, 64, 56, 128
    row = 32 if dtype == torch.float32 else 64
    assert C % row == 0
    w = (torch.randn(K, C, 3, 3, device=_dev()) / (C * 9) ** 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    dy = _nhwc(N, K, H // 2, H // 2, dtype=dtype)
    add = _nhwc(N, C, H, H, dtype=dtype)
    dx = conv_native.dgrad_stride2(dy, w, (N, C, H, H), 1, None, addend=add)
    assert dx is not None
    ref = torch.ops.aten.convolution_backward(dy.float(), torch.empty(N, C, H, H, device=_dev()), w.float(), None, (2, 2), (1, 1),
                                              (1, 1), False, (0, 0), 1, (True, False, False))[0] + add.float()
    tol = TOL[dtype] if dtype != torch.float32 else TF32_TOL
    assert (dx.float() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())


def test_residual_gradient_link_matches_unfused_backward(native_lib, monkeypatch):
    """Whole ResNet-18 backward with the shortcut gradients added in the data-gradient epilogues (GradLink) against
    the same network with autograd's separate adds: same gradients up to one 16-bit rounding per block, and no
    elementwise add kernel left for the eight blocks."""
    from pytorch_distributed_template_b200 import amp
    from pytorch_distributed_template_b200.models import create_model
    torch.manual_seed(21)
    dev = _dev()
    net = create_model("resnet18", num_classes=32).to(dev)
    net.train()
    x = torch.randn(32, 3, 96, 96, device=dev)
    y = torch.randint(0, 32, (32,), device=dev)
    grads = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("PDT_RESADD_FUSION", mode)
        net.zero_grad()
        torch.manual_seed(0)
        for m in net.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.reset_running_stats()
        with amp.autocast(True, torch.float16):
            loss = F.cross_entropy(net(x).float(), y)
        loss.backward()
        torch.cuda.synchronize()
        grads[mode] = torch.cat([p.grad.reshape(-1).clone() for p in net.parameters()])
    rel = ((grads["1"] - grads["0"]).norm() / grads["0"].norm()).item()
    assert rel < 2e-2, rel
    assert torch.isfinite(grads["1"]).all()


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("N,C,H,K,R,stride,pad", [(101, 256, 14, 256, 3, 1, 1),    # 155 M tiles: odd, the last pair has one empty half
                                                  (100, 128, 28, 256, 3, 2, 1),    # strided, 128 -> 256
                                                  (400, 512, 7, 512, 3, 1, 1),     # two N tiles per M tile
                                                  (160, 256, 14, 128, 3, 1, 1)])   # BN = 128 pairs
def test_conv_cta_pair_tiles(native_lib, dtype, N, C, H, K, R, stride, pad):
    """Shapes large enough (>= 148 M tiles, >= 9 k-blocks, N tile >= 128) to run as CTA pairs: one M = 256
    tcgen05.mma.cta_group::2 per k-step, each CTA holding half of the filter tile (k_conv_igemm<.., PAIR>)."""
    from pytorch_distributed_template_b200.ops import conv_native
    torch.manual_seed(7)
    x = _nhwc(N, C, H, H, dtype=dtype)
    w = (torch.randn(K, C, R, R, device=_dev()) / (C * R * R) ** 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    y = conv_native.fprop(x, w, stride, pad)
    tol = TOL[dtype]
    for sl in (slice(0, 24), slice(N - 24, N)):               # first and last images (the tail tiles)
        ref = F.conv2d(x[sl].float(), w.float(), None, stride, pad)
        assert (y[sl].float() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())
    # with the BatchNorm statistics epilogue
    bn = torch.nn.BatchNorm2d(K).to(_dev())
    y2, mean, invstd = conv_native.fprop_bnstats(x, w, stride, pad, bn.running_mean, bn.running_var, 0.1, 1e-5)
    yf = y2.float()
    assert torch.equal(y2, y)
    assert torch.allclose(mean, yf.mean((0, 2, 3)), rtol=2e-3, atol=2e-3)
    assert torch.allclose(invstd, torch.rsqrt(yf.var((0, 2, 3), unbiased=False) + 1e-5), rtol=5e-3, atol=5e-3)


def test_batched_dgrad_filter_layouts(native_lib):
    """One launch re-lays out every data-gradient filter of a model (64 x 64 tiles): against index_select / transpose,
    with channel counts that are not multiples of the tile and tap subsets of a stride-2 gradient."""
    import types
    from pytorch_distributed_template_b200.ops import conv_native
    torch.manual_seed(40)
    dev = torch.device("cuda", 0)
    st = types.SimpleNamespace(weights_version=0)
    cases = [(64, 64, 3, [2, 1, 0], [2, 1, 0]), (128, 64, 3, [1], [2, 0]), (72, 40, 3, [2, 0], [1]), (512, 256, 1, [0], [0]),
             (200, 136, 3, [2, 1, 0], [2, 1, 0]), (8, 8, 3, [0, 2], [0, 2])]
    ws = []
    for (K, C, R, rs, ss) in cases:
        w = torch.randn(K, C, R, R, device=dev).to(torch.float16).contiguous(memory_format=torch.channels_last)
        ws.append(w)
        conv_native.dgrad_weight(w, rs, ss, st)            # registers the job (and lays it out once on its own)
    conv_native.prepare_dgrad_weights(st)                    # all of them in one launch
    for w, (K, C, R, rs, ss) in zip(ws, cases):
        out = conv_native.dgrad_weight(w, rs, ss, st)
        ir, is_ = torch.tensor(rs, device=dev), torch.tensor(ss, device=dev)
        ref = w.index_select(2, ir).index_select(3, is_).transpose(0, 1)
        assert out.shape == ref.shape and torch.equal(out, ref.contiguous(memory_format=torch.channels_last)), (K, C, R)
    # a weight update followed by prepare() refreshes every buffer
    for w in ws:
        w.mul_(2)
    st.weights_version = 1
    conv_native.prepare_dgrad_weights(st)
    for w, (K, C, R, rs, ss) in zip(ws, cases):
        out = conv_native.dgrad_weight(w, rs, ss, st)
        ref = w.index_select(2, torch.tensor(rs, device=dev)).index_select(3, torch.tensor(ss, device=dev)).transpose(0, 1)
        assert torch.equal(out, ref.contiguous(memory_format=torch.channels_last))


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
def test_stem_filter_space_to_depth_kernels(native_lib, dtype, monkeypatch):
    """k_stem_weight_s2d / k_stem_weight_grad_unpack against the pad / reshape / permute formulation, for a
    channels_last and a plain-contiguous filter."""
    from pytorch_distributed_template_b200.ops import conv_native
    torch.manual_seed(41)
    dev = torch.device("cuda", 0)
    for fmt in (torch.channels_last, torch.contiguous_format):
        w = torch.randn(64, 3, 7, 7, device=dev).to(dtype).contiguous(memory_format=fmt)
        got = conv_native.stem_weight_s2d(w)
        monkeypatch.setenv("PDT_STEM_WEIGHT", "torch")
        want = conv_native.stem_weight_s2d(w)
        monkeypatch.setenv("PDT_STEM_WEIGHT", "native")
        assert got.shape == want.shape and torch.equal(got, want)
    dwp = torch.randn(64, 256, device=dev)
    got = conv_native.stem_weight_grad_from_s2d(dwp, 64)
    monkeypatch.setenv("PDT_STEM_WEIGHT", "torch")
    want = conv_native.stem_weight_grad_from_s2d(dwp, 64)
    assert got.shape == (64, 3, 7, 7) and torch.equal(got, want.contiguous())
