"""Deterministic weight gradients (the reference's --seed turns on cudnn.deterministic,
/root/reference/distributed.py:91-99): with PDT_DETERMINISTIC=1 split-K partials are folded in a fixed order, so
repeated launches are bit-identical; the default path accumulates with red.global.add and need not be.
Oracle for the values: aten convolution_backward in fp32."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(16, 64, 56, 56, 64, 3, 1, 1), (8, 128, 28, 28, 256, 3, 2, 1), (8, 256, 14, 14, 512, 1, 2, 0),
                                   (4, 512, 7, 7, 512, 3, 1, 1)])
def test_wgrad_is_bit_reproducible_and_correct(native_lib, monkeypatch, shape):
    monkeypatch.setenv("PDT_DETERMINISTIC", "1")
    from pytorch_distributed_template_b200.ops import conv_native
    N, C, H, W, K, R, stride, pad = shape
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(N, C, H, W, device="cuda", generator=g).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    P = (H + 2 * pad - R) // stride + 1
    dy = torch.randn(N, K, P, P, device="cuda", generator=g).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    outs = [conv_native.wgrad(x, dy, (K, C, R, R), stride, pad) for _ in range(3)]
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    ref = torch.ops.aten.convolution_backward(dy.float(), x.float(), torch.empty(K, C, R, R, device="cuda"), None, (stride, stride),
                                              (pad, pad), (1, 1), False, (0, 0), 1, (False, True, False))[1]
    err = (outs[0] - ref).abs().max() / ref.abs().max()
    assert err < 2e-3, err
    # it accumulates into its destination like the default path
    acc = outs[0].clone()
    conv_native.wgrad(x, dy, (K, C, R, R), stride, pad, out=acc)
    torch.cuda.synchronize()
    assert torch.allclose(acc, 2 * outs[0], rtol=1e-6, atol=1e-6)
