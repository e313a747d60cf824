"""T0 (CPU): host-side invariants of the op wrappers that need no GPU."""
import torch
import torch.nn.functional as F


def test_stem_filter_space_to_depth_round_trip_and_equivalence():
    """The [K,3,7,7] -> [K,4,4,16] filter transform (torch formulation, the oracle of the CUDA kernels): the inverse
    recovers the filter, and the 4x4 / 1 convolution over the space-to-depth image equals the 7x7 / 2 convolution
    (/root/reference/distributed.py builds torchvision's conv1: Conv2d(3, 64, 7, 2, 3))."""
    from pytorch_distributed_template_b200.ops import conv_native
    torch.manual_seed(0)
    K = 8
    w = torch.randn(K, 3, 7, 7)
    wp = conv_native.stem_weight_s2d(w)
    assert wp.shape == (K, 4, 4, 16)
    assert torch.count_nonzero(wp[..., 12:]) == 0                      # channel padding
    back = conv_native.stem_weight_grad_from_s2d(wp.reshape(K, 256), K)
    assert torch.equal(back, w)
    # the equivalent convolution: pad 3, space-to-depth 2x2 (dy, dx, c), then 4x4 / 1 with the transformed filter
    x = torch.randn(2, 3, 16, 12)
    xp = F.pad(x, (3, 3, 3, 3))
    n, c, h, wd = xp.shape
    xs = xp.reshape(n, c, h // 2, 2, wd // 2, 2).permute(0, 3, 5, 1, 2, 4).reshape(n, 12, h // 2, wd // 2)
    w44 = wp[..., :12].permute(0, 3, 1, 2).contiguous()               # [K, 12, 4, 4]
    assert torch.allclose(F.conv2d(xs, w44), F.conv2d(x, w, None, 2, 3), atol=1e-4)


def test_fused_pool_gating():
    from pytorch_distributed_template_b200.ops import bn
    x16 = torch.empty(2, 64, 9, 7, dtype=torch.float16)
    x32 = torch.empty(2, 64, 112, 112)
    assert bn._pool_fusable(x16, 2, 2, 0)                              # 16 bit: any window
    assert bn._pool_fusable(x32, 3, 2, 1)                              # fp32: the stem's window over an even map
    assert not bn._pool_fusable(x32, 2, 2, 0)
    assert not bn._pool_fusable(torch.empty(2, 64, 111, 112), 3, 2, 1)


def test_parity_taps_cover_every_filter_tap_once():
    """Stride-2 data gradient: the four parity classes of the input grid partition the filter taps."""
    from pytorch_distributed_template_b200.ops import conv_native
    for ksize, pad in ((3, 1), (1, 0), (7, 3)):
        seen = []
        for a in (0, 1):
            rs, _ = conv_native._parity_taps(56, ksize, pad, a)
            seen += rs
        assert sorted(seen) == list(range(ksize))
