"""T3: the cross-GPU kernel bodies run for ALL ranks inside one cooperative launch on ONE GPU
(rank = blockIdx.y; every rank's arena, pad and state a separate buffer), so shard ownership, the reduction
arithmetic, the inf consensus and the flag protocol are tested on the driver's single-GPU box.  Never emulated
as separate spinning launches (B200_PROFILING.md).  Oracle: plain torch arithmetic of
dist.all_reduce / world + torch.optim.SGD (/root/reference/distributed.py:262-263)."""
import ctypes
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(native_lib, W, seg_lens, dtype_shadow=torch.bfloat16, scale=1.0, seed=0):
    from pytorch_distributed_template_b200.ops import _native
    dev = torch.device("cuda", 0)
    total = sum(seg_lens)
    g = torch.Generator(device=dev).manual_seed(seed)
    pad_bytes = native_lib.pdt_comm_pad_bytes()
    r = {"pads": [torch.zeros(pad_bytes, dtype=torch.uint8, device=dev) for _ in range(W)],
         "grads": [torch.randn(total, device=dev, generator=g) * scale for _ in range(W)],
         "mom": [None] * W, "amp": [], "shadow": [torch.zeros(total, dtype=dtype_shadow, device=dev) for _ in range(W)]}
    p0 = torch.randn(total, device=dev, generator=g)
    m0 = torch.randn(total, device=dev, generator=g) * 0.1
    r["params"] = [p0.clone() for _ in range(W)]
    r["mom"] = [m0.clone() for _ in range(W)]
    for _ in range(W):
        st = torch.zeros(8, dtype=torch.float32, device=dev)
        st[0] = scale
        r["amp"].append(st)
    r["io"] = [torch.arange(16, dtype=torch.float32, device=dev) * (k + 1) for k in range(W)]
    e = _native.EmuRanks()
    for k in range(W):
        e.pad[k], e.amp_state[k], e.momentum[k], e.io[k] = (r["pads"][k].data_ptr(), r["amp"][k].data_ptr(),
                                                            r["mom"][k].data_ptr(), r["io"][k].data_ptr())
    r["emu"], r["p0"], r["m0"] = e, p0, m0
    offs, pos = [], 0
    for n in seg_lens:
        offs.append((pos, n))
        pos += n
    r["segs"] = _native.make_segments(offs)
    r["offs"] = offs
    return r


@pytest.mark.parametrize("W", [2, 3, 4, 5, 8])
def test_reduce_scatter_sgd_allgather_all_ranks_in_one_launch(native_lib, W):
    from pytorch_distributed_template_b200.ops import _native
    unit = 4 * W * 64 // math.gcd(4 * W, 64)
    seg_lens = [unit * 3, unit * 40, unit * 7]
    scale = 1024.0
    r = _setup(native_lib, W, seg_lens, scale=scale, seed=W)
    hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = native_lib.pdt_emu_step(r["emu"], W, _native.make_peers([t.data_ptr() for t in r["grads"]]),
                                 _native.make_peers([t.data_ptr() for t in r["params"]]),
                                 _native.make_peers([t.data_ptr() for t in r["shadow"]]), r["segs"], hyper.data_ptr(), 6, stream)
    assert rc == 0, rc
    grads0 = None
    torch.cuda.synchronize()
    # oracle (fp64 accumulate): mean over ranks of grad / scale, then torch.optim.SGD's arithmetic
    # (the kernel overwrote each rank's own shard of its gradient buffer with the mean, so rebuild the inputs)
    r2 = _setup(native_lib, W, seg_lens, scale=scale, seed=W)
    avg = (sum(t.double() for t in r2["grads"]) / (W * scale)).float()
    d_p = avg + 1e-4 * r2["p0"]
    buf = 0.9 * r2["m0"] + d_p
    expect = r2["p0"] - 0.1 * buf
    for k in range(W):
        assert torch.allclose(r["params"][k], expect, rtol=1e-5, atol=1e-6), (W, k)
        assert torch.equal(r["params"][k], r["params"][0])                       # all ranks bit-identical
        assert torch.equal(r["shadow"][k], r["params"][0].to(torch.bfloat16))    # shadow = rn(bf16) of what was stored
        assert int(r["amp"][k].view(torch.int32)[4]) == 0
        for off, n in r["offs"]:                                                 # momentum lives on the owner's shard only
            sh = n // W
            lo = off + k * sh
            assert torch.allclose(r["mom"][k][lo:lo + sh], buf[lo:lo + sh], rtol=1e-5, atol=1e-6)
            assert torch.allclose(r["grads"][k][lo:lo + sh], avg[lo:lo + sh], rtol=1e-5, atol=1e-7)
    # every pad is back to all-zero flags: the protocol is self-resetting (graph replays need no epoch)
    flag_words = 4 * 256 * 8
    for k in range(W):
        assert int(r["pads"][k].view(torch.int32)[:flag_words].abs().sum()) == 0


def test_inf_on_one_rank_skips_the_step_on_all_ranks(native_lib):
    from pytorch_distributed_template_b200.ops import _native
    W = 4
    r = _setup(native_lib, W, [4 * W * 16 * 8], scale=1.0, seed=3)
    r["grads"][2][5] = float("inf")
    hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device="cuda")
    rc = native_lib.pdt_emu_step(r["emu"], W, _native.make_peers([t.data_ptr() for t in r["grads"]]),
                                 _native.make_peers([t.data_ptr() for t in r["params"]]),
                                 _native.make_peers([t.data_ptr() for t in r["shadow"]]), r["segs"], hyper.data_ptr(), 4,
                                 torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    for k in range(W):
        assert torch.equal(r["params"][k], r["p0"]), k          # nobody stepped
        assert torch.equal(r["mom"][k], r["m0"])
        assert int(r["amp"][k].view(torch.int32)[4]) == 1       # consensus reached every rank
    assert sum(int(r["amp"][k].view(torch.int32)[2]) for k in range(W)) == 1    # only the owner of the shard saw it


@pytest.mark.parametrize("W", [2, 3, 8])
def test_small_allreduce_and_broadcast_all_ranks_in_one_launch(native_lib, W):
    from pytorch_distributed_template_b200.ops import _native
    r = _setup(native_lib, W, [4 * W * 16], seed=1)
    bufs = [torch.full((1024,), float(k), device="cuda") for k in range(W)]
    root = W - 1
    rc = native_lib.pdt_emu_small(r["emu"], W, 16, 1.0 / W, _native.make_peers([b.data_ptr() for b in bufs]), 4096, root,
                                  torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    want = torch.arange(16, dtype=torch.float32, device="cuda") * sum(range(1, W + 1)) / W
    for k in range(W):
        assert torch.allclose(r["io"][k], want, rtol=1e-6)
        assert torch.equal(bufs[k], torch.full((1024,), float(root), device="cuda"))
