"""T4: cross-GPU kernels on >= 2 real GPUs (fabric, reduce-scatter, SGD all-gather, bcast, SyncBN exchange).
Launched as one process per GPU; never emulated by time-slicing one GPU (B200_PROFILING.md)."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _launch(extra, nproc, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "_ddp_worker.py"), "--cuda"] + extra
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=dict(os.environ, PDT_RESULT_DIR=d))
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
        return [json.load(open(os.path.join(d, "rank{}.json".format(i)))) for i in range(nproc)]


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_ddp_native_fabric_matches_torch_ddp(native_lib):
    n = 8 if _ngpu() >= 8 else (4 if _ngpu() >= 4 else 2)
    outs = _launch([], n, 29661)
    assert len(outs) == n
    for o in outs:
        assert o["fabric"]["native"] and o["fabric"]["symmetric"], o
        assert o["init_synced"] and o["ranks_identical"], o
        assert o["first_param_rel"] < 1e-4, o
        assert o["worst_param_rel"] < 5e-2, o
        assert o["reduce_mean"] == pytest.approx([(n + 1) / 2, n + 1.0])


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_syncbn_native_matches_torch_syncbn(native_lib):
    outs = _launch(["--syncbn"], 2, 29662)
    for o in outs:
        assert o["ranks_identical"], o
        assert o["first_param_rel"] < 1e-3, o
        assert o["bn_buffer_rel"] < 1e-3, o


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_dataparallel_matches_torch_dataparallel(native_lib):
    """Single-process DataParallel: peer-copy scatter/gather, per-device reduce + SGD + weight push."""
    import copy
    import torch.nn.functional as F
    import torchvision
    from pytorch_distributed_template_b200 import optim
    from pytorch_distributed_template_b200.models import create_model
    from pytorch_distributed_template_b200.parallel import DataParallel
    n = 2
    torch.manual_seed(21)
    net = create_model("resnet18", num_classes=10)
    tv = torchvision.models.resnet18(num_classes=10)
    tv.load_state_dict(net.state_dict())
    ours = DataParallel(net, device_ids=list(range(n))).cuda()
    theirs = torch.nn.DataParallel(tv, device_ids=list(range(n))).cuda()
    o1 = optim.SGD(ours.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
    o2 = torch.optim.SGD(theirs.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(3)
    first = None
    for it in range(3):
        x = torch.randn(8 * n, 3, 64, 64, generator=g).cuda()
        y = torch.randint(0, 10, (8 * n,), generator=g).cuda()
        for model, opt in ((ours, o1), (theirs, o2)):
            model.train()
            out = model(x)
            assert out.shape == (8 * n, 10) and out.device.index == 0
            loss = F.cross_entropy(out, y)
            opt.zero_grad()
            loss.backward()
            opt.step()
        for d in range(n):
            torch.cuda.synchronize(d)
        worst = max(((p.detach() - q.detach()).abs().max() / (q.detach().abs().max() + 1e-12)).item()
                    for p, q in zip(ours.module.parameters(), theirs.module.parameters()))
        first = worst if first is None else first
    assert first < 1e-4, first
    assert worst < 5e-2, worst
    # replicas hold the same weights as the module (the optimizer kernel pushed them)
    for rep in ours.replicas[1:]:
        for p, q in zip(rep.parameters(), ours.module.parameters()):
            assert torch.equal(p.detach().cpu(), q.detach().cpu())
    assert all(k.startswith("module.") for k in ours.state_dict())
    ours.eval()
    with torch.no_grad():
        assert ours(torch.randn(4 * n, 3, 64, 64).cuda()).shape == (4 * n, 10)
