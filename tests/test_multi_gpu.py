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
