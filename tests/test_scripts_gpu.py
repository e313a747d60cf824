"""T5-lite: the three entrypoints end to end on the GPU with synthetic data (logs, checkpoints, amp,
SyncBN flag, DataParallel over every visible GPU)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ENV = dict(PDT_SYNTHETIC="1", PDT_SYNTH_TRAIN_SIZE="256", PDT_SYNTH_VAL_SIZE="64", PDT_MAX_STEPS="3",
           PDT_OUTPUT_EXISTS="delete", PDT_AMP_DTYPE="bf16")


def _run(cmd, tmp_path):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, **ENV), cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r


def _check(out, logger_lines):
    names = sorted(p.name for p in out.iterdir() if not p.name.startswith("events"))
    # model_best.pth.tar appears only when validation accuracy improved (it can stay 0 on random data)
    assert [n for n in names if n != "model_best.pth.tar"] == ["checkpoint.pth.tar", "experiment.log", "settings.log"], names
    log = (out / "experiment.log").read_text()
    for needle in logger_lines + ["=> creating model: resnet18", "||==> Train epoch: [0/1]", "||==> Val epoch: [0/1]",
                                  "||==> total_time_cost="]:
        assert needle in log, needle
    ck = torch.load(out / "checkpoint.pth.tar", map_location="cpu", weights_only=False)
    assert set(ck) == {"epoch", "arch", "state_dict", "best_acc1"} and len(ck["state_dict"]) == 122
    assert all(torch.isfinite(v.float()).all() for v in ck["state_dict"].values())
    import re
    m = re.search(r"\|\|==> Train epoch: \[0/1\]\tlr=0.100000\tce_loss=([0-9.naninf]+)", log)
    assert m and float(m.group(1)) == float(m.group(1)) and float(m.group(1)) < 50, log[-500:]


def test_distributed_amp_syncbn_script(tmp_path):
    n = min(2, torch.cuda.device_count())
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29681", "distributed_syncBN_amp.py", "--epochs", "1", "-b", str(64 * n), "-j", "2",
           "--outpath", str(tmp_path / "amp"), "--sync_batchnorm", "1", "-p", "1"]
    _run(cmd, tmp_path)
    _check(tmp_path / "amp_resnet18", ["=> using sync BN", "lr_scheduler: SGD MultiStepLR !!!"])


def test_distributed_fp32_script(tmp_path):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "1", "--master-addr",
           "127.0.0.1", "--master-port", "29682", "distributed.py", "--epochs", "1", "-b", "32", "-j", "0",
           "--outpath", str(tmp_path / "ddp")]
    _run(cmd, tmp_path)
    _check(tmp_path / "ddp_resnet18", [])


def test_dataparallel_script(tmp_path):
    cmd = [sys.executable, "dataparallel.py", "--epochs", "1", "-b", "32", "-j", "0", "--outpath", str(tmp_path / "dp")]
    _run(cmd, tmp_path)
    _check(tmp_path / "dp_resnet18", [])
