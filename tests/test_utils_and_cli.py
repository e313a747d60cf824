"""T0: CLI / utils / schedule / sampler / checkpoint parity with the reference."""
import argparse
import importlib.util
import os
import sys

import pytest
import torch

from pytorch_distributed_template_b200 import config, data, optim, utils

REF = "/root/reference"
HAVE_REF = os.path.exists(os.path.join(REF, "utils.py"))


def _ref_utils():
    spec = importlib.util.spec_from_file_location("ref_utils", os.path.join(REF, "utils.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


COMMON = {
    "data": "/mnt/cephfs/mixed/dataset/imagenet/", "arch": "resnet18", "workers": 8, "epochs": 5, "step": [3, 4],
    "start_epoch": 0, "batch_size": 1200, "lr": 0.1, "momentum": 0.9, "weight_decay": 1e-4, "print_freq": 10,
    "evaluate": False, "pretrained": False, "seed": None, "lr_scheduler": "steplr", "gamma": 0.1,
}


@pytest.mark.parametrize("mode,out,gpus", [("dp", "./output", "5,6,7"), ("ddp", "./output_ddp_test", "0,1,2"),
                                          ("ddp_amp", "./output_ddp_amp", "0,1,2")])
def test_parser_defaults(mode, out, gpus):
    a = config.build_parser(mode).parse_args([])
    for k, v in COMMON.items():
        assert getattr(a, k) == v, k
    assert a.outpath == out and a.gpus == gpus
    assert hasattr(a, "local_rank") == (mode != "dp")
    assert hasattr(a, "use_amp") == (mode == "ddp_amp")
    if mode == "ddp_amp":
        assert a.use_amp is True and a.sync_batchnorm is False


def test_parser_quirks_and_aliases():
    p = config.build_parser("ddp_amp")
    assert p.parse_args(["--evaluate", "False"]).evaluate is True          # type=bool quirk kept
    assert p.parse_args(["--use_amp", ""]).use_amp is False
    assert p.parse_args(["--local_rank=3"]).local_rank == 3
    assert p.parse_args(["--local-rank=2"]).local_rank == 2                  # torch>=2 launcher spelling (Q1)
    a = p.parse_args(["-a", "resnet50", "-j", "2", "-b", "64", "--lr", "0.2", "--wd", "1e-3", "-p", "5",
                      "--learning-rate", "0.3", "--weight-decay", "2e-3", "--start-epoch", "1", "--lr-scheduler", "x"])
    assert (a.arch, a.workers, a.batch_size, a.lr, a.weight_decay, a.print_freq, a.start_epoch, a.lr_scheduler) == \
        ("resnet50", 2, 64, 0.3, 2e-3, 5, 1, "x")
    assert config.normalise_steps("[30, 60]") == [30, 60] and config.normalise_steps([3, 4]) == [3, 4]
    with pytest.raises(SystemExit):
        p.parse_args(["-a", "not_a_model"])


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not mounted")
def test_parser_matches_reference_actions():
    """Every option string / dest / default / type of the reference parsers exists here."""
    for mode, script in [("dp", "dataparallel"), ("ddp", "distributed"), ("ddp_amp", "distributed_syncBN_amp")]:
        src = open(os.path.join(REF, script + ".py")).read()
        ours = {s: a for a in config.build_parser(mode)._actions for s in a.option_strings}
        import re
        for m in re.finditer(r"parser\.add_argument\(([^\n]*(?:\n\s+['\"][^\n]*)*)\)", src):
            flags = re.findall(r"'(--?[A-Za-z_\-]+)'", m.group(1).split("default")[0])
            for f in flags:
                assert f in ours, (script, f)


def test_lr_schedule_closed_form():
    p = torch.nn.Parameter(torch.zeros(4))
    opt = optim.SGD([p], 0.1, momentum=0.9, weight_decay=1e-4)
    sched = optim.MultiStepLR(opt, milestones=[3, 4], gamma=0.1)
    got = []
    for e in range(5):
        sched.step(e)
        got.append(utils.get_learning_rate(opt))
    assert got == pytest.approx([0.1, 0.1, 0.1, 0.01, 0.001])


def test_average_meter_and_accuracy_match_reference():
    m = utils.AverageMeter("Loss", ":.4e")
    for v, n in [(1.0, 2), (3.0, 6)]:
        m.update(v, n)
    assert m.avg == pytest.approx(2.5) and m.count == 8 and str(m).startswith("Loss 3.0000e+00")
    scores = torch.tensor([[0.1, 0.9], [0.8, 0.2], [0.3, 0.7]])
    tgt = torch.tensor([1, 1, 1])
    acc = utils.accuracy(scores, tgt, 1)
    assert acc.dim() == 0 and acc.item() == pytest.approx(2 / 3)
    if HAVE_REF:
        ref = _ref_utils()
        assert ref.accuracy(scores, tgt, 1).item() == pytest.approx(acc.item())
        r = ref.AverageMeter("Loss", ":.4e")
        r.update(1.0, 2); r.update(3.0, 6)
        assert str(r) == str(m)
    d = utils.DeviceMeter("x", ":f")
    d.update(torch.tensor(1.0), 2); d.update(3.0, 6)
    assert d.read()["avg"] == pytest.approx(2.5)


def test_logger_settings_output_dir(tmp_path, monkeypatch):
    out = tmp_path / "run"
    utils.output_process(str(out))
    (out / "stale").write_text("x")
    monkeypatch.setenv("PDT_OUTPUT_EXISTS", "quit")
    with pytest.raises(OSError):
        utils.output_process(str(out))
    monkeypatch.setenv("PDT_OUTPUT_EXISTS", "delete")
    utils.output_process(str(out))
    assert not (out / "stale").exists()
    ns = argparse.Namespace(outpath=str(out), lr=0.1, arch="resnet18")
    utils.write_settings(ns)
    assert (out / "settings.log").read_text() == "outpath: {}\nlr: 0.1\narch: resnet18\n".format(out)
    lg = utils.get_logger(str(out), "t")
    utils.ddp_print("hello", lg, 0)
    utils.ddp_print("hidden", None, 1)
    for h in lg.handlers:
        h.flush()
    txt = (out / "experiment.log").read_text()
    assert txt.rstrip().endswith("INFO: hello") and "hidden" not in txt


def test_checkpoint_layout_and_prefix_views(tmp_path):
    from pytorch_distributed_template_b200.models import create_model
    from pytorch_distributed_template_b200.parallel import DataParallel
    net = create_model("resnet18", num_classes=10)
    wrapped = DataParallel(net)
    assert all(k.startswith("module.") for k in wrapped.state_dict())
    sd = wrapped.module.state_dict()
    assert "conv1.weight" in sd and len(sd) == 122
    utils.save_checkpoint({"epoch": 1, "arch": "resnet18", "state_dict": sd, "best_acc1": 0.5}, True, str(tmp_path))
    assert (tmp_path / "checkpoint.pth.tar").exists() and (tmp_path / "model_best.pth.tar").exists()
    import torchvision
    tv = torchvision.models.resnet18(num_classes=10)
    ck = utils.load_checkpoint(str(tmp_path / "checkpoint.pth.tar"), tv)          # ours -> torchvision
    assert set(ck) == {"epoch", "arch", "state_dict", "best_acc1"}
    torch.save({"epoch": 2, "arch": "resnet18", "state_dict": utils.add_module_prefix(tv.state_dict()),
                "best_acc1": 0}, tmp_path / "tv.pth.tar")
    utils.load_checkpoint(str(tmp_path / "tv.pth.tar"), wrapped)                  # torch-DDP-style keys -> ours
    x = torch.randn(2, 3, 32, 32)
    net.eval(); tv.eval()
    assert torch.allclose(net(x), tv(x), atol=1e-5)


@pytest.mark.parametrize("n,world,drop", [(10, 3, False), (10, 4, True), (7, 2, False), (2, 4, False)])
def test_sampler_matches_torch(n, world, drop):
    from torch.utils.data.distributed import DistributedSampler
    ds = list(range(n))
    for rank in range(world):
        for epoch in (0, 3):
            a = data.ShardSampler(ds, world, rank, shuffle=True, seed=0, drop_last=drop)
            b = DistributedSampler(ds, world, rank, shuffle=True, seed=0, drop_last=drop)
            a.set_epoch(epoch); b.set_epoch(epoch)
            assert list(a) == list(b) and len(a) == len(b)


def test_synthetic_dataset_and_model_zoo():
    ds = data.SyntheticImageNet(10, 32, 10)
    x, y = ds[3]
    assert x.shape == (3, 32, 32) and 0 <= y < 10 and torch.equal(ds[3][0], x)
    names = config.model_names()
    assert "resnet18" in names and "vgg16" in names
    from pytorch_distributed_template_b200.models import create_model
    import torchvision
    for arch in ("resnet18", "resnet50"):
        ours, tv = create_model(arch), torchvision.models.__dict__[arch]()
        assert [(k, tuple(v.shape)) for k, v in ours.state_dict().items()] == \
               [(k, tuple(v.shape)) for k, v in tv.state_dict().items()]
    assert type(create_model("squeezenet1_0")).__module__.startswith("torchvision")


def test_uint8_input_pipeline_cpu():
    """Raw-pixel pipeline: synthetic uint8 dataset, device-side normalisation formula, model forward on
    uint8 == forward on the normalised floats (CPU path)."""
    import torch
    from pytorch_distributed_template_b200 import data
    from pytorch_distributed_template_b200.models import build_resnet
    from pytorch_distributed_template_b200.ops import input_norm
    ds = data.SyntheticImageNet(8, image_size=32, num_classes=10, uint8=True)
    img, target = ds[3]
    assert img.dtype == torch.uint8 and img.shape == (3, 32, 32) and 0 <= target < 10
    x8 = torch.stack([ds[i][0] for i in range(4)])
    xf = input_norm.normalize_u8(x8)
    mean = torch.tensor(input_norm.IMAGENET_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(input_norm.IMAGENET_STD).view(1, 3, 1, 1)
    assert torch.allclose(xf, (x8.float() / 255 - mean) / std, rtol=1e-5, atol=1e-5)
    torch.manual_seed(0)
    model = build_resnet("resnet18", num_classes=10).eval()
    assert getattr(model, "accepts_uint8", False)
    with torch.no_grad():
        assert torch.allclose(model(x8), model(xf), rtol=1e-4, atol=1e-4)
    input_norm.set_input_norm((0.5, 0.5, 0.5), (0.25, 0.25, 0.25))
    try:
        assert torch.allclose(input_norm.normalize_u8(x8), (x8.float() / 255 - 0.5) / 0.25, rtol=1e-5, atol=1e-5)
    finally:
        input_norm.set_input_norm(input_norm.IMAGENET_MEAN, input_norm.IMAGENET_STD)


def test_native_library_abi_matches_header():
    """The in-tree library (when built) reports the ABI number of include/pdt_b200.h; a stale build of an
    older checkout is refused with a clear message instead of failing later on a missing symbol."""
    import os
    from pytorch_distributed_template_b200.ops import _native
    want = _native._header_abi()
    assert isinstance(want, int) and want >= 1
    if not os.path.exists(_native.LIB_PATH):
        import pytest
        pytest.skip("library not built")
    lib = _native._load()
    assert lib is not None, _native._load_error
    assert lib.pdt_abi_version() == want


def test_start_sh_and_reduce_mean():
    """start.sh: LF endings, the reference's three runs through torchrun on 127.0.0.1; reduce_mean is importable
    from the scripts like in the reference and is the identity (divided by nprocs) without a process group."""
    import os
    import torch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    raw = open(os.path.join(root, "start.sh"), "rb").read()
    assert b"\r" not in raw
    text = raw.decode()
    for script in ("dataparallel.py", "distributed.py", "distributed_syncBN_amp.py"):
        assert script in text
    assert "torch.distributed.run" in text or "torchrun" in text
    assert "127.0.0.1" in text
    import distributed
    import distributed_syncBN_amp
    from pytorch_distributed_template_b200.utils import reduce_mean
    assert distributed.reduce_mean is reduce_mean and distributed_syncBN_amp.reduce_mean is reduce_mean
    t = torch.tensor([2.0, 4.0])
    out = reduce_mean(t, 2)
    assert torch.equal(out, torch.tensor([1.0, 2.0])) and torch.equal(t, torch.tensor([2.0, 4.0]))


def test_imagefolder_uint8_pipeline_equals_host_normalisation(tmp_path):
    """ImageFolder pipelines (the reference's transforms): the uint8 variant stops before ToTensor/Normalize and
    device-side normalisation of its output equals the host-side float pipeline (deterministic val transform)."""
    import numpy as np
    import torch
    PIL = pytest.importorskip("PIL.Image")
    from pytorch_distributed_template_b200 import data
    from pytorch_distributed_template_b200.ops import input_norm
    rng = np.random.RandomState(0)
    for split in ("train", "val"):
        for cls in ("a", "b"):
            d = tmp_path / split / cls
            d.mkdir(parents=True)
            for i in range(2):
                PIL.fromarray(rng.randint(0, 256, (300, 260, 3), dtype=np.uint8)).save(str(d / ("%d.png" % i)))
    tr8, va8 = data.imagefolder_datasets(str(tmp_path), uint8=True)
    trf, vaf = data.imagefolder_datasets(str(tmp_path), uint8=False)
    assert len(tr8) == len(trf) == 4 and len(va8) == 4
    x8, y8 = va8[1]
    xf, yf = vaf[1]
    assert x8.dtype == torch.uint8 and x8.shape == (3, 224, 224) and xf.dtype == torch.float32 and y8 == yf
    assert torch.allclose(input_norm.normalize_u8(x8[None])[0], xf, rtol=1e-5, atol=1e-5)
    t8, _ = tr8[0]
    assert t8.dtype == torch.uint8 and t8.shape == (3, 224, 224)
    loader = data.build_loader(va8, 2, 0, shuffle=False, pin=False)
    xb, yb = next(iter(loader))
    assert xb.dtype == torch.uint8 and xb.shape == (2, 3, 224, 224) and yb.dtype == torch.int64
