"""Reference arm: run the UNMODIFIED reference scripts installed in ``baseline/_ref``.

The reference (xiezheng-cs/PyTorch_Distributed_Template) does not launch as-is on
torch >= 2.0.  This shim changes nothing inside its files; it
prepares the process around them and then calls the script's own ``main()``:

  Q1  ``--local_rank=$LOCAL_RANK`` is put on argv (torchrun only sets the env var);
  Q2  a stub ``tensorboardX`` module is registered (not installed on this image);
  Q3  ``--outpath`` points at a fresh temp dir, so ``output_process`` never prompts;
  Q4  the fp32-DDP cell runs ``distributed_syncBN_amp.py --use_amp ''`` (empty string ->
      False): identical loop to ``distributed.py`` minus its debug ``break``;
  Q6  the caller sets CUDA_VISIBLE_DEVICES to the first WORLD_SIZE GPUs before CUDA
      initialises, because the script sizes the job with ``torch.cuda.device_count()``;
  data: ``datasets.ImageFolder`` / ``DataLoader`` *as seen by the script* are replaced by a
      synthetic source that yields pinned fp32 NCHW batches (BASELINE.json: synthetic data),
      and whose iterator brackets the timed steps with barrier + synchronize + CUDA events.

Everything between two ``next(loader)`` calls — H2D copy, forward, loss, barrier, the two
scalar all-reduces, ``.item()``, backward with DDP's bucketed NCCL all-reduce, optimizer /
GradScaler step — is the reference's own code on the stock torch/cuDNN/NCCL path.
"""
import importlib
import os
import sys
import tempfile
import time
import types

REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")

MODES = {
    # mode -> (module, extra argv)
    "dp": ("dataparallel", []),
    "ddp": ("distributed_syncBN_amp", ["--use_amp", ""]),
    "ddp_amp": ("distributed_syncBN_amp", []),
    "ddp_amp_syncbn": ("distributed_syncBN_amp", ["--sync_batchnorm", "1"]),
}


def available():
    return os.path.exists(os.path.join(REF_DIR, "distributed_syncBN_amp.py"))


class _NullWriter:
    def __init__(self, *a, **k):
        pass

    def add_scalar(self, *a, **k):
        pass

    def close(self):
        pass


class TimedSyntheticLoader:
    """Stands in for ``DataLoader(train_dataset, ...)`` inside the reference script."""

    def __init__(self, batch, warmup, steps, image_size=224, num_classes=1000, distinct=3):
        import torch
        self.torch = torch
        self.batch, self.warmup, self.steps = batch, warmup, steps
        g = torch.Generator().manual_seed(1234)
        self.images = [torch.randn(batch, 3, image_size, image_size, generator=g).pin_memory() for _ in range(distinct)]
        self.targets = [torch.randint(0, num_classes, (batch,), generator=g).pin_memory() for _ in range(distinct)]
        self.result = None

    def __len__(self):
        return self.warmup + self.steps

    def _fence(self):
        torch = self.torch
        import torch.distributed as dist
        torch.cuda.synchronize()
        if dist.is_available() and dist.is_initialized():
            dist.barrier()
            torch.cuda.synchronize()

    def __iter__(self):
        torch = self.torch
        start = stop = None
        t0 = 0.0
        for i in range(self.warmup + self.steps):
            if i == self.warmup:
                self._fence()
                start = torch.cuda.Event(enable_timing=True)
                stop = torch.cuda.Event(enable_timing=True)
                if self.on_timed_region is not None:
                    self.on_timed_region(True)
                t0 = time.perf_counter()
                start.record()
            yield self.images[i % len(self.images)], self.targets[i % len(self.targets)]
        stop.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        if self.on_timed_region is not None:
            self.on_timed_region(False)
        self._fence()
        self.result = {"device_ms": start.elapsed_time(stop), "wall_ms": wall * 1e3}

    on_timed_region = None


class _TinyLoader:
    def __init__(self, batch, image_size=224, num_classes=1000):
        import torch
        self.item = (torch.randn(batch, 3, image_size, image_size), torch.randint(0, num_classes, (batch,)))

    def __len__(self):
        return 1

    def __iter__(self):
        yield self.item


class _FakeFolder:
    """``datasets.ImageFolder(dir, transform)`` replacement: only its length is ever used
    (by DistributedSampler), the batches come from the loaders above."""

    def __init__(self, root, transform=None, *a, **k):
        self.root = root
        self.is_train = os.path.basename(os.path.normpath(root)) == "train"

    def __len__(self):
        return 1 << 20

    def __getitem__(self, i):
        raise RuntimeError("synthetic batches are produced by the loader")


def run_reference(mode, per_process_batch_total, steps, warmup, on_timed_region=None, extra_argv=None,
                  model_factory=None, print_freq=1000000):
    """Run one reference script for ``warmup + steps`` training iterations.

    ``per_process_batch_total`` is the value for ``-b`` (TOTAL batch, the script divides it).
    Returns the loader's timing dict (+ ``outpath``, the directory the script logged into).
    ``model_factory`` (parity tests): what ``models.__dict__[arch]()`` returns inside the script, so both arms
    start from the same weights; ``extra_argv`` appends flags of the script's own CLI (e.g. ``--lr``).
    """
    module_name, extra = MODES[mode]
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    if "tensorboardX" not in sys.modules:
        try:
            importlib.import_module("tensorboardX")
        except Exception:
            stub = types.ModuleType("tensorboardX")
            stub.SummaryWriter = _NullWriter
            sys.modules["tensorboardX"] = stub
    import torch  # noqa: F401
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    ref = importlib.import_module(module_name)
    assert os.path.dirname(os.path.abspath(ref.__file__)) == REF_DIR, ref.__file__

    ndev = torch.cuda.device_count()
    per_loader_batch = per_process_batch_total // (ndev if mode != "dp" else 1)
    holder = {}

    def fake_loader(dataset, batch_size=1, *a, **k):
        if getattr(dataset, "is_train", False):
            ld = TimedSyntheticLoader(batch_size, warmup, steps)
            ld.on_timed_region = on_timed_region
            holder["train"] = ld
            return ld
        return _TinyLoader(min(batch_size, 8))

    # names as the script itself sees them; torchvision / torch stay untouched
    ref.datasets = types.SimpleNamespace(ImageFolder=_FakeFolder)
    ref.DataLoader = fake_loader
    if model_factory is not None:
        ref.models = types.SimpleNamespace(resnet18=model_factory)
    out = tempfile.mkdtemp(prefix="pdt_ref_{}_".format(mode))
    argv = [module_name + ".py", "--epochs", "1", "-b", str(per_process_batch_total), "-j", "0",
            "--outpath", os.path.join(out, "run"), "--data", "/nonexistent", "-p", str(print_freq)] + extra \
        + list(extra_argv or [])
    if mode != "dp":
        argv.append("--local_rank={}".format(local_rank))
    old_argv = sys.argv
    sys.argv = argv
    try:
        ref.main()
    finally:
        sys.argv = old_argv
    res = dict(holder["train"].result)
    res["per_loader_batch"] = per_loader_batch
    res["world"] = world
    res["outpath"] = os.path.join(out, "run") + "_resnet18"
    return res
