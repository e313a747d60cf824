"""Running meters and top-k accuracy.

Parity: /root/reference/utils.py:65-69 (get_learning_rate), :78-102
(AverageMeter), :105-111 (accuracy, returns a 0-D *fraction*).
``DeviceMeter`` is new: it keeps sum/count on the GPU so the hot loop has no
per-iteration ``.item()`` and is read only at print time.
"""
import torch


def get_learning_rate(optimizer):
    return [g["lr"] for g in optimizer.param_groups][0]


class AverageMeter(object):
    """val / avg / sum / count with the reference's ``__str__`` format."""

    def __init__(self, name, fmt=":f"):
        self.name = name
        self.fmt = fmt
        self.reset()

    def reset(self):
        self.val = 0
        self.avg = 0
        self.sum = 0
        self.count = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count

    def __str__(self):
        fmtstr = "{name} {val" + self.fmt + "} ({avg" + self.fmt + "})"
        return fmtstr.format(**self.__dict__)


class DeviceMeter(object):
    """AverageMeter whose accumulators live in one device tensor.

    ``update`` enqueues two tiny in-place ops and never synchronises; ``avg``
    and ``val`` are device tensors, ``read()`` does the single D2H copy.
    """

    def __init__(self, name, fmt=":f", device=None):
        self.name = name
        self.fmt = fmt
        self.device = device
        self._acc = None  # [sum, count, last]
        self.reset()

    def reset(self):
        self._acc = torch.zeros(3, dtype=torch.float32, device=self.device)

    def update(self, val, n=1):
        v = val.detach().to(torch.float32).reshape(()) if torch.is_tensor(val) else torch.tensor(float(val))
        self._acc[0].add_(v.to(self._acc.device) * n)
        self._acc[1].add_(float(n))
        self._acc[2].copy_(v)

    @property
    def avg(self):
        return self._acc[0] / torch.clamp(self._acc[1], min=1.0)

    @property
    def val(self):
        return self._acc[2]

    def read(self):
        s, c, v = self._acc.tolist()
        return {"sum": s, "count": c, "val": v, "avg": s / c if c else 0.0}

    def __str__(self):
        r = self.read()
        fmtstr = "{name} {val" + self.fmt + "} ({avg" + self.fmt + "})"
        return fmtstr.format(name=self.name, val=r["val"], avg=r["avg"])


def accuracy(scores, targets, k=1):
    """Top-k hit fraction in [0, 1] as a 0-D tensor on ``scores.device``."""
    batch_size = targets.size(0)
    _, ind = scores.topk(k, 1, True, True)
    hit = ind.eq(targets.long().view(-1, 1).expand_as(ind))
    return hit.reshape(-1).float().sum() * (1.0 / batch_size)


def reduce_mean(tensor, nprocs):
    """Mean of ``tensor`` over the ranks: clone -> all_reduce(SUM) -> / nprocs, the reference's helper
    (/root/reference/distributed.py:78-82).  The trainer itself packs loss and accuracy into one 8-byte
    device all-reduce (``engine.StepRunner.update_metrics``); this is the same operation for user code."""
    import torch.distributed as dist
    rt = tensor.clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(rt, op=dist.ReduceOp.SUM)
    rt /= nprocs
    return rt
