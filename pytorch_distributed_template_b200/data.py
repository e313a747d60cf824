"""Input pipeline (SURVEY layer L7).

* ``imagefolder_datasets`` — the reference's ImageNet transforms
  (/root/reference/distributed.py:155-176).
* ``SyntheticImageNet`` — deterministic ImageNet-shaped tensors generated from
  the sample index; used by tests, the benchmark and ``PDT_SYNTHETIC=1``.
* ``ShardSampler`` — rank-strided shard of a per-epoch permutation with the
  exact index stream of torch's DistributedSampler
  (torch/utils/data/distributed.py:107-141), used for train AND val like
  distributed.py:167,177.
* ``PinnedPrefetcher`` — double-buffered pinned-host → device copies on a side
  stream, so H2D of batch i+1 overlaps compute of batch i (the reference
  cannot overlap because it drains the stream twice per step).
"""
import math
import os

import torch
from torch.utils.data import DataLoader, Dataset, Sampler

from .env import env_int


class SyntheticImageNet(Dataset):
    def __init__(self, length, image_size=224, num_classes=1000, seed=0, cache=64, uint8=False):
        self.length = int(length)
        self.image_size = int(image_size)
        self.num_classes = int(num_classes)
        self.seed = int(seed)
        # A small bank of distinct images; sample i is bank[i % cache] so a long
        # synthetic epoch does not spend its time in the host RNG.
        g = torch.Generator().manual_seed(self.seed)
        n = max(1, min(cache, self.length))
        if uint8:      # raw pixels, normalised on the device (ops/input_norm.py)
            self._bank = torch.randint(0, 256, (n, 3, self.image_size, self.image_size), generator=g, dtype=torch.uint8)
        else:
            self._bank = torch.randn(n, 3, self.image_size, self.image_size, generator=g)
        self.classes = list(range(self.num_classes))

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        if i < 0 or i >= self.length:
            raise IndexError(i)
        return self._bank[i % self._bank.size(0)], (i * 2654435761) % self.num_classes


def imagefolder_datasets(root, uint8=False):
    """The reference's two ImageFolder pipelines (distributed.py:112-131).  ``uint8``: stop at the cropped
    uint8 CHW image (``PILToTensor``); ToTensor's /255 and Normalize then run on the device with the same
    constants (ops/input_norm.py)."""
    import torchvision.datasets as datasets
    import torchvision.transforms as transforms
    from .ops.input_norm import IMAGENET_MEAN, IMAGENET_STD
    if uint8:
        tail = [transforms.PILToTensor()]
    else:
        tail = [transforms.ToTensor(), transforms.Normalize(mean=list(IMAGENET_MEAN), std=list(IMAGENET_STD))]
    train = datasets.ImageFolder(os.path.join(root, "train"), transforms.Compose([
        transforms.RandomResizedCrop(224), transforms.RandomHorizontalFlip()] + tail))
    val = datasets.ImageFolder(os.path.join(root, "val"), transforms.Compose([
        transforms.Resize(256), transforms.CenterCrop(224)] + tail))
    return train, val


def build_datasets(args, uint8=False):
    """ImageFolder under ``args.data`` or, with ``PDT_SYNTHETIC=1``, synthetic sets."""
    from .env import env_flag
    if env_flag("PDT_SYNTHETIC"):
        size = env_int("PDT_IMAGE_SIZE", 224)
        ncls = env_int("PDT_NUM_CLASSES", 1000)
        return (SyntheticImageNet(env_int("PDT_SYNTH_TRAIN_SIZE", 1281167), size, ncls, seed=1, uint8=uint8),
                SyntheticImageNet(env_int("PDT_SYNTH_VAL_SIZE", 50000), size, ncls, seed=2, uint8=uint8))
    return imagefolder_datasets(args.data, uint8=uint8)


class ShardSampler(Sampler):
    def __init__(self, dataset, num_replicas=None, rank=None, shuffle=True, seed=0, drop_last=False):
        import torch.distributed as dist
        if num_replicas is None:
            num_replicas = dist.get_world_size() if dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank() if dist.is_initialized() else 0
        if not 0 <= rank < num_replicas:
            raise ValueError("rank {} outside [0, {})".format(rank, num_replicas))
        self.n = len(dataset)
        self.num_replicas, self.rank = num_replicas, rank
        self.shuffle, self.seed, self.drop_last = shuffle, seed, drop_last
        self.epoch = 0
        if drop_last and self.n % num_replicas:
            self.num_samples = math.ceil((self.n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(self.n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch):
        self.epoch = int(epoch)

    def __len__(self):
        return self.num_samples

    def __iter__(self):
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            order = torch.randperm(self.n, generator=g).tolist()
        else:
            order = list(range(self.n))
        if self.drop_last:
            order = order[:self.total_size]
        else:
            pad = self.total_size - len(order)
            if pad > 0:  # wrap around, repeating the head as often as needed
                order += (order * math.ceil(pad / max(1, len(order))))[:pad]
        return iter(order[self.rank:self.total_size:self.num_replicas])


def build_loader(dataset, batch_size, workers, sampler=None, shuffle=False, pin=True):
    return DataLoader(dataset, batch_size=batch_size, shuffle=(shuffle and sampler is None),
                      num_workers=workers, pin_memory=pin and torch.cuda.is_available(), sampler=sampler,
                      persistent_workers=False)


class PinnedPrefetcher:
    """Iterates ``(images, target)`` already on ``device``.

    The next batch's H2D copy is issued on a side stream from pinned memory
    while the current batch computes; ``channels_last``/dtype conversion is
    left to the model's first kernel.
    """

    def __init__(self, loader, device):
        self.loader, self.device = loader, device
        self.stream = torch.cuda.Stream(device) if device.type == "cuda" else None

    def __len__(self):
        return len(self.loader)

    def _stage(self, batch):
        images, target = batch
        if self.stream is None:
            return images.to(self.device), torch.as_tensor(target).to(self.device)
        with torch.cuda.stream(self.stream):
            images = images.to(self.device, non_blocking=True)
            target = torch.as_tensor(target).to(self.device, non_blocking=True)
        return images, target

    def __iter__(self):
        it = iter(self.loader)
        try:
            nxt = self._stage(next(it))
        except StopIteration:
            return
        while nxt is not None:
            if self.stream is not None:
                torch.cuda.current_stream(self.device).wait_stream(self.stream)
                for t in nxt:
                    t.record_stream(torch.cuda.current_stream(self.device))
            cur = nxt
            try:
                nxt = self._stage(next(it))
            except StopIteration:
                nxt = None
            yield cur
