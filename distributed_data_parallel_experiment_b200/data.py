"""Data layer: sharded sampler, synthetic datasets, CIFAR-10 access and loaders.

* ``ShardedSampler`` reproduces ``torch.utils.data.DistributedSampler`` index-for-index
  (torch/utils/data/distributed.py:107-141) - the reference builds one with defaults at multigpu.py:153 and calls
  ``set_epoch`` at multigpu.py:103.
* ``prepare_dataloader`` mirrors multigpu.py:147-154 / singlegpu.py:174-180.
* ``SyntheticImages`` / ``SyntheticTokens`` give shape-exact random data for throughput runs (no network here).
* ``BatchedImageLoader`` is the fast path for in-memory image sets: whole-batch random-crop + flip on uint8 tensors
  into pinned buffers instead of 512 PIL calls per step on the training thread (the reference's num_workers=0
  pipeline).
"""
from __future__ import annotations

import math
from typing import Iterator, Optional

import torch
from torch.utils.data import DataLoader, Dataset, Sampler


def _dist_info():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(), dist.get_rank()
    return 1, 0


class ShardedSampler(Sampler[int]):
    """Rank-sharded sampler with per-epoch reshuffle; drop-in for DistributedSampler."""

    def __init__(self, dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        if num_replicas is None or rank is None:
            ws, rk = _dist_info()
            num_replicas = ws if num_replicas is None else num_replicas
            rank = rk if rank is None else rank
        if rank >= num_replicas or rank < 0:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset = dataset
        self.num_replicas = num_replicas
        self.rank = rank
        self.epoch = 0
        self.drop_last = drop_last
        n = len(self.dataset)
        if self.drop_last and n % self.num_replicas != 0:
            self.num_samples = math.ceil((n - self.num_replicas) / self.num_replicas)
        else:
            self.num_samples = math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas
        self.shuffle = shuffle
        self.seed = seed

    def __iter__(self) -> Iterator[int]:
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            indices = torch.randperm(n, generator=g).tolist()
        else:
            indices = list(range(n))
        if not self.drop_last:
            padding_size = self.total_size - len(indices)
            if padding_size <= len(indices):
                indices += indices[:padding_size]
            else:
                indices += (indices * math.ceil(padding_size / len(indices)))[:padding_size]
        else:
            indices = indices[: self.total_size]
        assert len(indices) == self.total_size
        indices = indices[self.rank: self.total_size: self.num_replicas]
        assert len(indices) == self.num_samples
        return iter(indices)

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch


def prepare_dataloader(dataset: Dataset, batch_size: int, distributed: Optional[bool] = None, **kw) -> DataLoader:
    """DataLoader as the reference builds it: sharded sampler when a process group exists, shuffle otherwise."""
    ws, _ = _dist_info()
    if distributed is None:
        distributed = ws > 1 or _pg_initialized()
    pin = kw.pop("pin_memory", torch.cuda.is_available())
    if distributed:
        return DataLoader(dataset, batch_size=batch_size, pin_memory=pin, shuffle=False,
                          sampler=ShardedSampler(dataset), **kw)
    return DataLoader(dataset, batch_size=batch_size, pin_memory=pin, shuffle=True, **kw)


def _pg_initialized() -> bool:
    import torch.distributed as dist
    return dist.is_available() and dist.is_initialized()


class SyntheticImages(Dataset):
    """CIFAR-shaped random images + labels, generated once (deterministic in ``seed``)."""

    def __init__(self, size: int = 50000, shape=(3, 32, 32), num_classes: int = 10, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        self.images = torch.rand(size, *shape, generator=g)
        self.labels = torch.randint(0, num_classes, (size,), generator=g)

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        return self.images[i], self.labels[i]


class SyntheticTokens(Dataset):
    """Random token sequences for language-model throughput runs: (input[T], target[T])."""

    def __init__(self, size: int, seq_len: int, vocab: int, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        self.tokens = torch.randint(0, vocab, (size, seq_len + 1), generator=g)

    def __len__(self):
        return self.tokens.shape[0]

    def __getitem__(self, i):
        t = self.tokens[i]
        return t[:-1], t[1:]


class SyntheticVectors(Dataset):
    """Random feature vectors + class labels (MLP config of BASELINE.json)."""

    def __init__(self, size: int, dim: int, num_classes: int, seed: int = 0, dtype=torch.float32):
        g = torch.Generator().manual_seed(seed)
        self.x = torch.randn(size, dim, generator=g).to(dtype)
        self.y = torch.randint(0, num_classes, (size,), generator=g)

    def __len__(self):
        return self.x.shape[0]

    def __getitem__(self, i):
        return self.x[i], self.y[i]


def get_training_data(root: str = "data/cifar10", download: bool = True):
    """CIFAR-10 train/test sets with the reference's transforms (multigpu.py:157-175).

    Differences that fix races the reference ignores: only rank 0 downloads, the other ranks
    wait on a barrier.  Raises if the data is absent and cannot be downloaded (this image has no network).
    """
    from torchvision.datasets import CIFAR10
    from torchvision.transforms import Compose, RandomCrop, RandomHorizontalFlip, ToTensor
    import torch.distributed as dist
    transforms = {"train": Compose([RandomCrop(32, padding=4), RandomHorizontalFlip(), ToTensor()]), "test": ToTensor()}
    ws, rank = _dist_info()
    dataset = {}
    if ws > 1:
        if rank == 0:
            for split in ["train", "test"]:
                CIFAR10(root=root, train=(split == "train"), download=download)
        dist.barrier()
        download = False
    for split in ["train", "test"]:
        dataset[split] = CIFAR10(root=root, train=(split == "train"), download=download, transform=transforms[split])
    return dataset["train"], dataset["test"]


class BatchedImageLoader:
    """Whole-batch augmentation loader over an in-memory uint8 image tensor [N,H,W,C] (or float [N,C,H,W]).

    Iterates ``(images[B,C,H,W] float32 in [0,1], labels[B] int64)`` from pinned staging buffers; random crop with
    zero padding and horizontal flip are applied with batched tensor ops.  ``sampler`` may be a ShardedSampler.
    """

    def __init__(self, images: torch.Tensor, labels: torch.Tensor, batch_size: int, sampler=None, shuffle: bool = True,
                 augment: bool = True, pad: int = 4, seed: int = 0, pin_memory: Optional[bool] = None):
        self.images = images
        self.labels = labels.to(torch.int64)
        self.batch_size = batch_size
        self.sampler = sampler
        self.shuffle = shuffle
        self.augment = augment
        self.pad = pad
        self.gen = torch.Generator().manual_seed(seed)
        self.pin = torch.cuda.is_available() if pin_memory is None else pin_memory
        self.dataset = self

    def __len__(self):
        n = len(self.sampler) if self.sampler is not None else self.images.shape[0]
        return math.ceil(n / self.batch_size)

    def _indices(self):
        if self.sampler is not None:
            return torch.tensor(list(iter(self.sampler)), dtype=torch.int64)
        n = self.images.shape[0]
        return torch.randperm(n, generator=self.gen) if self.shuffle else torch.arange(n)

    def __iter__(self):
        idx = self._indices()
        for s in range(0, idx.numel(), self.batch_size):
            b = idx[s: s + self.batch_size]
            x = self.images[b]
            if x.dtype == torch.uint8:  # NHWC uint8 -> NCHW float
                x = x.permute(0, 3, 1, 2).to(torch.float32).div_(255.0)
            else:
                x = x.clone()
            if self.augment:
                B, C, H, W = x.shape
                p = self.pad
                xp = torch.zeros(B, C, H + 2 * p, W + 2 * p, dtype=x.dtype)
                xp[:, :, p:p + H, p:p + W] = x
                oy = torch.randint(0, 2 * p + 1, (B,), generator=self.gen)
                ox = torch.randint(0, 2 * p + 1, (B,), generator=self.gen)
                rows = (oy[:, None] + torch.arange(H)[None, :])
                cols = (ox[:, None] + torch.arange(W)[None, :])
                flip = torch.rand(B, generator=self.gen) < 0.5
                cols = torch.where(flip[:, None], cols.flip(1), cols)
                x = xp[torch.arange(B)[:, None, None, None], torch.arange(C)[None, :, None, None],
                       rows[:, None, :, None], cols[:, None, None, :]]
            y = self.labels[b]
            if self.pin:
                x = x.pin_memory(); y = y.pin_memory()
            yield x, y
