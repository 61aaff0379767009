"""T0: ShardedSampler is index-for-index identical to torch's DistributedSampler (torch/utils/data/distributed.py)."""
import itertools

import pytest
import torch
from torch.utils.data.distributed import DistributedSampler

from distributed_data_parallel_experiment_b200.data import ShardedSampler, SyntheticImages, prepare_dataloader


class _DS:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@pytest.mark.parametrize("n,ws,shuffle,drop_last,seed",
                         list(itertools.product([1, 7, 10, 100, 1001], [1, 2, 3, 8], [True, False], [True, False], [0, 5])))
def test_index_parity(n, ws, shuffle, drop_last, seed):
    ds = _DS(n)
    for rank in range(ws):
        a = ShardedSampler(ds, num_replicas=ws, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        b = DistributedSampler(ds, num_replicas=ws, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        for epoch in (0, 1, 3):
            a.set_epoch(epoch)
            b.set_epoch(epoch)
            assert list(a) == list(b)
            assert len(a) == len(b)


def test_cifar_sized_shards():
    ds = _DS(50000)
    s = ShardedSampler(ds, num_replicas=2, rank=1)
    assert len(s) == 25000
    # ceil(25000/512) == 49 steps: the reference's hard-coded steps_per_epoch (multigpu.py:137)
    assert -(-len(s) // 512) == 49


def test_bad_rank():
    with pytest.raises(ValueError):
        ShardedSampler(_DS(4), num_replicas=2, rank=2)


def test_prepare_dataloader_single():
    ds = SyntheticImages(64)
    dl = prepare_dataloader(ds, 16, distributed=False)
    x, y = next(iter(dl))
    assert x.shape == (16, 3, 32, 32) and y.dtype == torch.int64 and len(dl) == 4
