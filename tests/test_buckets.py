"""T0: bucket assignment parity with torch's c10d rule, including the VGG 3-bucket layout."""
import random

import pytest
import torch
import torch.distributed as dist

from distributed_data_parallel_experiment_b200.models import VGG
from distributed_data_parallel_experiment_b200.parallel import compute_bucket_assignment_by_size


def _rand_tensors(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        numel = rng.choice([1, 10, 512, 4096, 70000, 300000, 2500000])
        dt = rng.choice([torch.float32, torch.float32, torch.float16])
        out.append(torch.empty(numel, dtype=dt))
    return out


@pytest.mark.parametrize("seed", range(8))
@pytest.mark.parametrize("limits", [[1 << 62], [1024 * 1024, 25 * 1024 * 1024], [4096, 65536, 1 << 20]])
def test_parity_default_order(seed, limits):
    ts = _rand_tensors(40, seed)
    if len({t.dtype for t in ts}) > 1 and len(limits) > 0:
        # leftover order across dtypes follows an unordered_map in torch; compare as sets of buckets
        ours, lim_o = compute_bucket_assignment_by_size(ts, limits)
        ref, lim_r = dist._compute_bucket_assignment_by_size(ts, limits)
        assert sorted(map(tuple, ours)) == sorted(map(tuple, ref))
    else:
        ours, lim_o = compute_bucket_assignment_by_size(ts, limits)
        ref, lim_r = dist._compute_bucket_assignment_by_size(ts, limits)
        assert ours == ref and lim_o == lim_r


def test_vgg_rebuilt_buckets():
    m = VGG()
    params = list(m.parameters())
    order = list(reversed(range(len(params))))      # gradient-ready order ~ reverse definition order
    tensors = [params[i] for i in order]
    ours, _ = compute_bucket_assignment_by_size(tensors, [1024 * 1024, 25 * 1024 * 1024], tensor_indices=order)
    ref, _ = dist._compute_bucket_assignment_by_size(tensors, [1024 * 1024, 25 * 1024 * 1024], [False] * len(order), order)
    assert ours == ref
    sizes = [sum(params[i].numel() * 4 for i in b) for b in ours]
    assert sizes == [9461800, 27148288, 303360]


def test_vgg_first_iteration_single_bucket():
    m = VGG()
    params = list(m.parameters())
    ours, _ = compute_bucket_assignment_by_size(params, [1 << 62])
    assert len(ours) == 1 and sum(p.numel() * 4 for p in params) == 36913448
