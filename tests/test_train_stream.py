"""Trainer.train_stream: the pipelined end-to-end loop yields the same losses, in order, as step-by-step train_step()."""
import copy

import pytest
import torch

import distributed_data_parallel_experiment_b200 as ddpx


def _toy(seed):
    torch.manual_seed(seed)
    m = ddpx.models.ToyLinear()
    opt = torch.optim.SGD(m.parameters(), lr=0.05)
    return m, opt


def test_train_stream_cpu_matches_train_step():
    torch.manual_seed(0)
    data = [(torch.randn(8, 20), torch.randn(8, 1)) for _ in range(5)]
    m1, o1 = _toy(1)
    m2, o2 = _toy(1)
    t1 = ddpx.Trainer(m1, [], o1, 0, 1, loss_fn="mse")
    t2 = ddpx.Trainer(m2, [], o2, 0, 1, loss_fn="mse")
    a = [t1.train_step(x, y) for x, y in data]
    b = list(t2.train_stream(iter(data)))
    assert a == b and len(b) == 5
    assert list(t2.train_stream(iter([]))) == []


@pytest.mark.gpu
def test_train_stream_gpu_matches_train_step():
    torch.manual_seed(0)
    data = [(torch.rand(64, 3, 32, 32).pin_memory(), torch.randint(0, 10, (64,)).pin_memory()) for _ in range(4)]
    torch.manual_seed(5)
    m1 = ddpx.models.VGG().cuda()
    m2 = copy.deepcopy(m1)
    o1 = torch.optim.SGD(m1.parameters(), lr=0.01, momentum=0.9)
    o2 = torch.optim.SGD(m2.parameters(), lr=0.01, momentum=0.9)
    t1 = ddpx.Trainer(m1, [], o1, 0, 1)
    t2 = ddpx.Trainer(m2, [], o2, 0, 1)
    a = [t1.train_step(x, y) for x, y in data]
    b = list(t2.train_stream(iter(data)))
    assert len(b) == 4 and all(v == v for v in b)
    # same kernels, same data: only fp32-atomic ordering differs between the two runs
    for u, v in zip(a, b):
        assert abs(u - v) < 2e-2 * max(1.0, abs(u)), (a, b)
