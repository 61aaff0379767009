"""SyncBatchNorm (optional extra): statistics over all ranks == BatchNorm over the concatenated batch."""
import os
import sys
import tempfile
from pathlib import Path

import pytest
import torch
import torch.multiprocessing as mp
from torch import nn

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

import distributed_data_parallel_experiment_b200 as ddpx  # noqa: E402


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = ddpx.nn.Conv2d(3, 6, 3, padding=1, bias=False)
        self.bn = ddpx.nn.BatchNorm2d(6)
        self.fc = ddpx.nn.Linear(6, 4)

    def forward(self, x):
        return self.fc(torch.relu(self.bn(self.conv(x))).mean([2, 3]))


def _port():
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _data(rank):
    g = torch.Generator().manual_seed(100 + rank)
    return torch.randn(5, 3, 8, 8, generator=g), torch.randint(0, 4, (5,), generator=g)


def _worker(rank, world, port, outdir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    ddpx.config.reset()
    ddpx.ddp_setup(rank, world, backend="gloo")
    torch.manual_seed(0)
    net = Net()
    ref = Net(); ref.load_state_dict(net.state_dict())
    ddpx.nn.convert_sync_batchnorm(net)
    assert isinstance(net.bn, ddpx.nn.SyncBatchNorm) and list(net.state_dict()) == list(ref.state_dict())
    w = ddpx.DataParallel(net)
    x, y = _data(rank)
    loss = ddpx.ops.cross_entropy(w(x), y)
    loss.backward()
    # oracle: ordinary BatchNorm over the concatenated global batch, in this process
    xs, ys = zip(*[_data(r) for r in range(world)])
    full = ddpx.ops.cross_entropy(ref(torch.cat(xs)), torch.cat(ys))
    full.backward()
    torch.save({"grad": {k: p.grad.clone() for k, p in net.named_parameters()},
                "ref_grad": {k: p.grad.clone() for k, p in ref.named_parameters()},
                "rm": net.bn.running_mean.clone(), "ref_rm": ref.bn.running_mean.clone(),
                "rv": net.bn.running_var.clone(), "ref_rv": ref.bn.running_var.clone(),
                "nbt": int(net.bn.num_batches_tracked)}, os.path.join(outdir, f"r{rank}.pt"))
    ddpx.teardown()


def test_syncbn_equals_full_batch_bn_on_gloo():
    world = 2
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, _port(), d), nprocs=world, join=True)
        outs = [torch.load(os.path.join(d, f"r{r}.pt"), weights_only=False) for r in range(world)]
    for o in outs:
        # DDP averages per-rank gradients of a loss that is a mean over the local batch: equal batch sizes => the
        # averaged gradient is the full-batch gradient
        for k, g in o["grad"].items():
            assert torch.allclose(g, o["ref_grad"][k], atol=2e-5, rtol=1e-4), k
        assert torch.allclose(o["rm"], o["ref_rm"], atol=1e-6) and torch.allclose(o["rv"], o["ref_rv"], atol=1e-5)
        assert o["nbt"] == 1
    assert torch.equal(outs[0]["rm"], outs[1]["rm"])          # identical statistics on every rank, without a broadcast


def test_convert_is_identity_without_process_group():
    net = Net()
    ddpx.nn.convert_sync_batchnorm(net)
    x, _ = _data(0)
    ref = Net(); ref.load_state_dict(net.state_dict())
    assert torch.allclose(net(x), ref(x), atol=1e-6)
