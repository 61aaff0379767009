"""T2 / T6: distributed semantics on CPU with 2 gloo processes, against torch's own DistributedDataParallel as oracle
(BASELINE.json config #1 runs here: toy Linear(20,1) + MSE, world_size 2, gloo)."""
import os
import sys
import tempfile
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

import distributed_data_parallel_experiment_b200 as ddpx  # noqa: E402
from distributed_data_parallel_experiment_b200.parallel import DataParallel  # noqa: E402


def _setup(rank, world, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    ddpx.config.reset()
    ddpx.ddp_setup(rank, world, backend="gloo")


class SmallNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = ddpx.nn.Conv2d(3, 8, 3, padding=1, bias=False)
        self.bn = ddpx.nn.BatchNorm2d(8)
        self.fc1 = ddpx.nn.Linear(8, 16)
        self.fc2 = ddpx.nn.Linear(16, 4)

    def forward(self, x):
        x = torch.relu(self.bn(self.conv(x))).mean([2, 3])
        return self.fc2(torch.relu(self.fc1(x)))


def _run(fn, world=2, *args):
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(fn, args=(world, port, d) + args, nprocs=world, join=True)
        outs = [torch.load(os.path.join(d, f"rank{r}.pt"), weights_only=False) for r in range(world)
                if os.path.exists(os.path.join(d, f"rank{r}.pt"))]
    return outs


# ---------------------------------------------------------------------------------------------- workers
def _w_parity(rank, world, port, outdir):
    _setup(rank, world, port)
    torch.manual_seed(100 + rank)                     # different init per rank: construction must adopt rank 0's
    ours = SmallNet()
    theirs = SmallNet()
    theirs.load_state_dict(ours.state_dict())
    w_ours = DataParallel(ours, bucket_cap_mb=0.0004)       # tiny caps -> several buckets after the rebuild
    w_theirs = nn.parallel.DistributedDataParallel(theirs, bucket_cap_mb=0.0004)
    init_equal = all(torch.equal(a, b) for a, b in zip(ours.state_dict().values(), theirs.state_dict().values()))
    o1 = torch.optim.SGD(ours.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-3)
    o2 = torch.optim.SGD(theirs.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-3)
    g = torch.Generator().manual_seed(rank)
    max_grad_diff = 0.0
    for step in range(5):
        x = torch.randn(6, 3, 8, 8, generator=g); y = torch.randint(0, 4, (6,), generator=g)
        for w, o in ((w_ours, o1), (w_theirs, o2)):
            o.zero_grad()
            ddpx.ops.cross_entropy(w(x), y).backward()
        for a, b in zip(ours.parameters(), theirs.parameters()):
            max_grad_diff = max(max_grad_diff, (a.grad - b.grad).abs().max().item())
        o1.step(); o2.step()
    torch.save({"init_equal": init_equal, "max_grad_diff": max_grad_diff,
                "state": {k: v.clone() for k, v in ours.state_dict().items()},
                "ref_state": {k: v.clone() for k, v in theirs.state_dict().items()},
                "log": w_ours.get_logging_data()}, os.path.join(outdir, f"rank{rank}.pt"))
    ddpx.teardown()


def _w_vgg_buckets(rank, world, port, outdir):
    _setup(rank, world, port)
    torch.manual_seed(rank)
    m = ddpx.models.VGG()
    w = DataParallel(m)
    first = w.get_logging_data()["bucket_sizes"]
    for _ in range(2):
        w(torch.randn(2, 3, 32, 32)).sum().backward()
        m.zero_grad()
    w(torch.randn(2, 3, 32, 32))
    torch.save({"first": first, "log": w.get_logging_data()}, os.path.join(outdir, f"rank{rank}.pt"))
    ddpx.teardown()


def _w_no_sync(rank, world, port, outdir):
    _setup(rank, world, port)
    torch.manual_seed(0)
    a, b = SmallNet(), SmallNet()
    b.load_state_dict(a.state_dict())
    a.eval(); b.eval()            # BN on running stats: the loss is then additive over micro-batches
    wa, wb = DataParallel(a), DataParallel(b)
    g = torch.Generator().manual_seed(10 + rank)
    x = torch.randn(8, 3, 8, 8, generator=g); y = torch.randint(0, 4, (8,), generator=g)
    ddpx.ops.cross_entropy(wa(x), y).backward()                       # one big batch
    with wb.no_sync():
        (ddpx.ops.cross_entropy(wb(x[:4]), y[:4]) / 2).backward()
    local_only = [p.grad.clone() for p in b.parameters()]
    (ddpx.ops.cross_entropy(wb(x[4:]), y[4:]) / 2).backward()           # synchronised: reduces the accumulated sum
    diff = max((p.grad - q.grad).abs().max().item() for p, q in zip(a.parameters(), b.parameters()))
    torch.save({"diff": diff, "local": local_only}, os.path.join(outdir, f"rank{rank}.pt"))
    ddpx.teardown()


def _w_buffers(rank, world, port, outdir):
    _setup(rank, world, port)
    torch.manual_seed(rank)
    m = SmallNet()
    w = DataParallel(m)
    g = torch.Generator().manual_seed(50 + rank)
    w(torch.randn(4, 3, 8, 8, generator=g)).sum().backward()      # local BN stats now differ per rank
    after_step = m.bn.running_mean.clone()
    seen = {}
    def _peek(mod, inp):
        seen.setdefault("rm", m.bn.running_mean.clone())

    h = m.conv.register_forward_pre_hook(_peek)
    w(torch.randn(4, 3, 8, 8, generator=g))
    h.remove()
    torch.save({"after_step": after_step, "at_forward": seen["rm"], "nbt": m.bn.num_batches_tracked.clone()},
               os.path.join(outdir, f"rank{rank}.pt"))
    ddpx.teardown()


def _w_toy_trainer(rank, world, port, outdir):
    _setup(rank, world, port)
    sys.path.insert(0, str(REPO))
    torch.manual_seed(0)
    from datautils import MyTrainDataset
    ds = MyTrainDataset(256)
    torch.manual_seed(1 + rank)
    model = ddpx.models.ToyLinear()
    ref_model = nn.Linear(20, 1)
    opt = torch.optim.SGD(model.parameters(), lr=1e-2)
    loader = ddpx.prepare_dataloader(ds, 32)
    t = ddpx.Trainer(model, loader, opt, rank, 1, loss_fn="mse", checkpoint_path=os.path.join(outdir, "checkpoint.pt"))
    ref_model.load_state_dict({"weight": model.net.weight.detach().clone(), "bias": model.net.bias.detach().clone()})
    ref = nn.parallel.DistributedDataParallel(ref_model)
    ropt = torch.optim.SGD(ref_model.parameters(), lr=1e-2)
    t.train(2)
    for epoch in range(2):
        loader.sampler.set_epoch(epoch)
        for x, y in loader:
            ropt.zero_grad(); torch.nn.functional.mse_loss(ref(x), y).backward(); ropt.step()
    diff = max((model.net.weight - ref_model.weight).abs().max().item(), (model.net.bias - ref_model.bias).abs().max().item())
    torch.save({"diff": diff, "ckpt": os.path.exists(os.path.join(outdir, "checkpoint.pt")), "w": model.net.weight.detach().clone()},
               os.path.join(outdir, f"rank{rank}.pt"))
    ddpx.teardown()


def _w_mismatch(rank, world, port, outdir):
    _setup(rank, world, port)
    m = nn.Linear(4, 3 if rank == 0 else 5)
    try:
        DataParallel(m)
        ok = False
    except RuntimeError as e:
        ok = "does not match rank 0" in str(e)
    torch.save({"ok": ok}, os.path.join(outdir, f"rank{rank}.pt"))
    ddpx.teardown()


def _w_fail(rank, world, port, outdir):
    _setup(rank, world, port)
    if rank == 1:
        raise ValueError("injected failure on rank 1")
    import time
    time.sleep(60)


# ---------------------------------------------------------------------------------------------- tests
def test_matches_torch_ddp():
    outs = _run(_w_parity)
    assert len(outs) == 2
    for o in outs:
        assert o["init_equal"]
        assert o["max_grad_diff"] < 1e-6
        for k in o["state"]:
            assert torch.allclose(o["state"][k].float(), o["ref_state"][k].float(), atol=1e-6), k
        assert o["log"]["has_rebuilt_buckets"] and o["log"]["num_buckets"] > 1
    # replicas stay identical (parameters) across ranks
    for k, v in outs[0]["state"].items():
        if "running" not in k and "num_batches" not in k:
            assert torch.equal(v, outs[1]["state"][k]), k


def test_vgg_bucket_layout_after_rebuild():
    outs = _run(_w_vgg_buckets)
    for o in outs:
        assert o["first"] == [36913448]                                # first iteration: one bucket
        assert o["log"]["rebuilt_bucket_sizes"] == [9461800, 27148288, 303360]


def test_no_sync_accumulation():
    outs = _run(_w_no_sync)
    for o in outs:
        assert o["diff"] < 1e-6
    # inside no_sync nothing was exchanged: the two ranks' local gradients differ
    assert any((a - b).abs().max().item() > 1e-6 for a, b in zip(outs[0]["local"], outs[1]["local"]))


def test_buffers_follow_rank0_each_forward():
    outs = _run(_w_buffers)
    assert not torch.allclose(outs[0]["after_step"], outs[1]["after_step"])        # per-rank batch stats (no SyncBN)
    assert torch.equal(outs[1]["at_forward"], outs[0]["after_step"])               # rank 0's stats broadcast pre-forward
    assert int(outs[1]["nbt"]) == 2


def test_toy_trainer_matches_torch_ddp():
    outs = _run(_w_toy_trainer)
    for o in outs:
        assert o["diff"] < 1e-6 and o["ckpt"]
    assert torch.equal(outs[0]["w"], outs[1]["w"])


def test_param_mismatch_detected():
    outs = _run(_w_mismatch)
    assert all(o["ok"] for o in outs)


def test_spawn_fail_fast():
    import socket, time
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    t0 = time.time()
    with pytest.raises(Exception) as ei:
        ddpx.spawn(_w_fail, args=(2, port, "/tmp"), nprocs=2)
    assert "injected failure on rank 1" in str(ei.value)
    assert time.time() - t0 < 45            # the sleeping sibling was terminated, not waited for
