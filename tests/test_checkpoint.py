"""T1: checkpoint.pt interchanges with the reference's VGG in both directions."""
import torch

import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200 import checkpoint
from distributed_data_parallel_experiment_b200.models import VGG
from conftest import reference_module


def test_state_dict_layout_matches_reference():
    ref = reference_module("multigpu")
    ours, theirs = VGG().state_dict(), ref.VGG().state_dict()
    assert list(ours.keys()) == list(theirs.keys())
    assert len(ours) == 50
    for k in ours:
        assert ours[k].shape == theirs[k].shape and ours[k].dtype == theirs[k].dtype, k
    assert ours["backbone.bn3.num_batches_tracked"].dtype == torch.int64
    assert getattr(ours, "_metadata", None) is not None
    assert set(ours._metadata.keys()) == set(theirs._metadata.keys())


def test_same_seed_same_init_as_reference():
    ref = reference_module("multigpu")
    torch.manual_seed(7); a = VGG()
    torch.manual_seed(7); b = ref.VGG()
    for (ka, va), (kb, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)


def test_our_checkpoint_loads_into_reference(tmp_path):
    ref = reference_module("multigpu")
    m = VGG()
    with torch.no_grad():
        m(torch.randn(4, 3, 32, 32))          # touch BN buffers
    path = str(tmp_path / "checkpoint.pt")
    checkpoint.save_checkpoint(m, path)
    sd = torch.load(path, weights_only=True)
    r = ref.VGG()
    res = r.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    x = torch.randn(3, 3, 32, 32)
    m.eval(); r.eval()
    assert torch.allclose(m(x), r(x), atol=1e-5)


def test_reference_checkpoint_loads_into_ours(tmp_path):
    ref = reference_module("multigpu")
    r = ref.VGG()
    path = str(tmp_path / "checkpoint.pt")
    torch.save(r.state_dict(), path)               # what the reference's _save_checkpoint writes (multigpu.py:110-112)
    m = VGG()
    checkpoint.load_checkpoint(m, path, strict=True)
    x = torch.randn(2, 3, 32, 32)
    m.eval(); r.eval()
    assert torch.allclose(m(x), r(x), atol=1e-5)


def test_wrapped_model_saves_unprefixed_keys(tmp_path):
    class Wrap(torch.nn.Module):
        def __init__(self, m):
            super().__init__(); self.module = m
    m = VGG()
    path = str(tmp_path / "c.pt")
    checkpoint.save_checkpoint(Wrap(m), path)
    sd = torch.load(path, weights_only=True)
    assert all(not k.startswith("module.") for k in sd)


def test_snapshot_roundtrip(tmp_path):
    m = ddpx.models.ToyLinear()
    opt = torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 0.5 ** s)
    for _ in range(3):
        opt.zero_grad(); m(torch.rand(4, 20)).sum().backward(); opt.step(); sched.step()
    p = str(tmp_path / "snap.pt")
    checkpoint.save_snapshot(p, m, opt, sched, epoch=3)
    m2 = ddpx.models.ToyLinear()
    opt2 = torch.optim.SGD(m2.parameters(), lr=0.1, momentum=0.9)
    sched2 = torch.optim.lr_scheduler.LambdaLR(opt2, lambda s: 0.5 ** s)
    assert checkpoint.load_snapshot(p, m2, opt2, sched2) == 3
    assert sched2.last_epoch == 3
    for a, b in zip(m.parameters(), m2.parameters()):
        assert torch.equal(a, b)
    assert torch.equal(opt.state_dict()["state"][0]["momentum_buffer"], opt2.state_dict()["state"][0]["momentum_buffer"])
