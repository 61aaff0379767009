"""T0: LR schedule values, model size, unit constants, CLI parsing, Trainer save cadence / argument forms."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200 import recipes
from distributed_data_parallel_experiment_b200.models import VGG, DeepNN, ToyLinear

REPO = Path(__file__).resolve().parent.parent


def test_lr_schedule_values():
    o = recipes.vgg_train_objs(world_size=2, synthetic=True, device=torch.device("cpu"), synthetic_size=0)
    lr0 = o.optimizer.param_groups[0]["lr"]
    assert float(lr0) == 0.0                      # LambdaLR ctor performs the initial step: first batch has lr 0
    lam = recipes.triangular_lr_lambda(49, 20)
    assert lam(0) == 0.0
    assert abs(lam(6 * 49) - 1.0) < 1e-12         # peak 0.4 at epoch 6
    assert lam(20 * 49) == 0.0 and lam(25 * 49) == 0.0
    assert abs(0.4 * lam(3 * 49) - 0.2) < 1e-12
    o1 = recipes.vgg_train_objs(world_size=1, synthetic=True, device=torch.device("cpu"), synthetic_size=0)
    for _ in range(98 * 6):
        o1.optimizer.step(); o1.scheduler.step()
    assert abs(o1.optimizer.param_groups[0]["lr"] - 0.4) < 1e-9


def test_model_size_and_units():
    m = VGG()
    assert sum(p.numel() for p in m.parameters()) == 9228362
    assert ddpx.get_model_size(m) == 9228362 * 32
    assert f"{ddpx.get_model_size(m) / ddpx.MiB:.2f}" == "35.20"
    assert ddpx.Byte == 8 and ddpx.KiB == 8192 and ddpx.GiB == 1024 * ddpx.MiB
    assert sum(p.numel() for p in DeepNN().parameters()) == 1186986
    assert len(list(m.buffers())) == 24


def test_vgg_shapes_cpu():
    m = VGG()
    y = m(torch.randn(2, 3, 32, 32))
    assert y.shape == (2, 10)
    assert DeepNN()(torch.randn(2, 3, 32, 32)).shape == (2, 10)


class _Loader(list):
    batch_size = 4


def _toy_loader(n=3):
    g = torch.Generator().manual_seed(0)
    return _Loader([(torch.rand(4, 20, generator=g), torch.rand(4, 1, generator=g)) for _ in range(n)])


def test_trainer_five_arg_form_and_save_cadence(tmp_path, monkeypatch, capsys):
    monkeypatch.chdir(tmp_path)
    model = ToyLinear()
    opt = torch.optim.SGD(model.parameters(), lr=1e-2)
    t = ddpx.Trainer(model, _toy_loader(), opt, 0, 2, loss_fn="mse", distributed=False)   # scheduler omitted
    saved = []
    orig = t._save_checkpoint
    t._save_checkpoint = lambda e: (saved.append(e), orig(e))
    t.train(5)
    assert saved == [0, 2, 4]                       # epoch % save_every == 0, including epoch 0
    assert (tmp_path / "checkpoint.pt").exists()
    out = capsys.readouterr().out
    assert "[GPU0] Epoch 0 | Batchsize: 4 | Steps: 3" in out
    assert "Epoch 4 | Training checkpoint saved at checkpoint.pt" in out


def test_trainer_save_every_zero_raises(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    model = ToyLinear()
    t = ddpx.Trainer(model, _toy_loader(), torch.optim.SGD(model.parameters(), lr=1e-2), 0, 0, loss_fn="mse", distributed=False)
    with pytest.raises(ZeroDivisionError):
        t.train(1)


def test_trainer_scheduler_stepped_per_batch():
    model = ToyLinear()
    opt = torch.optim.SGD(model.parameters(), lr=1.0)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0 / (1 + s))
    t = ddpx.Trainer(model, _toy_loader(3), opt, 0, 1, scheduler=sched, loss_fn="mse", distributed=False)
    t._run_epoch(0)
    assert sched.last_epoch == 3


def test_trainer_loss_decreases_toy():
    torch.manual_seed(0)
    model = ToyLinear()
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    t = ddpx.Trainer(model, _toy_loader(8), opt, 0, 100, loss_fn="mse", distributed=False)
    t._run_epoch(0)
    first = float(t.last_loss)
    for e in range(30):
        for x, y in t.train_data:
            t._run_batch(x, y)
    assert float(t.last_loss) < first


def test_gradient_accumulation_matches_big_batch():
    torch.manual_seed(0)
    a, b = ToyLinear(), ToyLinear()
    b.load_state_dict(a.state_dict())
    xs = torch.rand(8, 20); ys = torch.rand(8, 1)
    ta = ddpx.Trainer(a, [], torch.optim.SGD(a.parameters(), lr=0.1), 0, 1, loss_fn="mse", distributed=False)
    tb = ddpx.Trainer(b, [], torch.optim.SGD(b.parameters(), lr=0.1), 0, 1, loss_fn="mse", distributed=False, accum_steps=2)
    ta._run_batch(xs, ys)
    tb._run_batch(xs[:4], ys[:4]); tb._run_batch(xs[4:], ys[4:])
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-6)


@pytest.mark.parametrize("script", ["multigpu.py", "singlegpu.py"])
def test_cli_help_matches_reference_surface(script):
    r = subprocess.run([sys.executable, str(REPO / script), "--help"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0
    assert "total_epochs" in r.stdout and "save_every" in r.stdout and "--batch_size" in r.stdout
    assert "simple distributed training job" in r.stdout


def test_cli_requires_positionals():
    r = subprocess.run([sys.executable, str(REPO / "singlegpu.py")], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "total_epochs" in r.stderr


def test_datautils_toy_dataset():
    sys.path.insert(0, str(REPO))
    from datautils import MyTrainDataset
    ds = MyTrainDataset(2048)
    x, y = ds[5]
    assert len(ds) == 2048 and x.shape == (20,) and y.shape == (1,)


def test_algorithm_table_is_generated_from_sweep_rows(tmp_path):
    """bench/allreduce_sweep.emit_table: one-shot up to the largest size where it is not slower than two-shot (all smaller
    sizes agreeing), NVLS from the smallest size above which it beats two-shot by 3 % everywhere; other world sizes' entries
    are kept; and parallel/p2p.py reads the thresholds it wrote."""
    import importlib.util
    import json
    spec = importlib.util.spec_from_file_location("allreduce_sweep", REPO / "bench" / "allreduce_sweep.py")
    sweep = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sweep)
    rows = []
    for size, one, two, nv in [(1024, 10.0, 11.0, 11.5), (4096, 11.0, 11.1, 11.0), (16384, 14.0, 12.0, 11.9),
                               (1 << 20, 40.0, 30.0, 29.5), (1 << 24, 200.0, 100.0, 90.0), (1 << 26, None, 400.0, 350.0)]:
        if one is not None:
            rows.append({"impl": "p2p_oneshot+sgd", "bytes": size, "us": one})
        rows.append({"impl": "p2p_twoshot+sgd", "bytes": size, "us": two})
        rows.append({"impl": "p2p_nvls+sgd", "bytes": size, "us": nv})
    path = tmp_path / "table.json"
    path.write_text(json.dumps({"2": {"oneshot_max_bytes": 1 << 20, "nvls_min_bytes": None}}))
    entry = sweep.emit_table(rows, 8, str(path))
    assert entry == {"oneshot_max_bytes": 4096, "nvls_min_bytes": 1 << 24}      # 1 MB: NVLS only 1.7 % ahead -> not taken
    table = json.loads(path.read_text())
    assert table["2"]["oneshot_max_bytes"] == 1 << 20 and table["8"] == entry
    from distributed_data_parallel_experiment_b200.parallel import p2p
    shipped = json.loads((REPO / "distributed_data_parallel_experiment_b200" / "parallel" / "algo_table.json").read_text())
    assert p2p.oneshot_max_bytes(8) == shipped["8"]["oneshot_max_bytes"]
    assert p2p.nvls_min_bytes(8) == shipped["8"]["nvls_min_bytes"]


def test_select_mode_follows_the_table():
    """one-shot up to the table's threshold, the in-switch reduction only for multicast-bound buckets, two-shot otherwise;
    forced algorithms override the table."""
    from distributed_data_parallel_experiment_b200.parallel import p2p
    one = p2p.oneshot_max_bytes(8)
    assert p2p.select_mode("auto", one, 8) == 1 and p2p.select_mode("auto", one + 4, 8) == 0
    assert p2p.select_mode("auto", 1 << 27, 8, nvls=True) == 2 and p2p.select_mode("auto", 1 << 27, 8, nvls=False) == 0
    assert p2p.select_mode("twoshot", 64, 8) == 0 and p2p.select_mode("oneshot", 1 << 20, 8) == 1


def test_shards_partition_a_bucket():
    """FusedUpdate.shard_range restates the kernel's shard arithmetic: the ranks' ranges are disjoint, ordered, whole
    16-byte vectors, and cover the bucket (the last shards may be short or empty)."""
    from types import SimpleNamespace
    from distributed_data_parallel_experiment_b200.parallel.p2p import FusedUpdate
    for world in (2, 3, 8):
        fake = SimpleNamespace(comm=SimpleNamespace(world_size=world))
        for dtype, V in ((torch.float32, 4), (torch.bfloat16, 8)):
            for n in (V, 7 * V, 1856, 6787072, 8 * 1000 + V):
                got = [FusedUpdate.shard_range(fake, n, dtype, r) for r in range(world)]
                assert got[0][0] == 0 and got[-1][1] == n
                for (a0, a1), (b0, b1) in zip(got, got[1:]):
                    assert a1 == b0 and a0 <= a1 and a0 % V == 0
                assert all(b >= a for a, b in got)


def test_tail_bucket_split():
    """The reducer gives the last-ready parameters (at most DDPX_SPLIT_TAIL_BYTES) their own bucket on the fused backend
    and leaves torch's layout alone everywhere else."""
    from types import SimpleNamespace
    from distributed_data_parallel_experiment_b200.parallel.reducer import Reducer
    sizes = [64, 128 * 64 * 9, 128, 128, 64 * 27, 64, 64]          # elements, fp32: ... bn1, conv0 (1728), bn0 x2
    params = [torch.empty(s) for s in sizes]
    fused = SimpleNamespace(params=params, comm=SimpleNamespace(supports_fused_update=True))
    idx = [[0], [1, 2, 3, 4, 5, 6]]
    out = Reducer._split_tail(fused, idx)
    assert out[0] == [0] and out[1] + out[2] == idx[1] and len(out) == 3
    tail_bytes = sum(sizes[i] * 4 for i in out[2])
    assert 0 < tail_bytes <= 65536 and (sizes[out[1][-1]] * 4 + tail_bytes) > 65536        # as much as fits, no more
    plain = SimpleNamespace(params=params, comm=SimpleNamespace(supports_fused_update=False))
    assert Reducer._split_tail(plain, idx) == idx
