"""T0: configuration knobs, backend resolution, metrics helpers and the batched image loader (CPU only)."""
import os

import torch

import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200 import config, launch, metrics
from distributed_data_parallel_experiment_b200.data import BatchedImageLoader, ShardedSampler, SyntheticTokens, SyntheticVectors


def test_config_defaults_reproduce_reference_literals(monkeypatch):
    for k in list(os.environ):
        if k.startswith("DDPX_"):
            monkeypatch.delenv(k)
    config.reset()
    c = config.get()
    assert c.master_addr == "localhost" and c.master_port == 12355          # multigpu.py:30-31
    assert c.checkpoint_path == "checkpoint.pt" and c.bucket_cap_mb == 25.0  # multigpu.py:111, torch DDP default
    assert c.backend == "auto" and c.allreduce_algo == "auto"


def test_config_env_overrides(monkeypatch):
    monkeypatch.setenv("DDPX_BACKEND", "gloo")
    monkeypatch.setenv("DDPX_BUCKET_CAP_MB", "4")
    monkeypatch.setenv("DDPX_COMM_CTAS", "16")
    monkeypatch.setenv("DDPX_MASTER_PORT", "23456")
    config.reset()
    c = config.get()
    assert c.backend == "gloo" and c.bucket_cap_mb == 4.0 and c.comm_ctas == 16 and c.master_port == 23456
    assert launch.resolve_backend(None) == "gloo"
    config.reset()


def test_backend_auto_on_cpu():
    config.reset()
    if not torch.cuda.is_available():
        assert launch.resolve_backend(None) == "gloo"
    assert launch.resolve_backend("nccl") == "nccl"
    assert launch.world_size() == 1 and launch.rank() == 0 and not launch.is_distributed()


def test_step_timer_and_jsonl(tmp_path):
    t = metrics.StepTimer(torch.device("cpu"))
    for _ in range(3):
        t.start(); sum(range(1000)); t.stop()
    ms = t.times_ms()
    assert len(ms) == 3 and all(m >= 0 for m in ms)
    p = tmp_path / "m.jsonl"
    lg = metrics.JsonlLogger(str(p))
    lg.log(rank=0, step=1, device_ms=1.5); lg.log(rank=0, step=2, device_ms=1.4); lg.close()
    import json
    rows = [json.loads(l) for l in p.read_text().splitlines()]
    assert [r["step"] for r in rows] == [1, 2] and "ts" in rows[0]


def test_batched_image_loader_shapes_and_sharding():
    imgs = torch.randint(0, 256, (100, 32, 32, 3), dtype=torch.uint8)
    labels = torch.arange(100) % 10
    class DS:
        def __len__(self): return 100
    samp = ShardedSampler(DS(), num_replicas=2, rank=1, shuffle=True, seed=0)
    dl = BatchedImageLoader(imgs, labels, 16, sampler=samp, augment=True, pin_memory=False)
    assert len(dl) == 4                                     # ceil(50 / 16)
    seen = 0
    for x, y in dl:
        assert x.dtype == torch.float32 and x.shape[1:] == (3, 32, 32) and 0.0 <= float(x.min()) and float(x.max()) <= 1.0
        assert y.dtype == torch.int64
        seen += x.shape[0]
    assert seen == 50
    plain = BatchedImageLoader(imgs, labels, 25, shuffle=False, augment=False, pin_memory=False)
    x0, y0 = next(iter(plain))
    assert torch.allclose(x0, imgs[:25].permute(0, 3, 1, 2).float() / 255.0) and torch.equal(y0, labels[:25])


def test_synthetic_datasets():
    t = SyntheticTokens(8, 16, 100)
    x, y = t[3]
    assert x.shape == (16,) and torch.equal(x[1:], y[:-1])
    v = SyntheticVectors(4, 32, 10, dtype=torch.bfloat16)
    a, b = v[0]
    assert a.dtype == torch.bfloat16 and a.shape == (32,) and 0 <= int(b) < 10


def test_gpt2_and_mlp_cpu_forward_backward():
    cfg = ddpx.models.GPT2Config.tiny()
    m = ddpx.models.GPT2(cfg, dtype=torch.float32)
    idx = torch.randint(0, cfg.vocab_size, (2, 32))
    loss = m(idx, idx)
    loss.backward()
    assert torch.isfinite(loss) and m.wte.weight.grad is not None
    assert ddpx.models.GPT2(ddpx.models.GPT2Config.small(), dtype=None).num_params() == 124475904
    mlp = ddpx.models.MLP(64, 64, 3, 16, dtype=torch.float32)
    assert mlp(torch.randn(4, 64)).shape == (4, 16)


def test_trainer_writes_step_metrics_jsonl(tmp_path, monkeypatch):
    import json
    path = tmp_path / "m_{rank}.jsonl"
    monkeypatch.setenv("DDPX_METRICS_JSONL", str(path))
    config.reset()
    torch.manual_seed(0)
    m = ddpx.models.ToyLinear()
    opt = torch.optim.SGD(m.parameters(), lr=0.01)
    data = [(torch.randn(8, 20), torch.randn(8, 1)) for _ in range(4)]
    tr = ddpx.Trainer(m, data, opt, 0, 1, loss_fn="mse", accum_steps=2, checkpoint_path=str(tmp_path / "c.pt"))
    tr.train(1)
    rows = [json.loads(l) for l in (tmp_path / "m_0.jsonl").read_text().splitlines()]
    assert len(rows) == 2                                   # 4 micro-batches, 2 per optimizer step
    assert all(r["rank"] == 0 and r["samples"] == 16 and r["device_ms"] >= 0 and r["world_size"] == 1 for r in rows)
    monkeypatch.delenv("DDPX_METRICS_JSONL")
    config.reset()


def test_recipes_make_covers_every_cli_model():
    from distributed_data_parallel_experiment_b200 import recipes
    cpu = torch.device("cpu")
    o, loss = recipes.make("toy")
    assert loss == "mse" and len(o.train_set) == 2048
    o, loss = recipes.make("vgg", world_size=2, synthetic=True, synthetic_size=16, device=cpu)
    assert loss is None and sum(p.numel() for p in o.model.parameters()) == 9228362 and o.scheduler is not None
    o, loss = recipes.make("deepnn", synthetic=True, synthetic_size=16, device=cpu)
    assert sum(p.numel() for p in o.model.parameters()) == 1186986            # singlegpu.py:18-44
    o, loss = recipes.make("gpt2", synthetic=True, synthetic_size=4, device=cpu, dtype="fp32", seq_len=16)
    x, y = o.train_set[0]
    out = o.model(x[None])
    assert callable(loss) and torch.isfinite(loss(out, y[None]))
    import pytest
    with pytest.raises(ValueError):
        recipes.make("resnet")
    assert set(recipes.MODEL_CHOICES) >= {"vgg", "deepnn", "toy", "mlp", "gpt2", "gpt2-xl"}


def test_entrypoint_flags_parse(monkeypatch):
    """The extra flags exist on both scripts and default to the reference's behaviour."""
    import subprocess, sys
    from pathlib import Path
    repo = Path(__file__).resolve().parent.parent
    for script, extra in (("multigpu.py", ["--bucket-cap-mb", "--allreduce-algo", "--sync-bn", "--backend", "--nprocs"]),
                          ("singlegpu.py", [])):
        out = subprocess.run([sys.executable, str(repo / script), "--help"], capture_output=True, text=True, timeout=120)
        assert out.returncode == 0
        for flag in ["total_epochs", "save_every", "--batch_size", "--model", "--dtype", "--seq-len", "--accum-steps",
                     "--synthetic"] + extra:
            assert flag in out.stdout, (script, flag)
