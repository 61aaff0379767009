"""T6: failure detection - heartbeat watchdog, env-driven fault injection, spawn fail-fast, replica checksums (CPU, gloo)."""
import os
import socket
import sys
import time
from pathlib import Path

import pytest
import torch

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

import distributed_data_parallel_experiment_b200 as ddpx  # noqa: E402
from distributed_data_parallel_experiment_b200.utils import debug, watchdog  # noqa: E402


def test_watchdog_fires_when_progress_stops():
    got = []
    wd = watchdog.StepWatchdog(0.15, rank=3, describe=lambda: "bucket 2 of 3", on_timeout=got.append, poll_s=0.02).start()
    wd.beat("step 7 begins")
    time.sleep(0.5)
    wd.stop()
    assert wd.fired and len(got) == 1
    assert "rank3" in got[0] and "step 7 begins" in got[0] and "bucket 2 of 3" in got[0]


def test_watchdog_quiet_while_beating_or_paused():
    got = []
    wd = watchdog.StepWatchdog(0.2, on_timeout=got.append, poll_s=0.02).start()
    for _ in range(15):
        wd.beat("tick")
        time.sleep(0.03)
    wd.pause()
    time.sleep(0.4)
    wd.resume()
    wd.stop()
    assert not got and not wd.fired


def test_fault_plan_parsing():
    p = watchdog.FaultPlan.parse("kill:rank=1:step=3")
    assert (p.kind, p.rank, p.step) == ("kill", 1, 3)
    assert watchdog.FaultPlan.parse("") is None and watchdog.FaultPlan.parse(None) is None
    with pytest.raises(ValueError):
        watchdog.FaultPlan.parse("explode:rank=0")
    with pytest.raises(RuntimeError):
        watchdog.FaultPlan.parse("raise:rank=0:step=2").maybe_fire(0, 2)
    watchdog.FaultPlan.parse("raise:rank=0:step=2").maybe_fire(1, 2)        # other rank: nothing happens


def _port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _w_train(rank, world, port, diverge):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    ddpx.config.reset()
    ddpx.ddp_setup(rank, world, backend="gloo")
    torch.manual_seed(0)
    model = ddpx.models.ToyLinear()
    opt = torch.optim.SGD(model.parameters(), lr=0.01)
    tr = ddpx.Trainer(model, [], opt, rank, 1, loss_fn="mse")
    if diverge and rank == 1:
        with torch.no_grad():
            model.net.weight.add_(1.0)          # a replica that silently drifted
    g = torch.Generator().manual_seed(rank)
    for _ in range(5):
        tr._run_batch(torch.randn(8, 20, generator=g), torch.randn(8, 1, generator=g))
    ddpx.teardown()


def _spawn(fn, *args, env=None):
    old = {k: os.environ.get(k) for k in (env or {})}
    os.environ.update(env or {})
    try:
        ddpx.spawn(fn, args=(2, _port()) + args, nprocs=2)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def test_killed_rank_fails_the_job_fast():
    t0 = time.time()
    with pytest.raises(Exception) as ei:
        _spawn(_w_train, False, env={"DDPX_FAULT": "kill:rank=1:step=2"})
    assert "exit code 41" in str(ei.value) or "41" in str(ei.value)
    assert time.time() - t0 < 120


def test_hung_rank_is_caught_by_the_watchdog():
    with pytest.raises(Exception) as ei:
        _spawn(_w_train, False, env={"DDPX_FAULT": "hang:rank=1:step=2", "DDPX_WATCHDOG_S": "3"})
    assert str(watchdog.WATCHDOG_EXIT_CODE) in str(ei.value)


def test_replica_checksum_passes_in_sync_and_catches_divergence():
    _spawn(_w_train, False, env={"DDPX_DEBUG_CHECKSUM": "1"})
    with pytest.raises(Exception) as ei:
        _spawn(_w_train, True, env={"DDPX_DEBUG_CHECKSUM": "1"})
    assert "replica divergence" in str(ei.value)


def test_nvtx_range_is_a_noop_without_cuda():
    with debug.nvtx_range("x"):
        pass
    assert debug.replica_checksums([torch.ones(3), torch.full((2,), 2.0)]) == [7.0, 11.0]
