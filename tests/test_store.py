"""W5: native control-plane store - primitives in one process, barrier / all-gather across 3 processes."""
import os
import sys
import tempfile
from pathlib import Path

import pytest
import torch.multiprocessing as mp

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

from distributed_data_parallel_experiment_b200.parallel.store import NativeStore  # noqa: E402


def test_primitives_single_process():
    s = NativeStore("127.0.0.1", 0, rank=0, world_size=1)
    assert s.port > 0
    s.set("a", b"hello")
    assert s.get("a") == b"hello"
    big = bytes(range(256)) * 1000                     # larger than the first receive buffer
    s.set("big", big)
    assert s.get("big") == big
    assert s.add("n", 5) == 5 and s.add("n", -2) == 3
    assert s.num_keys() == 3
    assert s.delete("a") and not s.delete("a")
    with pytest.raises(TimeoutError):
        s.get("missing", timeout_s=0.05)
    s.barrier(); s.barrier()                            # world_size 1: returns at once, twice
    assert s.all_gather_object({"x": 1}) == [{"x": 1}]
    second = NativeStore("127.0.0.1", s.port, rank=1, world_size=2, is_server=False)
    second.set("from2", b"\x00\x01")
    assert s.get("from2") == b"\x00\x01"
    second.close(); s.close()


def test_consumed_generations_are_retired():
    """Barrier / gather keys of finished generations are deleted: a long job must not grow the server's map."""
    s = NativeStore("127.0.0.1", 0, rank=0, world_size=1)
    for _ in range(20):
        s.barrier("b")
        s.all_gather(b"x", "g")
    assert s.num_keys() <= 6          # the last two generations of each (2 barrier keys, 1 gather key per generation)
    s.close()


def test_stalled_client_does_not_freeze_the_server():
    """A peer that sends half a request and goes silent is dropped after the receive timeout; the others keep being served."""
    import socket
    import time
    s = NativeStore("127.0.0.1", 0, rank=0, world_size=1)
    bad = socket.create_connection(("127.0.0.1", s.port))
    bad.sendall(b"\x01")               # an opcode and nothing else
    t0 = time.time()
    s.set("k", b"v")                   # may wait for the stalled read to time out, but must come back
    assert s.get("k") == b"v"
    assert time.time() - t0 < 10.0
    bad.close(); s.close()


def _worker(rank, world, port_file, outdir):
    if rank == 0:
        s = NativeStore("127.0.0.1", 0, rank, world)
        with open(port_file + ".tmp", "w") as f:
            f.write(str(s.port))
        os.rename(port_file + ".tmp", port_file)
    else:
        import time
        while not os.path.exists(port_file):
            time.sleep(0.01)
        s = NativeStore("127.0.0.1", int(open(port_file).read()), rank, world)
    got = s.all_gather(f"handle-of-{rank}".encode() * (rank + 1))
    s.barrier("after_gather")
    order = s.broadcast_object([3, 1, 2] if rank == 0 else None)
    total = s.add("sum", rank + 1)
    s.barrier("end")
    final = int.from_bytes(s.get("sum"), "little", signed=True)
    with open(os.path.join(outdir, f"r{rank}"), "w") as f:
        f.write(repr((got, order, final)))
    s.barrier("bye")                                    # rank 0 (the server) must outlive everybody's last request
    s.close()


def test_barrier_and_all_gather_across_processes():
    world = 3
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, os.path.join(d, "port"), d), nprocs=world, join=True)
        outs = [eval(open(os.path.join(d, f"r{r}")).read()) for r in range(world)]
    expect = [f"handle-of-{r}".encode() * (r + 1) for r in range(world)]
    for got, order, final in outs:
        assert got == expect and order == [3, 1, 2] and final == 6


def _fd_child(conn):
    r, w = os.pipe()
    os.write(w, b"through-pidfd")
    conn.send((os.getpid(), r))
    conn.recv()                      # stay alive until the parent has duplicated the descriptor


def test_fd_passing_by_pidfd():
    """The NVLS allocator hands cuMem file descriptors to peers with pidfd_getfd (parallel/nvls.py)."""
    import multiprocessing as pmp
    from distributed_data_parallel_experiment_b200.parallel import nvls
    assert nvls.supported(0) in (True, False)           # never raises, with or without a GPU
    r, w = os.pipe()
    os.write(w, b"self")
    d = nvls._dup_fd_from(os.getpid(), r)
    assert os.read(d, 4) == b"self"
    ctx = pmp.get_context("spawn")
    parent, child = ctx.Pipe()
    p = ctx.Process(target=_fd_child, args=(child,))
    p.start()
    pid, fd = parent.recv()
    try:
        got = nvls._dup_fd_from(pid, fd)
    except OSError as e:                                 # ptrace-restricted sandboxes refuse pidfd_getfd
        parent.send(1); p.join()
        pytest.skip(f"pidfd_getfd not permitted here: {e}")
    assert os.read(got, 13) == b"through-pidfd"
    parent.send(1)
    p.join()
