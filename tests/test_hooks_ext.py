"""The C++ autograd-hook extension (csrc/torch_ext/autograd_hooks.cpp) on CPU tensors with a stand-in for the native
reducer: hooks fire once per parameter with the gradient's data pointer, only while armed, the first-launch and
all-launched callbacks run once per backward, and removed hooks stay silent."""
import ctypes

import pytest
import torch

from distributed_data_parallel_experiment_b200.parallel import hooks_ext

MARK = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p)


@pytest.fixture(scope="module")
def ext():
    m = hooks_ext.module()
    if m is None:
        pytest.skip("extension could not be built on this host")
    return m


def _backward(net):
    net(torch.randn(5, 4)).sum().backward()


def test_hooks_fire_once_per_parameter_while_armed(ext):
    net = torch.nn.Sequential(torch.nn.Linear(4, 8), torch.nn.Tanh(), torch.nn.Linear(8, 3))
    params = list(net.parameters())
    calls, events = [], []

    def mark(handle, pid, grad_ptr, stream):
        calls.append((handle, pid, grad_ptr, stream))
        flags = 0
        if len(calls) == 1:
            flags |= 1 | 4                      # "the first bucket was launched"
        if len(calls) == len(params):
            flags |= 1 | 2                      # "every bucket has been launched"
        return flags

    cb = MARK(mark)
    hs = ext.HookSet(params, lambda: events.append("first"), lambda: events.append("done"))
    hs.set_native(0xABC, ctypes.cast(cb, ctypes.c_void_p).value)
    _backward(net)                              # not armed: no_sync() behaviour
    assert calls == [] and events == []
    hs.arm(77)
    assert hs.armed()
    _backward(net)
    assert sorted(c[1] for c in calls) == list(range(len(params)))
    assert all(c[0] == 0xABC and c[3] == 77 for c in calls)
    for _, pid, ptr, _ in calls:
        assert ptr == params[pid].grad.data_ptr()
    assert events == ["first", "done"]
    assert not hs.armed()                       # disarmed itself once everything was launched
    calls.clear(); events.clear()
    _backward(net)
    assert calls == [] and events == []
    hs.arm(0)
    hs.remove()
    _backward(net)
    assert calls == []


def test_marked_twice_raises(ext):
    lin = torch.nn.Linear(4, 2)
    cb = MARK(lambda h, pid, g, s: -1)
    hs = ext.HookSet(list(lin.parameters()), None, lambda: None)
    hs.set_native(1, ctypes.cast(cb, ctypes.c_void_p).value)
    hs.arm(0)
    with pytest.raises(RuntimeError, match="marked ready twice"):
        lin(torch.randn(3, 4)).sum().backward()
    hs.remove()
