"""`ddpx_bn_train_fwd_static`: the BatchNorm training forward for callers that cannot hand over a workspace.

Some users of the C ABI have no way to allocate device memory at the point where they call into the library (they were
given their tensors by another component, their memory budget was fixed at start-up, or a cudaMalloc in their loop would
stall the device).  For them the per-channel accumulators sit in an 8 KiB array inside the library instead of a
caller-provided workspace.  This test drives that entry point directly through the C ABI (ctypes, raw pointers) and
checks that it produces what the workspace-taking entry point `ddpx_bn_train_fwd` produces - on two calls in a row, the
second with fewer channels and different data, so that anything the first call left in the shared array would show up.

The file name sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import ctypes

import pytest
import torch

from distributed_data_parallel_experiment_b200 import _lib

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _run_pair(lib, N, H, W, C, relu, pool, seed):
    """Run the existing entry point and the no-allocation variant on identical inputs; return both result sets."""
    torch.manual_seed(seed)
    dev = "cuda"
    x = torch.randn(N, H, W, C, device=dev) * 2.0 + 0.5              # NHWC, what the kernels read
    gamma = torch.rand(C, device=dev) + 0.5
    beta = torch.randn(C, device=dev) * 0.1
    Ho, Wo = (H // 2, W // 2) if pool else (H, W)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    outs = []
    for variant in ("workspace", "static"):
        y = torch.empty(N, Ho, Wo, C, device=dev)
        rm = torch.zeros(C, device=dev); rv = torch.ones(C, device=dev)
        nbt = torch.zeros((), dtype=torch.int64, device=dev)
        mr = torch.empty(2, C, device=dev)
        common = (_p(x), _p(y), _p(gamma), _p(beta), _p(rm), _p(rv), _p(nbt), ctypes.c_float(0.1), ctypes.c_float(1e-5),
                  N, H, W, C, relu, pool, _p(mr))
        if variant == "workspace":
            work = torch.full((2, C), 123.0, device=dev)             # garbage on purpose: the call must clear it
            rc = lib.ddpx_bn_train_fwd(*common, _p(work), stream)
        else:
            rc = lib.ddpx_bn_train_fwd_static(*common, stream)
        assert rc == 0, f"{variant}: return code {rc}"
        torch.cuda.synchronize()
        outs.append((y, rm, rv, nbt, mr))
    # independent oracle: plain PyTorch BatchNorm (+ReLU, +2x2 max-pool) in fp32 on the same data
    import torch.nn.functional as F
    xr = x.permute(0, 3, 1, 2).contiguous()
    rm_ref = torch.zeros(C, device=dev); rv_ref = torch.ones(C, device=dev)
    ref = F.batch_norm(xr, rm_ref, rv_ref, gamma, beta, True, 0.1, 1e-5)
    if relu:
        ref = F.relu(ref)
    if pool:
        ref = F.max_pool2d(ref, 2)
    outs.append((ref.permute(0, 2, 3, 1).contiguous(), rm_ref, rv_ref))
    return outs


def _close(a, b, tol=1e-5):
    return (a.double() - b.double()).abs().max().item() <= tol * max(1.0, b.double().abs().max().item())


def test_static_scratch_variant_matches_workspace_variant_twice_in_a_row():
    torch.cuda.set_device(0)
    lib = _lib.lib()
    # first call: 256 channels, ReLU + pool; second call right after it: fewer channels, other data, no pool
    for (N, H, W, C, relu, pool, seed) in [(8, 16, 16, 256, 1, 1, 0), (4, 8, 8, 64, 0, 0, 1)]:
        (y0, rm0, rv0, n0, mr0), (y1, rm1, rv1, n1, mr1), (yr, rmr, rvr) = _run_pair(lib, N, H, W, C, relu, pool, seed)
        assert _close(y1, y0) and _close(mr1, mr0) and _close(rm1, rm0) and _close(rv1, rv0)
        # and both agree with torch's own BatchNorm, output and running statistics
        assert _close(y1, yr, 1e-4) and _close(rm1, rmr, 1e-4) and _close(rv1, rvr, 1e-4)
        assert int(n0) == 1 and int(n1) == 1
        assert torch.isfinite(y1).all()


def test_static_scratch_variant_rejects_oversize():
    torch.cuda.set_device(0)
    lib = _lib.lib()
    C = 2048                                                         # beyond DDPX_BN_STATIC_MAX_C = 1024
    x = torch.zeros(1, 2, 2, C, device="cuda"); y = torch.empty_like(x)
    g = torch.ones(C, device="cuda"); b = torch.zeros(C, device="cuda")
    rm = torch.zeros(C, device="cuda"); rv = torch.ones(C, device="cuda")
    nbt = torch.zeros((), dtype=torch.int64, device="cuda"); mr = torch.empty(2, C, device="cuda")
    rc = lib.ddpx_bn_train_fwd_static(_p(x), _p(y), _p(g), _p(b), _p(rm), _p(rv), _p(nbt), ctypes.c_float(0.1),
                                      ctypes.c_float(1e-5), 1, 2, 2, C, 0, 0, _p(mr),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 7                                                   # DDPX_ERR_LIMIT
    torch.cuda.synchronize()
    assert int(nbt) == 0                                             # nothing was launched
