"""T3: every native sm_100a kernel against a plain PyTorch fp32 (or fp64) reference of the same op.

TF32 tensor-core paths are compared with a relative tolerance of a few 1e-3 (10-bit mantissa products, fp32
accumulation); pure fp32 kernels with ~1e-5.
"""
import ctypes

import pytest
import torch
import torch.nn.functional as F

import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200 import _lib, ops
from distributed_data_parallel_experiment_b200.ops import native

pytestmark = pytest.mark.gpu

TF32_RTOL = 4e-3


def rel_err(a, b):
    return (a.double() - b.double()).abs().max().item() / max(b.double().abs().max().item(), 1e-12)


def rel_l2(a, b):
    """Relative Frobenius error: robust to the few ReLU-mask / pool-argmax flips TF32 rounding causes at near-ties."""
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.fixture(autouse=True)
def _dev():
    torch.cuda.set_device(0)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.cuda.synchronize()
    assert native.read_error_word() == 0


def test_library_loaded_not_fallback():
    assert _lib.available()
    n0 = _lib.launches()
    ops.relu(torch.randn(1000, device="cuda"))
    assert _lib.launches() == n0 + 1


# ------------------------------------------------------------------------------------------------ GEMM
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,N,K", [(512, 768, 1024), (300, 1000, 520), (128, 64, 96), (4096, 256, 2048)])
def test_gemm_layouts(a_mn, b_mn, dtype, M, N, K):
    if dtype == torch.bfloat16 and (M % 8 or N % 8 or K % 8):
        M, N, K = (M + 7) // 8 * 8, (N + 7) // 8 * 8, (K + 7) // 8 * 8
    torch.manual_seed(0)
    A = torch.randn(M, K, device="cuda").to(dtype)
    B = torch.randn(N, K, device="cuda").to(dtype)
    ref = A.double() @ B.double().t()
    As = A.t().contiguous() if a_mn else A
    Bs = B.t().contiguous() if b_mn else B
    D = torch.empty(M, N, device="cuda", dtype=torch.float32)
    native.gemm(As, Bs, D, M, N, K, As.stride(0), Bs.stride(0), N, a_mn, b_mn)
    assert rel_err(D, ref) < (TF32_RTOL if dtype == torch.float32 else 1e-4)


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gemm_cta_pairs_large(a_mn, b_mn, dtype):
    """Shapes big enough for the 256x256 CTA-pair (cta_group::2) tiles, including a ragged M and bf16 output + bias."""
    torch.manual_seed(0)
    M, N, K = 2304 + 40, 2048, 768
    A = torch.randn(M, K, device="cuda").to(dtype)
    B = torch.randn(N, K, device="cuda").to(dtype)
    bias = torch.randn(N, device="cuda")
    ref = A.double() @ B.double().t() + bias.double()
    As = A.t().contiguous() if a_mn else A
    Bs = B.t().contiguous() if b_mn else B
    D = torch.empty(M, N, device="cuda", dtype=dtype)
    native.gemm(As, Bs, D, M, N, K, As.stride(0), Bs.stride(0), N, a_mn, b_mn, bias, 0)
    assert rel_err(D, ref) < (TF32_RTOL if dtype == torch.float32 else 1e-2)


def test_gemm_epilogue_bias_gelu_bf16_out():
    M, N, K = 384, 1024, 512
    A = torch.randn(M, K, device="cuda").bfloat16(); B = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(N, device="cuda")
    ref = F.gelu(A.double() @ B.double().t() + bias.double(), approximate="tanh")
    D = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    native.gemm(A, B, D, M, N, K, K, K, N, 0, 0, bias, 2)
    assert rel_err(D, ref) < 1e-2


def test_gemm_split_k_accumulate():
    M, N, K = 256, 256, 8192
    A = torch.randn(M, K, device="cuda"); B = torch.randn(N, K, device="cuda")
    D = torch.zeros(M, N, device="cuda")
    native.gemm(A, B, D, M, N, K, K, K, N, 0, 0, None, 0, 8)
    assert rel_err(D, A.double() @ B.double().t()) < TF32_RTOL


def test_gemm_small_odd_shapes():
    A = torch.randn(37, 20, device="cuda"); B = torch.randn(1, 20, device="cuda"); bias = torch.randn(1, device="cuda")
    C = torch.empty(37, 1, device="cuda")
    native.gemm_small(A, B, C, 37, 1, 20, 20, 1, 20, 1, 1, bias, 0)
    assert torch.allclose(C, A @ B.t() + bias, atol=1e-5)


# ------------------------------------------------------------------------------------------------ conv
@pytest.mark.parametrize("N,H,W,Cin,Cout", [(4, 32, 32, 64, 128), (4, 16, 16, 128, 256), (8, 8, 8, 256, 512),
                                              (32, 4, 4, 512, 512), (4, 32, 32, 128, 64), (16, 16, 16, 64, 32)])
def test_conv3x3_fprop_dgrad_wgrad(N, H, W, Cin, Cout):
    torch.manual_seed(1)
    x = torch.randn(N, Cin, H, W, device="cuda", requires_grad=True)
    w = (torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.05).requires_grad_()
    b = torch.randn(Cout, device="cuda", requires_grad=True)
    # gradients are compared without the ReLU: TF32 rounding flips the mask of near-zero outputs, which is not a bug
    y = ops.conv3x3(x, w, b, relu=False)
    g = torch.randn_like(y)
    y.backward(g)
    xr = x.detach().double().requires_grad_(); wr = w.detach().double().requires_grad_(); br = b.detach().double().requires_grad_()
    yr = F.conv2d(xr, wr, br, padding=1)
    yr.backward(g.double())
    assert rel_err(y, yr) < TF32_RTOL
    assert rel_err(ops.conv3x3(x.detach(), w.detach(), b.detach(), relu=True), F.relu(yr)) < TF32_RTOL
    assert rel_err(x.grad, xr.grad) < TF32_RTOL
    assert rel_err(w.grad, wr.grad) < TF32_RTOL
    assert rel_err(b.grad, br.grad) < 1e-4


def test_conv_stem_rgb():
    torch.manual_seed(2)
    x = torch.randn(16, 3, 32, 32, device="cuda")
    w = torch.randn(64, 3, 3, 3, device="cuda", requires_grad=True)
    y = ops.conv3x3(x, w, None)
    g = torch.randn_like(y)
    y.backward(g)
    wr = w.detach().double().requires_grad_()
    yr = F.conv2d(x.double(), wr, padding=1)
    yr.backward(g.double())
    assert rel_err(y, yr) < TF32_RTOL and rel_err(w.grad, wr.grad) < TF32_RTOL      # stem runs on TF32 tensor cores too


@pytest.mark.parametrize("Cin,Cout,H,pool", [(3, 64, 32, False), (64, 128, 32, True), (256, 256, 16, True), (512, 512, 4, False)])
@pytest.mark.parametrize("training", [True, False])
def test_conv_bn_relu_pool_block(Cin, Cout, H, pool, training):
    torch.manual_seed(3)
    N = 8
    x = torch.randn(N, Cin, H, H, device="cuda", requires_grad=Cin != 3)
    w = (torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.1).requires_grad_()
    gamma = (torch.rand(Cout, device="cuda") + 0.5).requires_grad_(); beta = (torch.randn(Cout, device="cuda") * 0.1).requires_grad_()
    rm = torch.randn(Cout, device="cuda") * 0.1; rv = torch.rand(Cout, device="cuda") + 0.5
    nbt = torch.tensor(3, device="cuda")
    rm_r, rv_r = rm.clone().double(), rv.clone().double()
    y = ops.conv_bn_act(x, w, gamma, beta, rm, rv, nbt, training, 0.1, 1e-5, relu=True, pool=pool)
    g = torch.randn_like(y)
    y.backward(g)
    xr = x.detach().double().requires_grad_(x.requires_grad); wr = w.detach().double().requires_grad_()
    gr = gamma.detach().double().requires_grad_(); br = beta.detach().double().requires_grad_()
    yr = F.batch_norm(F.conv2d(xr, wr, padding=1), rm_r, rv_r, gr, br, training, 0.1, 1e-5)
    yr = F.relu(yr)
    if pool:
        yr = F.max_pool2d(yr, 2)
    yr.backward(g.double())
    tol = 3e-2        # BN backward amplifies TF32 rounding through 1/sigma; mask/argmax flips at near-ties
    assert rel_l2(y, yr) < 1e-2
    assert rel_l2(w.grad, wr.grad) < tol
    assert rel_l2(gamma.grad, gr.grad) < tol and rel_l2(beta.grad, br.grad) < tol
    if x.requires_grad:
        assert rel_l2(x.grad, xr.grad) < tol
    if training:
        assert rel_err(rm, rm_r) < 3e-3 and rel_err(rv, rv_r) < 3e-3 and int(nbt) == 4
    else:
        assert int(nbt) == 3


def test_standalone_bn_relu_pool_mean_flatten():
    torch.manual_seed(4)
    x = torch.randn(8, 32, 16, 16, device="cuda", requires_grad=True)
    bn = ddpx.nn.BatchNorm2d(32).cuda()
    ref = torch.nn.BatchNorm2d(32).cuda().double()
    y = ops.spatial_mean(ops.max_pool2x2(ops.relu(bn(x))))
    y.sum().backward()
    xr = x.detach().double().requires_grad_()
    yr = F.max_pool2d(F.relu(ref(xr)), 2).mean([2, 3])
    yr.sum().backward()
    assert rel_err(y, yr) < 1e-5 and rel_err(x.grad, xr.grad) < 1e-4
    assert rel_err(bn.running_var, ref.running_var) < 1e-5
    f = ops.flatten_nchw(ops.max_pool2x2(x.detach()))
    assert torch.equal(f, torch.flatten(F.max_pool2d(x.detach(), 2), 1))


# ------------------------------------------------------------------------------------------------ linear / losses / optim
@pytest.mark.parametrize("shape", [(512, 512, 10), (256, 20, 1), (512, 2048, 512)])
def test_linear_fwd_bwd(shape):
    M, K, N = shape
    torch.manual_seed(5)
    x = torch.randn(M, K, device="cuda", requires_grad=True)
    lin = ddpx.nn.Linear(K, N).cuda()
    y = lin(x, act=1)
    g = torch.randn_like(y)
    y.backward(g)
    xr = x.detach().double().requires_grad_()
    wr = lin.weight.detach().double().requires_grad_(); br = lin.bias.detach().double().requires_grad_()
    yr = F.relu(F.linear(xr, wr, br)); yr.backward(g.double())
    tol = 3e-2       # L2: the ReLU mask of near-zero pre-activations may flip under TF32 rounding
    assert rel_l2(y, yr) < tol and rel_l2(x.grad, xr.grad) < tol
    assert rel_l2(lin.weight.grad, wr.grad) < tol and rel_l2(lin.bias.grad, br.grad) < tol


@pytest.mark.parametrize("rows,cols,dtype", [(512, 10, torch.float32), (64, 50304, torch.bfloat16), (300, 1000, torch.float32)])
def test_softmax_cross_entropy(rows, cols, dtype):
    torch.manual_seed(6)
    logits = (torch.randn(rows, cols, device="cuda") * 3).to(dtype).requires_grad_()
    t = torch.randint(0, cols, (rows,), device="cuda")
    loss = ops.cross_entropy(logits, t)
    (loss * 2.0).backward()
    lr = logits.detach().double().requires_grad_()
    ref = F.cross_entropy(lr, t); (ref * 2.0).backward()
    assert abs(loss.item() - ref.item()) < 1e-3 * max(1.0, abs(ref.item()))
    assert rel_err(logits.grad, lr.grad) < (1e-5 if dtype == torch.float32 else 1e-2)


def test_mse_loss():
    p = torch.randn(128, 1, device="cuda", requires_grad=True); t = torch.rand(128, 1, device="cuda")
    loss = ops.mse_loss(p, t); loss.backward()
    pr = p.detach().clone().requires_grad_(); ref = F.mse_loss(pr, t); ref.backward()
    assert torch.allclose(loss, ref, atol=1e-6) and torch.allclose(p.grad, pr.grad, atol=1e-7)


def _clone_model(m):
    import copy
    return copy.deepcopy(m)


@pytest.mark.parametrize("nesterov", [False, True])
def test_sgd_multi_tensor_matches_torch(nesterov):
    torch.manual_seed(7)
    a = torch.nn.Sequential(torch.nn.Linear(33, 17), torch.nn.Linear(17, 5000)).cuda()
    b = _clone_model(a)
    oa = torch.optim.SGD(a.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4, nesterov=nesterov)
    ob = torch.optim.SGD(b.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4, nesterov=nesterov)
    ddpx.optim.fuse(oa)
    sched_a = torch.optim.lr_scheduler.LambdaLR(oa, lambda s: float(s > 0))      # lr == 0 on the first step
    sched_b = torch.optim.lr_scheduler.LambdaLR(ob, lambda s: float(s > 0))
    for step in range(4):
        x = torch.randn(8, 33, device="cuda")
        for m, o, s in ((a, oa, sched_a), (b, ob, sched_b)):
            o.zero_grad(); m(x).square().mean().backward(); o.step(); s.step()
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-6, rtol=1e-5)
        assert torch.allclose(oa.state[p]["momentum_buffer"], ob.state[q]["momentum_buffer"], atol=1e-6, rtol=1e-5)


def test_adamw_multi_tensor_matches_torch():
    torch.manual_seed(8)
    a = torch.nn.Sequential(torch.nn.Linear(33, 17), torch.nn.Linear(17, 3000)).cuda()
    b = _clone_model(a)
    oa = torch.optim.AdamW(a.parameters(), lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1)
    ob = torch.optim.AdamW(b.parameters(), lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1)
    ddpx.optim.fuse(oa)
    for step in range(5):
        x = torch.randn(8, 33, device="cuda")
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad(); m(x).square().mean().backward(); o.step()
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-6, rtol=1e-4)


# ------------------------------------------------------------------------------------------------ whole model
def test_vgg_forward_backward_matches_torch_reference():
    from conftest import reference_module
    ref = reference_module("multigpu")
    torch.manual_seed(9)
    ours = ddpx.models.VGG().cuda()
    theirs = ref.VGG().cuda().double()
    theirs.load_state_dict(ours.state_dict())
    x = torch.rand(64, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (64,), device="cuda")
    lo = ops.cross_entropy(ours(x), y); lo.backward()
    lt = F.cross_entropy(theirs(x.double()), y); lt.backward()
    assert abs(lo.item() - lt.item()) < 5e-3
    # yardstick: torch's own fp32 model with TF32 convolutions (what the reference runs) against the same fp64 oracle
    torch.backends.cudnn.allow_tf32 = True
    tf32 = ref.VGG().cuda(); tf32.load_state_dict(ours.state_dict())
    F.cross_entropy(tf32(x), y).backward()
    torch.backends.cudnn.allow_tf32 = False
    for (n, p), (_, q), (_, r) in zip(ours.named_parameters(), theirs.named_parameters(), tf32.named_parameters()):
        assert rel_l2(p.grad, q.grad) < max(3.0 * rel_l2(r.grad, q.grad), 2e-2), n
    for (n, b1), (_, b2) in zip(ours.named_buffers(), theirs.named_buffers()):
        assert rel_err(b1.double(), b2.double()) < 1e-2, n
    ours.eval(); theirs.eval()
    assert rel_err(ours(x), theirs(x.double())) < 1e-2


def test_trainer_step_loss_goes_down_and_checkpoint_interchange(tmp_path):
    from conftest import reference_module
    ref = reference_module("multigpu")
    torch.manual_seed(10)
    objs = ddpx.recipes.vgg_train_objs(world_size=1, synthetic=True, synthetic_size=0)
    for g in objs.optimizer.param_groups:
        g["lr"] = 0.05
    t = ddpx.Trainer(objs.model, [], objs.optimizer, 0, 1, scheduler=None, distributed=False,
                     checkpoint_path=str(tmp_path / "checkpoint.pt"))
    x = torch.rand(128, 3, 32, 32).pin_memory(); y = torch.randint(0, 10, (128,)).pin_memory()
    losses = [t.train_step(x, y) for _ in range(12)]
    assert losses[-1] < losses[0]
    t._save_checkpoint(0)
    r = ref.VGG()
    r.load_state_dict(torch.load(str(tmp_path / "checkpoint.pt"), map_location="cpu", weights_only=True), strict=True)


def test_deepnn_runs_natively():
    torch.manual_seed(11)
    m = ddpx.models.DeepNN().cuda().eval()
    x = torch.rand(16, 3, 32, 32, device="cuda")
    import copy
    cpu = copy.deepcopy(m).cpu()
    assert rel_err(m(x), cpu(x.cpu()).cuda()) < 2e-2


@pytest.mark.gpu
@pytest.mark.parametrize("N,H,W,Cin,Cout", [(8, 8, 8, 64, 128), (4, 16, 16, 128, 256), (16, 4, 4, 256, 256)])
def test_conv_wgrad_staging_layout_matches_direct(N, H, W, Cin, Cout):
    """ddpx_conv3x3_wgrad_packed + ddpx_unpack_conv_wgrad == ddpx_conv3x3_wgrad (same MMAs, different reduction layout)."""
    from distributed_data_parallel_experiment_b200.ops import native
    torch.manual_seed(3)
    x = torch.randn(N, H, W, Cin, device="cuda")
    dy = torch.randn(N, H, W, Cout, device="cuda")
    direct = torch.zeros(Cout, Cin, 3, 3, device="cuda")
    native.conv3x3_wgrad(dy, x, direct, N, H, W, Cin, Cout)
    staging = torch.zeros(9 * Cout * Cin, device="cuda")
    native.conv3x3_wgrad_packed(dy, x, staging, N, H, W, Cin, Cout)
    out = torch.empty(Cout, Cin, 3, 3, device="cuda")
    native.unpack_conv_wgrad(staging, out, Cout, Cin)
    assert rel_err(out, direct) < 1e-5
    ref = torch.nn.grad.conv2d_weight(x.permute(0, 3, 1, 2).double(), (Cout, Cin, 3, 3), dy.permute(0, 3, 1, 2).double(), padding=1)
    assert rel_err(out, ref) < TF32_RTOL
