"""T3 (transformer side): LayerNorm, GELU-fused Linear, embedding, causal attention, padded cross-entropy, AdamW bf16,
and the MLP / GPT-2 models end to end against torch references."""
import copy

import pytest
import torch
import torch.nn.functional as F

import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200.ops import functional as FX
from distributed_data_parallel_experiment_b200.ops import native

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.fixture(autouse=True)
def _dev():
    torch.cuda.set_device(0)
    yield
    torch.cuda.synchronize()
    assert native.read_error_word() == 0
    assert native.attn_error_word() == 0


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("rows,C", [(64, 768), (200, 1600), (33, 128)])
def test_layernorm(dtype, tol, rows, C):
    torch.manual_seed(0)
    x = torch.randn(rows, C, device="cuda").to(dtype).requires_grad_()
    w = (torch.rand(C, device="cuda") + 0.5).to(dtype).requires_grad_(); b = torch.randn(C, device="cuda").to(dtype).requires_grad_()
    y = FX.layer_norm(x, w, b, 1e-5)
    g = torch.randn_like(y)
    y.backward(g)
    xr, wr, br = (t.detach().double().requires_grad_() for t in (x, w, b))
    yr = F.layer_norm(xr, (C,), wr, br, 1e-5); yr.backward(g.double())
    assert rel_l2(y, yr) < tol and rel_l2(x.grad, xr.grad) < tol
    assert rel_l2(w.grad, wr.grad) < tol and rel_l2(b.grad, br.grad) < tol


@pytest.mark.parametrize("rows,C", [(2048, 1600), (8192, 768), (300, 512)])
def test_layernorm_backward_with_residual_gradient(rows, C):
    """ddpx_layernorm_bwd_res at the shapes that take the two-kernel form (few rows per warp) and the fused form: dx with
    the residual branch's gradient added, dgamma / dbeta overwritten (no zeroing), against float64 torch."""
    torch.manual_seed(5)
    x = torch.randn(rows, C, device="cuda").bfloat16()
    w = (torch.rand(C, device="cuda") + 0.5).bfloat16(); b = torch.randn(C, device="cuda").bfloat16()
    dy = torch.randn(rows, C, device="cuda").bfloat16(); dres = torch.randn(rows, C, device="cuda").bfloat16()
    y = torch.empty_like(x); stats = torch.empty(2, rows, device="cuda")
    native.layernorm_fwd(x, w, b, y, stats[0], stats[1], rows, C, 1e-5)
    dx = torch.empty_like(x)
    dgb = torch.full((2, C), 7.0, device="cuda")                     # garbage: must be overwritten
    scratch = torch.empty(native.layernorm_bwd_scratch_floats(rows, C), device="cuda")
    native.layernorm_bwd(x, dy, w, stats[0], stats[1], dx, dgb[0], dgb[1], rows, C, scratch, dres)
    xr, wr, br = (t.double().requires_grad_() for t in (x, w, b))
    F.layer_norm(xr, (C,), wr, br, 1e-5).backward(dy.double())
    assert rel_l2(dx, xr.grad + dres.double()) < 1e-2
    assert rel_l2(dgb[0], wr.grad) < 5e-3 and rel_l2(dgb[1], br.grad) < 5e-3


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 5e-3), (torch.bfloat16, 2e-2)])
def test_linear_gelu(dtype, tol):
    torch.manual_seed(1)
    x = torch.randn(256, 512, device="cuda").to(dtype).requires_grad_()
    lin = ddpx.nn.Linear(512, 1024).cuda().to(dtype)
    y = lin(x, act=2)
    g = torch.randn_like(y)
    y.backward(g)
    xr = x.detach().double().requires_grad_(); wr = lin.weight.detach().double().requires_grad_(); br = lin.bias.detach().double().requires_grad_()
    yr = F.gelu(F.linear(xr, wr, br), approximate="tanh"); yr.backward(g.double())
    assert rel_l2(y, yr) < tol and rel_l2(x.grad, xr.grad) < tol
    assert rel_l2(lin.weight.grad, wr.grad) < tol and rel_l2(lin.bias.grad, br.grad) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 5e-3), (torch.bfloat16, 3e-2)])
# the last case has 600 work items (more than the persistent kernels' 148 / 296 CTAs: several rounds of the snake schedule,
# the last one partial) and a ragged sequence length
@pytest.mark.parametrize("B,T,H,hd", [(2, 128, 2, 64), (2, 256, 4, 64), (1, 1024, 2, 64), (2, 200, 3, 64), (1, 128, 2, 32),
                                      (3, 1000, 25, 64)])
def test_causal_attention(dtype, tol, B, T, H, hd):
    torch.manual_seed(2)
    C = H * hd
    qkv = torch.randn(B, T, 3 * C, device="cuda").to(dtype).requires_grad_()
    y = FX.causal_attention(qkv, H)
    g = torch.randn_like(y)
    y.backward(g)
    qr = qkv.detach().double().requires_grad_()
    q, k, v = qr.split(C, dim=2)
    q = q.view(B, T, H, hd).transpose(1, 2); k = k.view(B, T, H, hd).transpose(1, 2); v = v.view(B, T, H, hd).transpose(1, 2)
    yr = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, C)
    yr.backward(g.double())
    assert rel_l2(y, yr) < tol
    assert rel_l2(qkv.grad, qr.grad) < tol
    C1 = H * hd          # each of dq / dk / dv on its own (a wrong third would hide inside the norm of the whole)
    for part in range(3):
        assert rel_l2(qkv.grad[..., part * C1:(part + 1) * C1], qr.grad[..., part * C1:(part + 1) * C1]) < tol


def test_flash_attention_matches_materialised_path(monkeypatch):
    """The fused kernels against the older batched-GEMM + softmax path on the same bf16 inputs."""
    from distributed_data_parallel_experiment_b200.ops import transformer as TR
    torch.manual_seed(5)
    B, T, H = 2, 384, 4
    qkv = torch.randn(B, T, 3 * H * 64, device="cuda").bfloat16()
    g = torch.randn(B, T, H * 64, device="cuda").bfloat16()
    res = []
    for flash in (True, False):
        monkeypatch.setattr(TR, "_FLASH", flash)
        x = qkv.clone().requires_grad_()
        y = FX.causal_attention(x, H)
        y.backward(g)
        res.append((y.detach(), x.grad))
    assert rel_l2(res[0][0], res[1][0]) < 2e-2 and rel_l2(res[0][1], res[1][1]) < 3e-2


@pytest.mark.parametrize("B,T,C,H", [(2, 256, 128, 2), (1, 200, 256, 4), (4, 1024, 768, 12)])
def test_gpt2_block_single_call_matches_per_op_path(monkeypatch, B, T, C, H):
    """The block as one autograd node over two native calls (blocks_lm.cu) against the same block run op by op."""
    from distributed_data_parallel_experiment_b200.ops import transformer as TR
    from distributed_data_parallel_experiment_b200.models.gpt2 import Block, GPT2Config
    torch.manual_seed(11)
    blk = Block(GPT2Config(n_embd=C, n_head=H, n_layer=1)).cuda().bfloat16()
    for p in blk.parameters():
        torch.nn.init.normal_(p, std=0.05 if p.dim() > 1 else 0.1)      # non-trivial biases and LayerNorm parameters
    blk.ln_1.weight.data.add_(1.0); blk.ln_2.weight.data.add_(1.0)
    x = torch.randn(B, T, C, device="cuda").bfloat16()
    g = torch.randn(B, T, C, device="cuda").bfloat16()
    res = []
    for fused in (True, False):
        monkeypatch.setattr(TR, "_BLOCK_CALLS", fused)
        blk.zero_grad()
        xi = x.clone().requires_grad_()
        y = blk(xi)
        y.backward(g)
        res.append((y.detach(), xi.grad, {n: p.grad.clone() for n, p in blk.named_parameters()}))
    # oracle: the same block in float64 torch math
    P = {n: p.detach().double().requires_grad_() for n, p in blk.named_parameters()}
    xr = x.double().requires_grad_()
    h = F.layer_norm(xr, (C,), P["ln_1.weight"], P["ln_1.bias"], blk.ln_1.eps)
    q, k, v = F.linear(h, P["attn.c_attn.weight"], P["attn.c_attn.bias"]).split(C, dim=2)
    q, k, v = (t.view(B, T, H, C // H).transpose(1, 2) for t in (q, k, v))
    a = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, C)
    x1 = xr + F.linear(a, P["attn.c_proj.weight"], P["attn.c_proj.bias"])
    h2 = F.layer_norm(x1, (C,), P["ln_2.weight"], P["ln_2.bias"], blk.ln_2.eps)
    m = F.linear(F.gelu(F.linear(h2, P["mlp.c_fc.weight"], P["mlp.c_fc.bias"]), approximate="tanh"),
                 P["mlp.c_proj.weight"], P["mlp.c_proj.bias"])
    yr = x1 + m
    yr.backward(g.double())
    for fused, (y, dx, grads) in zip((True, False), res):
        assert rel_l2(y, yr) < 1e-2, fused
        assert rel_l2(dx, xr.grad) < 3e-2, fused
    # every parameter gradient of the single-call path is as close to the float64 oracle as the per-op path's is (both are
    # chains of bf16 roundings: they differ from each other by about as much as each differs from the oracle)
    for n in res[0][2]:
        e_fused, e_ops = rel_l2(res[0][2][n], P[n].grad), rel_l2(res[1][2][n], P[n].grad)
        assert e_fused < 4e-2 and e_fused < 2.0 * e_ops + 5e-3, (n, e_fused, e_ops)


@pytest.mark.parametrize("Mo,No,Kr,split", [(768, 3072, 8192, 2), (1600, 1600, 2048, 3), (200, 96, 4096, 8), (2304, 768, 4096, 1)])
def test_splitk_gemm_with_in_kernel_fixup(Mo, No, Kr, split):
    """dW = d^T x (both operands MN-major) split along the token dimension; the last-arriving epilogue warp converts the fp32
    partial sums to bf16 and leaves workspace and counters at zero, so the same scratch serves the next call."""
    torch.manual_seed(7)
    d = (torch.randn(Kr, Mo, device="cuda") * 0.1).bfloat16()
    x = (torch.randn(Kr, No, device="cuda") * 0.1).bfloat16()
    ws = torch.zeros(Mo * No, device="cuda", dtype=torch.float32)
    cnt = torch.zeros(16 * ((Mo + 127) // 128 * ((No + 63) // 64) + 1), device="cuda", dtype=torch.int32)
    ref = d.double().t() @ x.double()
    for _ in range(2):                                  # second call reuses the scratch the first one left behind
        out = torch.empty(Mo, No, device="cuda", dtype=torch.bfloat16)
        native.gemm_splitk_bf16(d, x, out, Mo, No, Kr, Mo, No, No, 1, 1, split, ws, cnt)
        torch.cuda.synchronize()
        assert rel_l2(out, ref) < 1e-2
        assert float(ws.abs().max()) == 0.0 and int(cnt.abs().max()) == 0


@pytest.mark.parametrize("M,N,K", [(2048, 1600, 1600), (300, 200, 136), (1024, 3072, 768)])
def test_gemm_with_fused_epilogue_passes(M, N, K):
    """Residual add, multiplication by gelu'(pre-activation) and the GELU with two outputs, all inside the GEMM's store
    (whole tiles take the staged vector path, the odd shape the per-element one)."""
    torch.manual_seed(11)
    x = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    b = (torch.randn(N, device="cuda") * 0.1).bfloat16()
    aux = torch.randn(M, N, device="cuda").bfloat16()
    lin = x.double() @ w.double().t() + b.double()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    native.gemm_fused(x, w, out, M, N, K, K, K, N, bias=b, aux=aux, aux_op=1)
    assert rel_l2(out, lin + aux.double()) < 6e-3
    native.gemm_fused(x, w, out, M, N, K, K, K, N, aux=aux, aux_op=2)
    a = aux.double().requires_grad_()
    F.gelu(a, approximate="tanh").sum().backward()
    assert rel_l2(out, (lin - b.double()) * a.grad) < 8e-3
    pre = torch.empty_like(out)
    native.gemm_fused(x, w, out, M, N, K, K, K, N, bias=b, act=2, out2=pre)
    torch.cuda.synchronize()
    assert rel_l2(pre, lin) < 6e-3
    assert rel_l2(out, F.gelu(lin, approximate="tanh")) < 8e-3
    assert native.read_error_word() == 0


def test_embedding_sum_and_padded_ce():
    torch.manual_seed(3)
    V, Vp, C, B, T = 1000, 1024, 128, 4, 64
    wte = (torch.randn(Vp, C, device="cuda") * 0.1).bfloat16().requires_grad_()
    wpe = (torch.randn(128, C, device="cuda") * 0.1).bfloat16().requires_grad_()
    idx = torch.randint(0, V, (B, T), device="cuda")
    tgt = torch.randint(0, V, (B, T), device="cuda")
    x = FX.embedding_sum(idx, wte, wpe)
    logits = ddpx.ops.linear(x, wte, None)
    loss = FX.cross_entropy_padded(logits, tgt, V)
    loss.backward()
    wr = wte.detach().double().requires_grad_(); pr = wpe.detach().double().requires_grad_()
    xr = F.embedding(idx, wr) + pr[:T].unsqueeze(0)
    lr = F.cross_entropy((xr @ wr.t())[..., :V].reshape(-1, V), tgt.reshape(-1)); lr.backward()
    assert abs(loss.item() - lr.item()) < 2e-2
    assert rel_l2(wte.grad, wr.grad) < 5e-2 and rel_l2(wpe.grad, pr.grad) < 5e-2


def test_mlp_model_bf16_step_matches_torch():
    torch.manual_seed(4)
    m = ddpx.models.MLP(256, 512, 4, 128, dtype=torch.bfloat16).cuda()
    ref = copy.deepcopy(m).double()
    x = torch.randn(128, 256, device="cuda").bfloat16(); y = torch.randint(0, 128, (128,), device="cuda")
    loss = ddpx.ops.cross_entropy(m(x), y); loss.backward()
    h = x.double()
    for i, l in enumerate(ref.layers):
        h = F.linear(h, l.weight, l.bias)
        if i < len(ref.layers) - 1:
            h = F.relu(h)
    lr = F.cross_entropy(h, y); lr.backward()
    assert abs(loss.item() - lr.item()) < 3e-2
    for p, q in zip(m.parameters(), ref.parameters()):
        assert rel_l2(p.grad, q.grad) < 6e-2


def test_gpt2_tiny_matches_cpu_reference_and_trains():
    torch.manual_seed(5)
    cfg = ddpx.models.GPT2Config.tiny()
    m = ddpx.models.GPT2(cfg, dtype=torch.float32).cuda()
    ref = copy.deepcopy(m).cpu()                     # CPU path = plain torch ops with the same parameters
    idx = torch.randint(0, cfg.vocab_size, (2, 128), device="cuda"); tgt = torch.randint(0, cfg.vocab_size, (2, 128), device="cuda")
    loss = m(idx, tgt); loss.backward()
    lr = ref(idx.cpu(), tgt.cpu()); lr.backward()
    assert abs(loss.item() - lr.item()) < 2e-2
    for (n, p), (_, q) in zip(m.named_parameters(), ref.named_parameters()):
        assert rel_l2(p.grad.cpu(), q.grad) < 5e-2, n
    # bf16 + native AdamW: the loss must go down on a fixed batch
    mb = ddpx.models.GPT2(cfg, dtype=torch.bfloat16).cuda()
    opt = ddpx.optim.fuse(torch.optim.AdamW(mb.parameters(), lr=3e-3, weight_decay=0.0))
    losses = []
    for _ in range(10):
        opt.zero_grad(); l = mb(idx, tgt); l.backward(); opt.step(); losses.append(l.item())
    assert losses[-1] < losses[0] - 0.3


def test_adamw_bf16_matches_torch():
    torch.manual_seed(6)
    a = torch.nn.Linear(64, 256).cuda().bfloat16(); b = copy.deepcopy(a)
    oa = ddpx.optim.fuse(torch.optim.AdamW(a.parameters(), lr=1e-2, weight_decay=0.1)); ob = torch.optim.AdamW(b.parameters(), lr=1e-2, weight_decay=0.1)
    for _ in range(3):
        x = torch.randn(16, 64, device="cuda").bfloat16()
        for mdl, o in ((a, oa), (b, ob)):
            o.zero_grad(); mdl(x).float().square().mean().backward(); o.step()
    for p, q in zip(a.parameters(), b.parameters()):
        assert rel_l2(p, q) < 2e-2


def test_dropout_mask_consistency():
    x = torch.ones(1 << 16, device="cuda", requires_grad=True)
    y = FX.dropout(x, 0.25)
    y.sum().backward()
    keep = (y != 0).float().mean().item()
    assert abs(keep - 0.75) < 0.02
    assert torch.equal((x.grad != 0), (y != 0)) and torch.allclose(y[y != 0], torch.tensor(1 / 0.75, device="cuda"))
