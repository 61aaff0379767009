"""Optimizer bridge: keeps the user's stock ``torch.optim.SGD`` / ``AdamW`` object (built on ``model.parameters()``
*before* wrapping, multigpu.py:129-134) as the source of truth for hyper-parameters and state, but runs the update
with one hand-written multi-tensor kernel instead of torch's ``_foreach`` kernels (sgd.py:279-301, K14).

``param_groups[...]['lr']`` is read every step, so ``LambdaLR`` works unchanged - including its ``numpy.float64`` zero
learning rate on the very first step, where momentum buffers must still be initialised.
State lives where ``optimizer.state_dict()`` expects it (``momentum_buffer`` / ``exp_avg`` / ``exp_avg_sq`` / ``step``).
"""
from __future__ import annotations

import types

import torch


def supported(optimizer) -> bool:
    if type(optimizer) is torch.optim.SGD:
        return all(not g.get("maximize", False) for g in optimizer.param_groups)
    if type(optimizer) is torch.optim.AdamW:
        return all(not g.get("amsgrad", False) and not g.get("maximize", False) for g in optimizer.param_groups)
    return False


def _native_step_sgd(opt, grad_scale: float = 1.0):
    from .ops import native
    for group in opt.param_groups:
        params, grads, bufs = [], [], []
        mom = float(group["momentum"])
        first_flags = set()
        for p in group["params"]:
            if p.grad is None:
                continue
            st = opt.state[p]
            first = False
            if mom != 0.0:
                if st.get("momentum_buffer") is None:
                    st["momentum_buffer"] = torch.empty_like(p, memory_format=torch.preserve_format)
                    first = True
                bufs.append(st["momentum_buffer"])
            first_flags.add(first)
            params.append(p.data)
            grads.append(p.grad if p.grad.is_contiguous() else p.grad.contiguous())
        if not params:
            continue
        if len(first_flags) > 1:
            raise RuntimeError("native SGD: parameters with and without momentum state in one group are not supported")
        native.sgd_multi(params, grads, bufs if mom != 0.0 else None, float(group["lr"]), mom, float(group["dampening"]),
                         float(group["weight_decay"]), bool(group["nesterov"]), first_flags.pop(), grad_scale)


def _native_step_adamw(opt, grad_scale: float = 1.0):
    from .ops import native
    for group in opt.param_groups:
        params, grads, m, v = [], [], [], []
        step_val = None
        for p in group["params"]:
            if p.grad is None:
                continue
            st = opt.state[p]
            if len(st) == 0:
                st["step"] = torch.tensor(0.0, dtype=torch.float32)
                st["exp_avg"] = native.zero_(torch.empty_like(p, memory_format=torch.preserve_format))
                st["exp_avg_sq"] = native.zero_(torch.empty_like(p, memory_format=torch.preserve_format))
            st["step"] += 1
            s = int(st["step"].item()) if st["step"].device.type == "cpu" else None
            if s is None:
                raise RuntimeError("native AdamW keeps `step` on the host (capturable=False)")
            if step_val is None:
                step_val = s
            elif step_val != s:
                raise RuntimeError("native AdamW: all parameters of a group must share the step count")
            params.append(p.data)
            grads.append(p.grad if p.grad.is_contiguous() else p.grad.contiguous())
            m.append(st["exp_avg"])
            v.append(st["exp_avg_sq"])
        if not params:
            continue
        b1, b2 = group["betas"]
        native.adamw_multi(params, grads, m, v, float(group["lr"]), float(b1), float(b2), float(group["eps"]),
                           float(group["weight_decay"]), step_val, grad_scale)


def fuse(optimizer: torch.optim.Optimizer) -> torch.optim.Optimizer:
    """Re-route ``optimizer.step()`` to the native multi-tensor kernel for fp32 CUDA parameters.  Idempotent; a no-op for
    optimizers / devices / dtypes the kernels do not cover (those keep torch's own step)."""
    if getattr(optimizer, "_ddpx_fused", False) or not supported(optimizer):
        return optimizer
    ps = [p for g in optimizer.param_groups for p in g["params"]]
    dtypes = {p.dtype for p in ps}
    if not ps or not all(p.is_cuda for p in ps) or len(dtypes) != 1 or dtypes.pop() not in (torch.float32, torch.bfloat16):
        return optimizer
    inner = _native_step_sgd if type(optimizer) is torch.optim.SGD else _native_step_adamw
    orig_step = optimizer.step

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        self._opt_called = True          # what LambdaLR's own wrapper records (lr_scheduler.py:165)
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if getattr(self, "_ddpx_skip_step", False):     # the fused all-reduce kernel already applied this update
            self._ddpx_skip_step = False
            return loss
        inner(self)
        return loss

    step._wrapped_by_lr_sched = True     # keep lr_scheduler's call-order check quiet: we preserve its bookkeeping
    optimizer._ddpx_orig_step = orig_step
    optimizer.step = types.MethodType(step, optimizer)
    optimizer._ddpx_fused = True
    return optimizer
