"""Optimiser step and regularisers either side of the per-ray hot path.

`FusedAdam` is a drop-in for the reference's `torch.optim.Adam(grad_vars, betas=(0.9, 0.99))` (train.py:272-273):
same constructor, same `param_groups` (train.py:389-392 decays `param_group["lr"]` in place, honoured every
step), same `state` keys (`step`, `exp_avg`, `exp_avg_sq`) so checkpoints interchange — but `step()` is ONE
kernel launch over a segment table (`ftrf_adam_step`) instead of a chain of per-tensor passes.

`TVLoss` restates loss.py:120-141 as the torch module callers construct (`tvreg = TVLoss()`, train.py:253);
`TensorVMSplit.TV_loss_density(reg)` / `TV_loss_app(reg)` recognise it and run `ftrf_reg_factors` instead.
"""
import ctypes as C

import torch
import torch.nn as nn

from ._abi import FtrfAdamSegment, check, lib


from ._abi import current_stream_handle as _stream  # noqa: E402


class TVLoss(nn.Module):
    """loss.py:120-141.  Generic torch arithmetic (any 4-D tensor); the VM planes take the fused kernel."""

    def __init__(self, TVLoss_weight=1):
        super().__init__()
        self.TVLoss_weight = TVLoss_weight

    def forward(self, x):
        batch_size, _, h_x, w_x = x.shape
        count_h = x.shape[1] * (h_x - 1) * w_x
        count_w = x.shape[1] * h_x * (w_x - 1)
        h_tv = (x[:, :, 1:, :] - x[:, :, :h_x - 1, :]).pow(2).sum()
        w_tv = (x[:, :, :, 1:] - x[:, :, :, :w_x - 1]).pow(2).sum()
        return self.TVLoss_weight * 2 * (h_tv / count_h + w_tv / count_w) / batch_size


def _same_layout(a: torch.Tensor, b: torch.Tensor) -> bool:
    return a.shape == b.shape and a.stride() == b.stride()


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam (no weight decay, no amsgrad) with the whole step in one CUDA launch.

    guard: a model (its `overflow_flag()`), a callable returning a 1-element fp32 CUDA tensor or None, or such a tensor.
    A non-zero value on the device turns the step into a no-op (parameters and moments untouched; the host-side step
    count still advances, so the bias corrections of later steps are one step further than torch's would be) — what
    makes `train_step(wait="counters", allreduce=...)` safe to follow with `step()` without looking at
    `check_steps()`: a step whose lists overflowed on any rank is dropped on every rank."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, guard=None):
        if not 0.0 <= eps or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError("FusedAdam: invalid betas / eps")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self._table = None
        self._table_key = None
        self.guard = guard

    def _guard_tensor(self):
        g = self.guard
        if g is None:
            return None
        if hasattr(g, "overflow_flag"):
            g = g.overflow_flag()
        elif callable(g):
            g = g()
        if g is None:
            return None
        if not (torch.is_tensor(g) and g.is_cuda and g.dtype == torch.float32 and g.numel() == 1):
            raise RuntimeError("FusedAdam: guard must be a 1-element fp32 CUDA tensor")
        return g

    def _init_state(self, p):
        st = self.state[p]
        if len(st) == 0:
            st['step'] = torch.tensor(0.0, dtype=torch.float32)                       # host scalar, as torch keeps it
            st['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # one launch per distinct (betas, eps, step) combination; normally exactly one
        batches = {}
        keep = []
        guard = self._guard_tensor()
        for group in self.param_groups:
            b1, b2 = group['betas']
            for p in group['params']:
                g = p.grad
                if g is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or g.dtype != torch.float32 or g.is_sparse:
                    raise RuntimeError("FusedAdam: fp32 CUDA parameters with dense fp32 gradients only")
                st = self._init_state(p)
                if not _same_layout(g, p):              # all four arrays must share one element order
                    g = torch.empty_like(p, memory_format=torch.preserve_format).copy_(g)
                    keep.append(g)
                if not (_same_layout(st['exp_avg'], p) and _same_layout(st['exp_avg_sq'], p)):
                    for k in ('exp_avg', 'exp_avg_sq'):
                        st[k] = torch.empty_like(p, memory_format=torch.preserve_format).copy_(st[k])
                st['step'] += 1
                key = (float(b1), float(b2), float(group['eps']), int(st['step']))
                batches.setdefault(key, []).append((p, g, st, float(group['lr'])))
        for (b1, b2, eps, t), items in batches.items():
            ptr_key = tuple((p.data_ptr(), g.data_ptr(), st['exp_avg'].data_ptr()) for p, g, st, _ in items)
            if self._table_key != ptr_key or len(batches) > 1:
                table = (FtrfAdamSegment * len(items))()
                for s, (p, g, st, _) in zip(table, items):
                    s.param, s.grad = p.data_ptr(), g.data_ptr()
                    s.exp_avg, s.exp_avg_sq = st['exp_avg'].data_ptr(), st['exp_avg_sq'].data_ptr()
                    s.n = p.numel()
                self._table, self._table_key = table, ptr_key
            table = self._table
            for s, it in zip(table, items):
                s.lr = it[3]
            with torch.cuda.device(items[0][0].device):        # launch where the parameters live
                check(lib.ftrf_adam_step_guarded(table, len(items), b1, b2, eps, t, None if guard is None else guard.data_ptr(),
                                                 _stream()), "adam_step")
        return loss
