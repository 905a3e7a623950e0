"""Fused optimizer tail of the train step: ``AdamW`` and ``AdaBelief`` behind the objects the
reference's ``get_optimizer`` returns (functions/__init__.py:5-50), and ``clip_grad_norm_`` as called at
runners/diffusion.py:155-160 -- each ONE or TWO library launches per parameter group over a device chunk table instead of a
Python loop of ~12 torch ops per tensor (AdaBelief.py:82-142 walks ~330 tensors that way).

Both classes are ``torch.optim.Optimizer`` subclasses (LambdaLR, ``zero_grad``, ``param_groups``, ``state_dict`` work) and keep
the reference's per-parameter state layout, so a ``ckpt.pth`` optimizer entry loads either way:
AdamW ``{step (tensor), exp_avg, exp_avg_sq}`` (torch.optim.AdamW), AdaBelief ``{step (int), exp_avg, exp_avg_var}``
(AdaBelief.py:97-109).  CUDA fp32 parameters only; there is no CPU path.
"""
from __future__ import annotations

import numpy as np
import torch
from torch.optim.optimizer import Optimizer

from .. import _C

CHUNK = 65536   # elements per CTA


class ChunkTable:
    """Device table of {param, grad, m, v, count} records for one list of tensors; rebuilt only when a pointer moves."""

    def __init__(self):
        self.key = None
        self.table = None
        self.total = 0

    def get(self, columns, device, check=None):
        """columns: list of 4 lists of equally-shaped tensors (None entries allowed as whole columns).  ``check`` (dtype /
        device / contiguity validation) runs only when a pointer moved and the table is rebuilt."""
        ref = next(c for c in columns if c is not None)
        n_t = len(ref)
        ptrs = np.zeros((n_t, 4), dtype=np.int64)
        for j, col in enumerate(columns):
            if col is not None:
                ptrs[:, j] = [t.data_ptr() for t in col]
        numel = np.array([t.numel() for t in ref], dtype=np.int64)
        key = (ptrs.tobytes(), numel.tobytes())
        if key == self.key:
            return self.table, self.total
        if check is not None:
            check()
        nch = (numel + CHUNK - 1) // CHUNK
        owner = np.repeat(np.arange(n_t), nch)
        first = np.repeat(np.cumsum(nch) - nch, nch)
        off = (np.arange(int(nch.sum())) - first) * CHUNK
        rec = np.zeros((owner.size, 8), dtype=np.int64)
        rec[:, :4] = np.where(ptrs[owner] != 0, ptrs[owner] + 4 * off[:, None], 0)
        rec[:, 4] = np.minimum(CHUNK, numel[owner] - off)
        self.table = torch.from_numpy(rec).to(device)
        self.total = int(numel.sum())
        self.key = key
        return self.table, self.total


def _check(tensors, what):
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError(f"ddim_audio_b200.optim: {what} must be CUDA tensors (there is no CPU path)")
        if t.dtype != torch.float32 or not t.is_contiguous():
            raise RuntimeError(f"ddim_audio_b200.optim: {what} must be contiguous float32")


_clip_tables = {}


@torch.no_grad()
def clip_grad_norm_(parameters, max_norm, norm_type=2.0):
    """``torch.nn.utils.clip_grad_norm_(parameters, max_norm)`` (L2 only): grads scaled in place by
    ``min(1, max_norm / (total_norm + 1e-6))``; returns ``total_norm`` as a 0-d device tensor.  Two launches."""
    if float(norm_type) != 2.0:
        raise NotImplementedError("ddim_audio_b200.clip_grad_norm_: only the L2 norm the reference uses")
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    parameters = list(parameters)
    grads = [p.grad for p in parameters if p.grad is not None]
    if not grads:
        return torch.tensor(0.0)
    if not grads[0].is_cuda:
        raise RuntimeError("ddim_audio_b200.optim: gradients must be CUDA tensors (there is no CPU path)")
    dev = grads[0].device
    key = (dev, tuple(id(p) for p in parameters))        # one table per parameter list: two groups of equal length must not share (and rebuild) one
    slot = _clip_tables.setdefault(key, (ChunkTable(), torch.zeros(1, dtype=torch.float64, device=dev)))
    table, total = slot[0].get([None, grads, None, None], dev, check=lambda: _check(grads, "gradients"))
    norm = torch.empty((), dtype=torch.float32, device=dev)
    _C.call("ddimb_grad_clip", _C.ptr(table), table.shape[0], total, float(max_norm), _C.ptr(slot[1]), _C.ptr(norm))
    return norm


class _Fused(Optimizer):
    KIND = -1
    SECOND = ""

    def __init__(self, params, defaults):
        super().__init__(params, defaults)
        self._tables = {}

    def _init_state(self, p, state):
        raise NotImplementedError

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            if group.get("amsgrad", False):
                raise NotImplementedError("ddim_audio_b200.optim: amsgrad is not on the reference's configured path (audio.yml:90,101)")
            grads = [p.grad for p in ps]
            if not ps[0].is_cuda:
                raise RuntimeError("ddim_audio_b200.optim: parameters must be CUDA tensors (there is no CPU path)")
            states = [self.state[p] for p in ps]
            for p, st in zip(ps, states):
                if len(st) == 0:
                    self._init_state(p, st)
            step = self._bump(states)
            beta1, beta2 = group["betas"]

            def check():
                _check(ps, "parameters")
                _check(grads, "gradients")

            table, total = self._tables.setdefault(gi, ChunkTable()).get(
                [[p.data for p in ps], grads, [st["exp_avg"] for st in states], [st[self.SECOND] for st in states]], ps[0].device, check)
            _C.call("ddimb_optim_step", _C.ptr(table), table.shape[0], total, self.KIND, float(group["lr"]), float(beta1), float(beta2),
                    float(group["eps"]), float(group["weight_decay"]), step)
        _C.bump_param_epoch()       # the masters changed behind torch's version counters: Model re-packs its bf16 operands
        return self._result(loss)


class AdamW(_Fused):
    """``torch.optim.AdamW(params, lr, betas, eps, weight_decay, amsgrad=False)`` (functions/__init__.py:15-23)."""
    KIND, SECOND = 0, "exp_avg_sq"

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, amsgrad=False):
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, amsgrad=amsgrad))

    def _init_state(self, p, st):
        st["step"] = torch.tensor(0.0, dtype=torch.float32)
        st["exp_avg"] = torch.zeros_like(p.data, memory_format=torch.preserve_format)
        st["exp_avg_sq"] = torch.zeros_like(p.data, memory_format=torch.preserve_format)

    def _bump(self, states):
        steps = [st["step"] for st in states]
        torch._foreach_add_(steps, 1)
        lo, hi = int(steps[0]), int(steps[-1])
        if lo != hi:
            raise RuntimeError("ddim_audio_b200.optim: parameters of one group must share a step count")
        return lo

    def _result(self, loss):
        return loss


class AdaBelief(_Fused):
    """``clip_opt.AdaBelief`` as the reference configures it (functions/__init__.py:24-42; AdaBelief.py:5-162): decoupled,
    non-fixed weight decay, no amsgrad, no rectify, ``clip_step=None``.  ``step()`` returns ``{"loss", "norm": None}`` (:162)."""
    KIND, SECOND = 1, "exp_avg_var"

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, weight_decouple=True,
                 fixed_decay=False, rectify=False, clip_step=None, norm_ord=2):
        if rectify:
            raise NotImplementedError(f"AdaBelief does not support rectify={rectify}. use Fasle instead")   # AdaBelief.py:32-35
        if not weight_decouple or fixed_decay or clip_step is not None:
            raise NotImplementedError("ddim_audio_b200.AdaBelief: only weight_decouple=True, fixed_decay=False, clip_step=None "
                                      "(the reference's configuration, functions/__init__.py:36-41, audio.yml:103) has a kernel")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, amsgrad=amsgrad,
                                      weight_decouple=weight_decouple, fixed_decay=fixed_decay, rectify=rectify, clip_step=clip_step,
                                      norm_ord=norm_ord))

    def _init_state(self, p, st):
        st["step"] = 0
        st["exp_avg"] = torch.zeros_like(p.data, memory_format=torch.preserve_format)
        st["exp_avg_var"] = torch.zeros_like(p.data, memory_format=torch.preserve_format)

    def _bump(self, states):
        for st in states:
            st["step"] += 1
        if states[0]["step"] != states[-1]["step"]:
            raise RuntimeError("ddim_audio_b200.optim: parameters of one group must share a step count")
        return int(states[0]["step"])

    def _result(self, loss):
        return {"loss": loss, "norm": None}
