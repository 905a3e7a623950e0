"""``EMAHelper`` behind the reference's ``models/ema.py`` (:4-60): same methods, same ``shadow`` dict (name -> tensor, what the
reference stores as the fifth entry of ``ckpt.pth``), same arithmetic bit for bit -- but ``update`` is ONE library launch
(``ddimb_ema_update``) over a device chunk table built at ``register`` instead of three torch ops per parameter.
CUDA fp32 parameters only; there is no CPU path."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import _C

CHUNK = 16384   # elements per CTA


def _unwrap(module):
    return module.module if isinstance(module, (nn.DataParallel, nn.parallel.DistributedDataParallel)) else module


class EMAHelper(object):
    def __init__(self, mu=0.999):
        self.mu = mu
        self.shadow = {}
        self._named = None      # (module, [(name, parameter)]) of the last update
        self._table = None      # (device int64 [nchunks, 4], total elements, key of the (param, shadow) pointers it was built for)

    def register(self, module):
        """:9-14."""
        for name, param in _unwrap(module).named_parameters():
            if param.requires_grad:
                self.shadow[name] = param.data.clone()
        self._table = None

    def _chunks(self, module):
        inner = _unwrap(module)
        if self._named is None or self._named[0] is not inner:                      # the module walk is the slow part: once per model
            self._named = (inner, [(n, p) for n, p in inner.named_parameters()])
        pairs = [(p.data, self.shadow[n]) for n, p in self._named[1] if p.requires_grad]
        key = tuple((p.data_ptr(), s.data_ptr(), p.numel()) for p, s in pairs)
        if self._table is not None and self._table[2] == key:
            return self._table
        rows, total = [], 0
        for p, s in pairs:
            if not (p.is_cuda and s.is_cuda):
                raise RuntimeError("ddim_audio_b200.EMAHelper: parameters and shadow must be CUDA tensors (there is no CPU path)")
            if p.dtype != torch.float32 or s.dtype != torch.float32 or not p.is_contiguous() or not s.is_contiguous() or s.shape != p.shape:
                raise RuntimeError("ddim_audio_b200.EMAHelper: expected contiguous float32 parameter / shadow pairs of equal shape")
            n = p.numel()
            total += n
            for off in range(0, n, CHUNK):
                rows.append((p.data_ptr() + 4 * off, s.data_ptr() + 4 * off, min(CHUNK, n - off), 0))
        if not rows:
            raise RuntimeError("ddim_audio_b200.EMAHelper: nothing registered")
        table = torch.tensor(rows, dtype=torch.int64).to(pairs[0][0].device)
        self._table = (table, total, key)
        return self._table

    @torch.no_grad()
    def update(self, module):
        """:16-23 -- ``shadow = (1 - mu) * param + mu * shadow`` for every trainable parameter, one launch."""
        table, total, _ = self._chunks(module)
        _C.call("ddimb_ema_update", _C.ptr(table), table.shape[0], total, float(self.mu))

    def ema(self, module):
        """:25-30."""
        for name, param in _unwrap(module).named_parameters():
            if param.requires_grad:
                param.data.copy_(self.shadow[name].data)
        _C.bump_param_epoch()       # ``.data.copy_`` leaves ``_version`` alone: tell Model its packed operands are stale
        mark = getattr(_unwrap(module), "mark_params_dirty", None)
        if mark is not None:
            mark()

    def ema_copy(self, module):
        """:32-46."""
        inner = _unwrap(module)
        module_copy = type(inner)(inner.config).to(next(inner.parameters()).device)
        module_copy.load_state_dict(inner.state_dict())
        self.ema(module_copy)
        return module_copy

    def state_dict(self):
        return self.shadow

    def load_state_dict(self, state_dict):
        self.shadow = state_dict
        self._table = None
