"""DDIM sampler behind the reference's ``generalized_steps(x, seq, model, alpha, select_index, **kwargs)``
(functions/denoising.py:10-52).

Per step: one ``model(xt, t)`` call and ONE fused update kernel (``ddimb_ddim_step``: x0 prediction, the
alpha_next combination, the eta/sigma noise term and both optional snapshots in a single pass over ``xt``)
instead of the reference's five elementwise launches.  Snapshots leave the GPU through a side stream into
pinned host memory, so the loop never blocks on a device->host copy (the reference's ``.to("cpu")`` does).

Stated differences from the literal reference text:
  * the working tensor is a clone -- the reference aliases (and overwrites) a CUDA fp32 ``x``, so its ``xs[0]``
    ends up holding the final sample; here ``xs[0]`` is the untouched input;
  * the N(0,1) draw of :42 happens only when its coefficient c1 is non-zero (eta > 0), or comes from the
    ``noise=`` keyword (an iterable of per-step tensors) when the caller wants a reproducible stream.
"""
from __future__ import annotations

import torch

from .. import ops


def _ddim_scalars(alpha_list, i, j, eta):
    """functions/denoising.py:22-24,36-40 -- float64 host arithmetic, as the reference."""
    at = alpha_list[int(i) + 1]
    at_next = alpha_list[int(j) + 1]
    c1 = eta * ((1 - at / at_next) * (1 - at_next) / (1 - at)) ** 0.5
    c2 = ((1 - at_next) - c1 ** 2) ** 0.5
    return at, at_next, c1, c2


class _SnapshotPipe:
    """device buffer -> pinned host tensor on a side stream."""

    def __init__(self, device):
        self.stream = torch.cuda.Stream(device=device)
        self.device = device

    def push(self, dev_tensor):
        host = torch.empty(dev_tensor.shape, dtype=dev_tensor.dtype, pin_memory=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ev)
            host.copy_(dev_tensor, non_blocking=True)
            dev_tensor.record_stream(self.stream)
        return host

    def push_into(self, dev_tensor, host_view):
        """Same, into a caller-owned (pinned) host tensor or a dim-0 slice of one."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ev)
            host_view.copy_(dev_tensor, non_blocking=True)
            dev_tensor.record_stream(self.stream)

    def drain(self):
        self.stream.synchronize()


def ddim_loop_device(xt, seq, model, alpha_list, eta=0.0, select_index=None, noise=None, on_snapshot=None):
    """The loop body of functions/denoising.py:21-50 on a DEVICE-resident fp32 ``xt`` (updated in place).

    ``on_snapshot(x0_buf, xs_buf)`` receives device copies of the two tensors the reference appends at :29-34 and
    :45-50 for the selected steps (``select_index`` None = every step; negative indices allowed); with
    ``on_snapshot=None`` no snapshot is taken.  Returns ``xt``."""
    device = xt.device
    seq = list(seq)
    seq_next = [-1] + seq[:-1]
    noise_it = iter(noise) if noise is not None else None
    t = torch.empty(xt.size(0), dtype=torch.int64, device=device)
    L = len(seq)
    for index, (i, j) in enumerate(zip(reversed(seq), reversed(seq_next))):
        t.fill_(int(i))
        at, at_next, c1, c2 = _ddim_scalars(alpha_list, i, j, eta)
        et = model(xt, t)
        selected = on_snapshot is not None and (select_index is None or index in select_index or index - L in select_index)
        z = None
        if noise_it is not None:
            z = next(noise_it).to(device=device, dtype=torch.float32).contiguous()
        elif c1 != 0:
            z = torch.randn_like(xt)
        x0_buf = torch.empty_like(xt) if selected else None
        xs_buf = torch.empty_like(xt) if selected else None
        ops.ddim_step(xt, et.contiguous(), z, at, at_next, c1, c2, x0_out=x0_buf, xs_out=xs_buf)
        if selected:
            on_snapshot(x0_buf, xs_buf)
    return xt


def _chunk_bounds(B, n):
    n = max(1, min(int(n), B))
    base, rem = divmod(B, n)
    out, c0 = [], 0
    for k in range(n):
        c1 = c0 + base + (1 if k < rem else 0)
        out.append((c0, c1))
        c0 = c1
    return out


def generalized_steps(x, seq, model, alpha, select_index, **kwargs):
    """functions/denoising.py:10-52.  Returns ``(xs, x0_preds)``: lists of CPU fp32 tensors, ``xs[0]`` being the input.

    A HOST ``x`` is processed as ``chunks`` independent sample groups (keyword, default 4 when eta == 0 and no explicit
    noise stream is given; samples never interact -- SURVEY.md section 8e): every group's host->device copy is queued up front
    on a copy stream and its snapshots return on another, so the PCIe traffic of one group hides behind the denoiser of the
    next instead of bracketing the whole batch.  A CUDA ``x`` runs as one group."""
    if not torch.cuda.is_available():
        raise RuntimeError("ddim_audio_b200.generalized_steps needs a CUDA device: there is no CPU path")
    eta = kwargs.get("eta", 0)
    noise = kwargs.get("noise", None)
    with torch.no_grad():
        alpha_list = [1.0] + alpha.to("cpu", torch.float32).numpy().tolist()          # :12
        device = x.device if x.is_cuda else torch.device("cuda", torch.cuda.current_device())
        B = x.size(0)
        nchunk = 1
        if not x.is_cuda and noise is None and eta == 0:
            nchunk = kwargs.get("chunks", 4 if B >= 8 else 1)
        bounds = _chunk_bounds(B, nchunk)
        cur = torch.cuda.current_stream(device)
        pipe = _SnapshotPipe(device)
        # ---- inputs: all host->device copies are queued now, on their own stream
        staged = []
        if x.is_cuda:
            staged.append((x.to(dtype=torch.float32, copy=True).contiguous(), None))
        else:
            h2d = torch.cuda.Stream(device=device)
            for c0, c1 in bounds:
                with torch.cuda.stream(h2d):
                    xc = x[c0:c1].to(device=device, dtype=torch.float32, non_blocking=True).contiguous()
                    ev = torch.cuda.Event()
                    ev.record(h2d)
                staged.append((xc, ev))
        # ---- snapshots of step k land in slices [c0:c1] of one full-batch pinned tensor per list entry
        out_x0, out_xs = [], []
        for (c0, c1), (xt, ev) in zip(bounds, staged):
            if ev is not None:
                cur.wait_event(ev)
                xt.record_stream(cur)
            k = [0]

            def snap(x0_buf, xs_buf, c0=c0, c1=c1, k=k):
                if k[0] == len(out_x0):
                    out_x0.append(torch.empty((B,) + tuple(x0_buf.shape[1:]), dtype=torch.float32, pin_memory=True))
                    out_xs.append(torch.empty((B,) + tuple(xs_buf.shape[1:]), dtype=torch.float32, pin_memory=True))
                pipe.push_into(x0_buf, out_x0[k[0]][c0:c1])                             # :29-34
                pipe.push_into(xs_buf, out_xs[k[0]][c0:c1])                             # :45-50
                k[0] += 1

            ddim_loop_device(xt, seq, model, alpha_list, eta, select_index, noise, snap)
        pipe.drain()
    return [x] + out_xs, out_x0
