"""DDIM sampler behind the reference's ``generalized_steps(x, seq, model, alpha, select_index, **kwargs)``
(functions/denoising.py:10-52).

Per step: one ``model(xt, t)`` call and ONE fused update kernel (``ddimb_ddim_step``: x0 prediction, the
alpha_next combination, the eta/sigma noise term and both optional snapshots in a single pass over ``xt``)
instead of the reference's five elementwise launches.  Snapshots leave the GPU through a side stream into
pinned host memory, so the loop never blocks on a device->host copy (the reference's ``.to("cpu")`` does).

Runtime around the loop (none of it changes a result):
  * sample groups -- samples never interact, so a batch that does not fit beside its activations
    (the reference's ``sample_sequence`` makes ALL of ``num_samples`` one batch, runners/diffusion.py:377-398) is walked as
    sequential groups sized from free device memory; a HOST ``x`` is also pipelined group by group over PCIe;
  * CUDA graph -- for small groups (launch-bound: ~260 kernels of a few microseconds each) the denoiser forward is captured
    once per (group shape, weight version) and replayed; the per-step update stays a plain launch with host scalars;
  * ``out=(xs, x0_preds)`` -- recycle the pinned snapshot tensors of an earlier call instead of pinning fresh host memory.

Stated differences from the literal reference text:
  * the working tensor is a clone -- the reference aliases (and overwrites) a CUDA fp32 ``x``, so its ``xs[0]``
    ends up holding the final sample; here ``xs[0]`` is the untouched input;
  * the N(0,1) draw of :42 happens only when its coefficient c1 is non-zero (eta > 0), or comes from the
    ``noise=`` keyword (an iterable of per-step tensors) when the caller wants a reproducible stream.
    ``match_reference_rng=True`` restores the reference's generator consumption: one ``torch.randn_like(x)`` per step even at
    eta = 0 (drawn and discarded -- the add is skipped, not the draw), one sample group, so the CUDA generator ends in the
    state the reference leaves it in.
"""
from __future__ import annotations

import torch

from .. import ops

GRAPH_MAX_POSITIONS = 4 * 8192          # B * T up to which the forward is replayed from a CUDA graph by default
BYTES_PER_POSITION = 340                # device bytes per (sample, frame, bin) of a forward: skips + block temporaries + fp32 x / eps / xt
                                        # (measured peak at x[64,2,8192,256]: 37.4 GB = 279 B per position; kept with a margin)


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

    def push_into(self, dev_tensor, host_view):
        """Into a caller-owned (pinned) host tensor or a dim-0 slice of one."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ev)
            host_view.copy_(dev_tensor, non_blocking=True)
            dev_tensor.record_stream(self.stream)

    def drain(self):
        self.stream.synchronize()


class _GraphedForward:
    """``model(xt, t)`` captured once for fixed ``xt`` / ``t`` tensors (the loop updates both in place) and replayed per step."""

    def __init__(self, model, xt, t):
        cur = torch.cuda.current_stream(xt.device)
        side = torch.cuda.Stream(device=xt.device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(2):                  # packs weights, builds constants, sets the kernels' shared-memory attributes
                model(xt, t)
        cur.wait_stream(side)
        self.pack = getattr(model, "_packed", None)     # keeps the captured operand tensors alive; identity = weight version
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.et = model(xt, t)
        self.xt, self.t = xt, t

    def __call__(self):
        self.graph.replay()
        return self.et


def _graphed_forward(model, shape, device):
    """Per-model cache of captured forwards keyed by group shape; a new weight version (re-pack) drops the stale capture."""
    cache = model.__dict__.setdefault("_ddim_graphs", {})
    key = (tuple(shape), str(device))
    ent = cache.get(key)
    pack = model._packed_for(False)
    if ent is not None and ent.pack is pack:
        return ent
    cache.pop(key, None)
    while len(cache) >= 2:                       # each capture pins a forward's activations: keep the cache small
        cache.pop(next(iter(cache)))
    xt = torch.zeros(shape, dtype=torch.float32, device=device)
    t = torch.zeros(shape[0], dtype=torch.int64, device=device)
    ent = _GraphedForward(model, xt, t)
    cache[key] = ent
    return ent


def ddim_loop_device(xt, seq, model, alpha_list, eta=0.0, select_index=None, noise=None, on_snapshot=None, match_reference_rng=False,
                     graph=None):
    """The loop body of functions/denoising.py:21-50 on a DEVICE-resident fp32 ``xt`` (updated in place).

    ``on_snapshot(x0_buf, xs_buf)`` receives device copies of the two tensors the reference appends at :29-34 and
    :45-50 for the selected steps (``select_index`` None = every step; negative indices allowed); with
    ``on_snapshot=None`` no snapshot is taken.  ``graph``: None = replay a captured forward when the group is small
    (launch-bound), True / False force it.  Returns the tensor holding the final ``xt``."""
    device = xt.device
    seq = list(seq)
    seq_next = [-1] + seq[:-1]
    noise_it = iter(noise) if noise is not None else None
    L = len(seq)
    B, T = xt.size(0), xt.size(2)
    can_graph = hasattr(model, "_packed_for") and not (torch.is_grad_enabled() and any(p.requires_grad for p in model.parameters()))
    use_graph = can_graph and (graph if graph is not None else (B * T <= GRAPH_MAX_POSITIONS and L >= 2))
    fwd = None
    if use_graph:
        fwd = _graphed_forward(model, xt.shape, device)
        fwd.xt.copy_(xt)
        xt, t = fwd.xt, fwd.t
    else:
        t = torch.empty(B, dtype=torch.int64, device=device)
    for index, (i, j) in enumerate(zip(reversed(seq), reversed(seq_next))):
        t.fill_(int(i))
        at, at_next, c1, c2 = _ddim_scalars(alpha_list, i, j, eta)
        et = fwd() if fwd is not None else model(xt, t)
        selected = on_snapshot is not None and (select_index is None or index in select_index or index - L in select_index)
        z = None
        if noise_it is not None:
            z = next(noise_it).to(device=device, dtype=torch.float32).contiguous()
        elif c1 != 0 or match_reference_rng:
            z = torch.randn_like(xt)                                                  # :42 (the reference draws every step)
            if c1 == 0:
                z = None                                                              # drawn for the generator state, never added
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


def max_group_size(shape, device, reserve=0.85):
    """Largest number of samples of ``shape = [B, C, T, F]`` whose forward fits in the device memory that is free right now
    (free as the driver reports it plus what torch's allocator holds unused), from the level-0 footprint of a forward."""
    free, _ = torch.cuda.mem_get_info(device)
    free += torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device)
    per_sample = int(shape[2]) * int(shape[3]) * BYTES_PER_POSITION
    return max(1, int(free * reserve) // per_sample)


def _snapshot_count(L, select_index):
    if select_index is None:
        return L
    return sum(1 for k in range(L) if k in select_index or k - L in select_index)


def generalized_steps(x, seq, model, alpha, select_index, **kwargs):
    """functions/denoising.py:10-52.  Returns ``(xs, x0_preds)``: lists of CPU fp32 tensors, ``xs[0]`` being the input.

    Keywords beyond the reference's ``eta``: ``noise`` (iterable of per-step N(0,1) tensors), ``match_reference_rng``,
    ``chunks`` (number of sample groups; default: 4 for a host ``x`` with B >= 8 so the PCIe traffic of one group hides behind
    the denoiser of the next, else as many as device memory requires), ``graph`` (see ``ddim_loop_device``), ``out`` (the
    ``(xs, x0_preds)`` lists of an earlier call with the same shapes: their pinned tensors are overwritten and returned)."""
    if not torch.cuda.is_available():
        raise RuntimeError("ddim_audio_b200.generalized_steps needs a CUDA device: there is no CPU path")
    eta = kwargs.get("eta", 0)
    noise = kwargs.get("noise", None)
    match_rng = bool(kwargs.get("match_reference_rng", False))
    graph = kwargs.get("graph", None)
    out = kwargs.get("out", None)
    with torch.no_grad():
        alpha_list = [1.0] + alpha.to("cpu", torch.float32).numpy().tolist()          # :12
        device = x.device if x.is_cuda else torch.device("cuda", torch.cuda.current_device())
        B = x.size(0)
        seq = list(seq)
        nchunk = 1
        if B > 0:
            need = -(-B // max_group_size(x.shape, device))                           # groups device memory forces
            if match_rng:
                if need > 1:
                    raise RuntimeError(f"generalized_steps(match_reference_rng=True) runs the batch as ONE group like the reference, but {B} "
                                       f"samples of {tuple(x.shape[1:])} do not fit in device memory")
            else:
                want = kwargs.get("chunks", 4 if (not x.is_cuda and B >= 8 and eta == 0 and noise is None) else 1)
                nchunk = max(int(need), int(want))
        if noise is not None and nchunk > 1:
            noise = [n for n in noise]                                                # sliced per group below
        bounds = _chunk_bounds(B, nchunk) if B > 0 else []
        cur = torch.cuda.current_stream(device)
        pipe = _SnapshotPipe(device)
        # ---- inputs: host->device copies of all groups are queued now, on their own stream (bounded by the group size)
        staged = []
        if x.is_cuda:
            for c0, c1 in bounds:
                staged.append((None, None))
        else:
            h2d = torch.cuda.Stream(device=device)
            for c0, c1 in bounds:
                with torch.cuda.stream(h2d):
                    xc = x[c0:c1].to(device=device, dtype=torch.float32, non_blocking=True).contiguous()
                    ev = torch.cuda.Event()
                    ev.record(h2d)
                staged.append((xc, ev))
        # ---- snapshots of step k land in slices [c0:c1] of one full-batch pinned tensor per list entry
        nsnap = _snapshot_count(len(seq), select_index)
        shape = (B,) + tuple(x.shape[1:])
        out_xs, out_x0 = [], []
        if out is not None:
            rx, r0 = list(out[0][1:]), list(out[1])
            ok = len(rx) == nsnap and len(r0) == nsnap and all(v.shape == shape and v.dtype == torch.float32 and v.is_pinned() for v in rx + r0)
            if not ok:
                raise RuntimeError("generalized_steps(out=...): expected the pinned (xs, x0_preds) lists of a call with the same shapes")
            out_xs, out_x0 = rx, r0
        else:
            for _ in range(nsnap):
                out_x0.append(torch.empty(shape, dtype=torch.float32, pin_memory=True))
                out_xs.append(torch.empty(shape, dtype=torch.float32, pin_memory=True))
        for (c0, c1), (xt, ev) in zip(bounds, staged):
            if xt is None:
                xt = x[c0:c1].to(dtype=torch.float32, copy=True).contiguous()
            else:
                cur.wait_event(ev)
                xt.record_stream(cur)
            k = [0]

            def snap(x0_buf, xs_buf, c0=c0, c1=c1, k=k):
                pipe.push_into(x0_buf, out_x0[k[0]][c0:c1])                             # :29-34
                pipe.push_into(xs_buf, out_xs[k[0]][c0:c1])                             # :45-50
                k[0] += 1

            gnoise = noise if (noise is None or nchunk == 1) else [n[c0:c1] for n in noise]
            ddim_loop_device(xt, seq, model, alpha_list, eta, select_index, gnoise, snap if nsnap else None, match_rng, graph)
            del xt
        pipe.drain()
    return [x] + out_xs, out_x0
