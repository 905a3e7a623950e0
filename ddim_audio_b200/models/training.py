"""Training pass of the denoiser: ``Model.forward`` with gradients enabled runs ``DenoiserFn`` -- ONE autograd node over the
whole network whose forward saves bf16 activations + fp64 GroupNorm statistics and whose backward walks the network in
reverse entirely in libddim_b200 kernels (autograd of reference models/diffusion.py:237-294).

Gradients are written straight into views of one flat fp32 buffer (``model.flat_grad``) so data-parallel training reduces
them with a single collective.  Dropout (13 sites, models/diffusion.py:129,144 + hf modeling_fnet.py:226-231) uses the
library's counter-based mask, regenerated in backward from the saved seed.
"""
from __future__ import annotations

import math

import torch

from .. import ops

GN_EPS = 1e-6
_SEED_STRIDE = 0x9E3779B1


class _Saved:
    """Everything backward needs from one forward."""


def _flat_views(model, device):
    # every view starts on a 16-byte boundary (128-bit accesses in the optimizer / clip kernels); the <= 3 zero floats of
    # padding per parameter ride along in the data-parallel all-reduce
    params = model._params()
    names = model.__dict__["_param_names"]
    total = sum((p.numel() + 3) // 4 * 4 for p in params)
    flat = torch.zeros(total, dtype=torch.float32, device=device)
    views, off = {}, 0
    for name, p in zip(names, params):
        views[name] = flat[off:off + p.numel()].view(p.shape)
        off += (p.numel() + 3) // 4 * 4
    return flat, views


# ------------------------------------------------------------------------------------------------ forward
def _block_fwd(h, st_h, pb, temb, temb_off, temb_stride, stats, nxt):
    """models/diffusion.py:42-56 with materialised operands and PRE-activation conv outputs (their SiLU derivative is needed
    in backward): a0 = silu(GN0(h)); u0 = conv0(a0) + temb; n1 = GN1(silu(u0)); u1 = conv1(n1) + b; out = h + GN2(silu(u1))."""
    s0, s1, s2 = stats[nxt()], stats[nxt()], stats[nxt()]
    a0 = ops.gn_apply(h, st_h, pb["g0"], pb["b0"], 1, GN_EPS)
    u0 = ops.conv3x3(a0, pb["w0"], bias_b=temb[:, temb_off:], bias_b_stride=temb_stride, act=3, stats=s0, wimg=pb["w0_img"])
    n1 = ops.gn_apply(u0, s0, pb["g1"], pb["b1"], 3, GN_EPS)
    u1 = ops.conv3x3(n1, pb["w1"], bias=pb["cb1"], act=3, stats=s1, wimg=pb["w1_img"])
    out = ops.gn_apply(u1, s1, pb["g2"], None, 4, GN_EPS, res=h, stats_out=s2)
    return out, s2, dict(h=h, st_h=st_h, a0=a0, u0=u0, s0=s0, n1=n1, u1=u1, s1=s1, pb=pb, temb_off=temb_off)


def forward_train(model, x, t):
    from .diffusion import GN_GROUPS, _fourier_operands, sinusoid_table

    cfg = model.config
    B, Cin, T, F = x.shape
    nl = model.levels
    pk = model._packed_for(True)
    dev = x.device
    sv = _Saved()
    sv.x, sv.t, sv.pk = x, t, pk
    p_drop = model.transformer.embedding.dropout_p if model.training else 0.0
    sv.p_drop = p_drop
    # the step's seed: drawn on the host (a CPU draw: legal under stream capture, where its value is baked into the launches) plus, once a
    # captured step exists (GraphedLossBackward), a device word the kernels add -- that word is what changes between replays, the host
    # part is what changes between eagerly issued steps
    sv.seed_dev = model.__dict__.get("_seed_dev")
    sv.seed = int(torch.randint(0, 2 ** 62, (1,)).item()) if p_drop > 0 else 0

    # ---- timestep embedding, pre-activations kept (:107-120)
    lin = model.temb.weight
    sv.z1 = ops.linear_small(model.temb.te, lin[0].weight, lin[0].bias, act=0, idx=t)
    sv.z2 = ops.linear_small(sv.z1, lin[1].weight, lin[1].bias, act=2)
    temb = ops.linear_small(sv.z2, lin[2].weight, lin[2].bias, act=2)
    temb_stride = temb.shape[1]
    offs = [0]
    for w in model.embedding_size:
        offs.append(offs[-1] + w)
    rb_i = 0
    n_stats = 2 + 2 * (nl - 1) + 3 * len(model.embedding_size)
    stats = torch.zeros((n_stats, B, GN_GROUPS, 2), dtype=torch.float64, device=dev)
    counter = [0]

    def nxt():
        counter[0] += 1
        return counter[0] - 1

    # ---- down path
    conv0 = model.down_modules[0]
    st = stats[nxt()]
    h = ops.conv_in(x, conv0.weight, conv0.bias, st, wimg=pk["conv_in_img"])
    hidden = [h]
    sv.down = []
    for ent in pk["down"]:
        rec = dict(ent=ent, blocks=[], h_in=h)
        if ent["down"] is not None:
            st = stats[nxt()]
            h = ops.conv_down4x4(h, ent["down"]["w"], ent["down"]["b"], st, wimg=ent["down"]["w_img"])
        for pb in ent["blocks"]:
            h, st, blk = _block_fwd(h, st, pb, temb, offs[rb_i], temb_stride, stats, nxt)
            rec["blocks"].append(blk)
            rb_i += 1
        hidden.append(h)
        sv.down.append(rec)
    sv.hidden = list(hidden)

    # ---- FNet bottleneck
    tr = model.transformer
    S, Wt, Cl = h.shape[1], h.shape[2], h.shape[3]
    D, H = Cl * Wt, tr.hidden
    rows = B * S
    te = model._const(("te", S, D, dev), lambda: sinusoid_table(2 ** math.ceil(math.log2(S)) if S > 1 else 1, D)[:S].contiguous().to(dev))
    wh, a2, Sp = model._const(("dft", S, H, dev), lambda: _fourier_operands(S, H, dev))
    ut = model._const(("ut", B, S, H, dev), lambda: torch.zeros((B, 2 * H, 2 * Sp), dtype=torch.bfloat16, device=dev))
    sv.te, sv.dft, sv.dims = te, (wh, a2, ut), (B, S, Wt, Cl, H)
    sv.tok = ops.tokens_ln(h, te, tr.embedding.LayerNorm.weight, tr.embedding.LayerNorm.bias, tr.ln_eps)
    xr = ops.linear(sv.tok, pk["proj_w"], tr.embedding.projection.bias, out_f32=True)
    site = 0
    if p_drop > 0:
        xr = ops.dropout(xr, p_drop, sv.seed + site * _SEED_STRIDE, out=xr, seed_dev=sv.seed_dev)
    xbf = torch.empty((rows, 2 * H), dtype=torch.bfloat16, device=dev)
    ops.split_bf16(xr, xbf, True)
    sv.layers = []
    for lay, pw in zip(tr.encoder.layer, pk["fnet"]):
        site += 1
        y = ops.fourier_mix(xbf, 2 * H, wh, a2, ut, B, S, H, 3)
        ln1 = lay.fourier.output.LayerNorm
        hbf = torch.empty((rows, H), dtype=torch.bfloat16, device=dev)
        hr = ops.layernorm512(xr, y, ln1.weight, ln1.bias, tr.ln_eps, hbf, False)
        u = ops.linear(hbf, pw["w1"], lay.intermediate.dense.bias, act=0)            # pre-gelu, bf16
        gact = ops.gelu(u)
        v = ops.linear(gact, pw["w2"], lay.output.dense.bias, out_f32=True)
        if p_drop > 0:
            v = ops.dropout(v, p_drop, sv.seed + site * _SEED_STRIDE, out=v, seed_dev=sv.seed_dev)
        ln2 = lay.output.LayerNorm
        xbf = torch.empty((rows, 2 * H), dtype=torch.bfloat16, device=dev)
        xr_next = ops.layernorm512(v, hr, ln2.weight, ln2.bias, tr.ln_eps, xbf, True)
        sv.layers.append(dict(lay=lay, pw=pw, xr=xr, y=y, hr=hr, hbf=hbf, u=u, gact=gact, v=v, site=site))
        xr = xr_next
    sv.xbf_final = xbf
    tok_out = ops.linear(xbf, pk["out_w"], tr.compute_out.bias, out_f32=True, M=rows, K=H)
    st = stats[nxt()]
    h = ops.tokens_to_nhwc(tok_out, hidden.pop(), st)

    # ---- up path
    sv.up = []
    for ent in pk["up"]:
        rec = dict(ent=ent, blocks=[])
        for pb in ent["blocks"]:
            h, st, blk = _block_fwd(h, st, pb, temb, offs[rb_i], temb_stride, stats, nxt)
            rec["blocks"].append(blk)
            rb_i += 1
        if ent["up"] is not None:
            rec["h_in"] = h
            st = stats[nxt()]
            h = ops.conv_up4x4(h, ent["up"]["w"], ent["up"]["b"], skip=hidden.pop(), stats=st, wimg=ent["up"]["w_img"])
        sv.up.append(rec)
    conv_last = model.up_modules[-1]
    sv.h_last, sv.skip0 = h, hidden.pop()
    sv.stats = stats
    return ops.conv_out(h, sv.skip0, conv_last.weight, conv_last.bias, wimg=pk["conv_out_img"]), sv


# ------------------------------------------------------------------------------------------------ backward
def _block_bwd(blk, g, G, dtemb):
    """Backward of one residual block; ``g`` = dL/d(out) (bf16 NHWC).  Returns dL/d(h)."""
    pb = blk["pb"]
    pre = pb["prefix"]
    u1, u0, h = blk["u1"], blk["u0"], blk["h"]
    positions = h.shape[1] * h.shape[2]
    # out = h + GN2(silu(u1))
    R2 = ops.gn_bwd_reduce(g, u1, blk["s1"], pb["g2"], None, 1, GN_EPS)
    du1, sums1 = ops.gn_bwd_apply(g, u1, blk["s1"], R2, pb["g2"], None, 1, GN_EPS, sums=True)
    ops.gn_bwd_param(R2, blk["s1"], G[pre + ".norm.2.weight"], None, positions, GN_EPS)
    ops.sums_scatter(sums1, out_c=G[pre + ".conv.1.bias"])
    # u1 = conv1(n1) + b
    ops.conv3x3_wgrad(blk["n1"], du1, G[pre + ".conv.1.weight"])
    dn1 = ops.conv3x3(du1, pb["w1d"], wimg=pb["w1d_img"])
    # n1 = GN1(silu(u0))
    R1 = ops.gn_bwd_reduce(dn1, u0, blk["s0"], pb["g1"], pb["b1"], 1, GN_EPS)
    du0, sums0 = ops.gn_bwd_apply(dn1, u0, blk["s0"], R1, pb["g1"], pb["b1"], 1, GN_EPS, sums=True, out=dn1)
    ops.gn_bwd_param(R1, blk["s0"], G[pre + ".norm.1.weight"], G[pre + ".norm.1.bias"], positions, GN_EPS)
    ops.sums_scatter(sums0, out_b=dtemb[:, blk["temb_off"]:], ldb=dtemb.shape[1])
    # u0 = conv0(a0) + temb
    ops.conv3x3_wgrad(blk["a0"], du0, G[pre + ".conv.0.weight"])
    da0 = ops.conv3x3(du0, pb["w0d"], out=du1, wimg=pb["w0d_img"])
    # a0 = silu(GN0(h)); + residual path
    R0 = ops.gn_bwd_reduce(da0, h, blk["st_h"], pb["g0"], pb["b0"], 0, GN_EPS)
    dh, _ = ops.gn_bwd_apply(da0, h, blk["st_h"], R0, pb["g0"], pb["b0"], 0, GN_EPS, add=g, out=da0)
    ops.gn_bwd_param(R0, blk["st_h"], G[pre + ".norm.0.weight"], G[pre + ".norm.0.bias"], positions, GN_EPS)
    return dh


def backward_train(model, sv, gout):
    """``gout`` = dL/d(output) fp32 [B,2,T,F].  Returns (flat gradient buffer, {parameter name: view})."""
    pk = sv.pk
    dev = gout.device
    flat, G = _flat_views(model, dev)
    B, S, Wt, Cl, H = sv.dims
    rows = B * S
    tr = model.transformer
    dtemb = torch.zeros((B, sum(model.embedding_size)), dtype=torch.float32, device=dev)
    nup = len(model.up_modules) - 1

    # ---- conv_out (:199-208,284)
    ops.conv_out_wgrad(sv.h_last, sv.skip0, gout, G[f"up_modules.{nup}.weight"], G[f"up_modules.{nup}.bias"])
    zero_b = model._const(("zero32", dev), lambda: torch.zeros(32, dtype=torch.float32, device=dev))
    g = ops.conv_in(gout, pk["conv_out_wd"], zero_b, None, wimg=pk["conv_out_wd_img"])          # d(h + skip0), bf16 NHWC
    skipgrad = {0: g}

    # ---- up path, reversed
    k = 1                                                            # index of the hidden tensor the next Upsample added
    for rec in reversed(sv.up):
        ent = rec["ent"]
        if ent["up"] is not None:
            pre = ent["up"]["prefix"]
            skipgrad[k] = g                                          # out = convT(h_in) + hidden[k]
            k += 1
            ops.conv_up4x4_wgrad(rec["h_in"], g, G[pre + ".conv.weight"])
            ops.colsum(g, G[pre + ".conv.bias"])
            g = ops.conv_down4x4(g, ent["up"]["wd"], None, wimg=ent["up"]["wd_img"])
        for blk in reversed(rec["blocks"]):
            g = _block_bwd(blk, g, G, dtemb)

    # ---- FNet bottleneck, reversed (:267-279)
    skipgrad[k] = g                                                  # h = tokens(tok_out) + hidden[-1]
    g_tok = ops.nhwc_to_tokens(g)
    ops.colsum(g_tok, G["transformer.compute_out.bias"])
    ops.linear_wgrad(sv.xbf_final, g_tok, G["transformer.compute_out.weight"], K=H)
    dy1 = ops.linear(g_tok, pk["out_wt"], None, out_f32=True)
    dy2 = None
    wh, a2, ut = sv.dft
    for L, rec in reversed(list(enumerate(sv.layers))):
        lay, pw = rec["lay"], rec["pw"]
        pre = f"transformer.encoder.layer.{L}"
        ln2, ln1 = lay.output.LayerNorm, lay.fourier.output.LayerNorm
        dvbf = torch.empty((rows, H), dtype=torch.bfloat16, device=dev)
        ds2 = ops.layernorm512_bwd(rec["v"], rec["hr"], dy1, dy2, ln2.weight, G[pre + ".output.LayerNorm.weight"],
                                   G[pre + ".output.LayerNorm.bias"], tr.ln_eps, dsbf=dvbf, write_lo=False)
        dv = ds2
        if sv.p_drop > 0:
            dv = ops.dropout(ds2, sv.p_drop, sv.seed + rec["site"] * _SEED_STRIDE, seed_dev=sv.seed_dev)
            ops.add_split_bf16(dv, None, dvbf, False)
        ops.colsum(dv, G[pre + ".output.dense.bias"])
        ops.linear_wgrad(rec["gact"], dvbf, G[pre + ".output.dense.weight"])
        dgact = ops.linear(dvbf, pw["w2t"], None)
        du = ops.gelu(rec["u"], dgact)
        ops.colsum(du, G[pre + ".intermediate.dense.bias"])
        ops.linear_wgrad(rec["hbf"], du, G[pre + ".intermediate.dense.weight"])
        dhr = ops.linear(du, pw["w1t"], None, out_f32=True)
        ds1bf = torch.empty((rows, 2 * H), dtype=torch.bfloat16, device=dev)
        ds1 = ops.layernorm512_bwd(rec["xr"], rec["y"], dhr, ds2, ln1.weight, G[pre + ".fourier.output.LayerNorm.weight"],
                                   G[pre + ".fourier.output.LayerNorm.bias"], tr.ln_eps, dsbf=ds1bf, write_lo=True)
        dy1, dy2 = ds1, ops.fourier_mix(ds1bf, 2 * H, wh, a2, ut, B, S, H, 3)      # the DFT matrices are symmetric
    dxbf = torch.empty((rows, H), dtype=torch.bfloat16, device=dev)
    dxr = torch.empty_like(dy1)
    ops.add_split_bf16(dy1, dy2, dxbf, False, sum_out=dxr)
    if sv.p_drop > 0:
        dxr = ops.dropout(dxr, sv.p_drop, sv.seed, out=dxr, seed_dev=sv.seed_dev)
        ops.add_split_bf16(dxr, None, dxbf, False)
    ops.colsum(dxr, G["transformer.embedding.projection.bias"])
    ops.linear_wgrad(sv.tok, dxbf, G["transformer.embedding.projection.weight"])
    dtok = ops.linear(dxbf, pk["proj_wt"], None, out_f32=True)
    emb_ln = tr.embedding.LayerNorm
    g = ops.tokens_ln_bwd(sv.hidden[-1], sv.te, emb_ln.weight, dtok, skipgrad[k], G["transformer.embedding.LayerNorm.weight"],
                          G["transformer.embedding.LayerNorm.bias"], tr.ln_eps)

    # ---- down path, reversed (:254-264)
    for rec in reversed(sv.down):
        ent = rec["ent"]
        for blk in reversed(rec["blocks"]):
            g = _block_bwd(blk, g, G, dtemb)
        if ent["down"] is not None:
            pre = ent["down"]["prefix"]
            k -= 1
            ops.conv_down4x4_wgrad(rec["h_in"], g, G[pre + ".conv.weight"])
            ops.colsum(g, G[pre + ".conv.bias"])
            g = ops.conv_up4x4(g, ent["down"]["wd"], None, skip=skipgrad[k], wimg=ent["down"]["wd_img"])
    assert k == 1
    ops.conv_in_wgrad(sv.x, g, skipgrad[0], G["down_modules.0.weight"], G["down_modules.0.bias"])

    # ---- timestep embedding MLP (:107-120)
    lin = model.temb.weight
    ops.linear_small_wgrad(dtemb, sv.z2, G["temb.weight.2.weight"], G["temb.weight.2.bias"], True)
    dz2 = ops.linear_small_dgrad(dtemb, lin[2].weight, sv.z2)
    ops.linear_small_wgrad(dz2, sv.z1, G["temb.weight.1.weight"], G["temb.weight.1.bias"], True)
    dz1 = ops.linear_small_dgrad(dz2, lin[1].weight, sv.z1)
    ops.linear_small_wgrad(dz1, model.temb.te, G["temb.weight.0.weight"], G["temb.weight.0.bias"], False, idx=sv.t)
    return flat, G


class DenoiserFn(torch.autograd.Function):
    """forward(model, x, t, *parameters) -> eps; the parameters are passed so autograd routes their gradients here."""

    @staticmethod
    def forward(ctx, model, x, t, *params):
        out, sv = forward_train(model, x, t)
        ctx.model, ctx.sv = model, sv
        return out

    @staticmethod
    def backward(ctx, gout):
        model, sv = ctx.model, ctx.sv
        ctx.sv = None
        flat, G = backward_train(model, sv, gout.contiguous().float())
        model.flat_grad = flat
        model._params()
        grads = tuple(G[name] for name in model.__dict__["_param_names"])
        del G
        return (None, None, None) + grads


# ------------------------------------------------------------------------------------------------ captured loss + backward
def loss_and_grads(model, x0, t, e, a, world=1):
    """noise_estimation_loss forward + backward WITHOUT the autograd engine (functions/losses.py:4-18 and its derivative): returns
    (loss, flat gradient buffer, {parameter name: view into it}) of loss / world.  The same library calls as the autograd path
    (functions/losses.py, DenoiserFn) in the same order."""
    x0 = x0.contiguous().float()
    e = e.contiguous().float()
    t = t.long().contiguous()
    x = ops.qsample(x0, e, a, t)                                                      # :12-13
    out, sv = forward_train(model, x, t)                                              # :14
    out = out.contiguous()
    per = ops.sqerr_loss(e, out).to(torch.float32)                                    # :15-18
    loss = per.mean(dim=0)
    g = torch.full((per.numel(),), 1.0 / (per.numel() * world), dtype=torch.float32, device=x0.device)
    gout = ops.sqerr_loss_bwd(e, out, g)
    flat, G = backward_train(model, sv, gout.contiguous().float())
    return loss, flat, G


class GraphedLossBackward:
    """``loss = noise_estimation_loss(model, x, t, e, alphas); (loss / world).backward()`` for ONE batch shape, captured in a CUDA graph
    and replayed per step.  A train step at x[32,2,1024,256] is ~1100 launches of 8-400 us: issued from Python the short ones (FNet, the
    deep levels, the GroupNorm parameter sums) leave the GPU waiting for the host; a replay does not.

    What the capture contains: the re-pack of the bf16 operand copies from the fp32 masters (``_RepackPlan.refresh``: the optimizer
    changes the masters between replays), forward diffusion, the whole forward and backward (``loss_and_grads``: the library calls of
    the autograd path without the autograd engine, whose AccumulateGrad nodes are tied to the stream they were created on), the zeroing
    of the flat gradient buffer.  What stays outside: drawing t / e (copied into static input buffers), the dropout seed (a device word
    the kernels read: ``ddimb_dropout`` seed_dev), the data-parallel all-reduce, clipping, optimizers, EMA.  After the capture every
    ``p.grad`` is a static view into ``model.flat_grad``: callers must not ``zero_grad(set_to_none=True)`` (``keep_grads`` puts the
    tensors back if they did)."""

    def __init__(self, model, alphas, x, t, e, world=1):
        dev = x.device
        self.model, self.world = model, world
        self.params = model._params()
        self.a = alphas.to(device=dev, dtype=torch.float32).contiguous()
        self.x, self.t, self.e = x.detach().float().clone(), t.detach().long().clone(), e.detach().float().clone()
        self.seed = torch.zeros(1, dtype=torch.int64, device=dev)
        model.__dict__["_seed_dev"] = self.seed
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):                  # builds the re-pack plan, sets the kernels' shared-memory attributes, sizes the allocator pools
                loss_and_grads(model, self.x, self.t, self.e, self.a, world)
        cur.wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph), torch.no_grad():
            model.mark_params_dirty()           # the re-pack launches belong to every replay
            self.loss, self.flat_grad, G = loss_and_grads(model, self.x, self.t, self.e, self.a, world)
        model._params()
        self.grads = [G[name] for name in model.__dict__["_param_names"]]
        self.keep_grads()

    def keep_grads(self):
        for p, g in zip(self.params, self.grads):
            if p.grad is not g:
                p.grad = g
        self.model.flat_grad = self.flat_grad

    def __call__(self, x, t, e):
        self.x.copy_(x, non_blocking=True)
        self.t.copy_(t, non_blocking=True)
        self.e.copy_(e, non_blocking=True)
        self.seed.random_(0, 2 ** 62)
        self.keep_grads()
        self.graph.replay()
        return self.loss
