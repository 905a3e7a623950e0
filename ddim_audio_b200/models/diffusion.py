"""B200-native denoiser behind the reference's ``Model(config).forward(input, t)`` (models/diffusion.py:170-294).

The module tree exists to own the parameters under the reference's 389 ``state_dict`` keys (and to consume the
RNG in the reference's construction order, so ``torch.manual_seed(s); Model(config)`` initialises identically);
no ``torch.nn`` forward is ever called.  ``forward`` drives the sm_100a kernels of libddim_b200 through
``ddim_audio_b200.ops``: activations live in HBM as NHWC bf16, GroupNorm statistics are produced by the
kernel that writes a tensor, convolutions / Linears / the FNet Fourier mixing are tcgen05 GEMMs.
There is no CPU path: a non-CUDA input raises.
"""
from __future__ import annotations

import math
from typing import List

import torch
import torch.nn as nn

from .. import _C, ops

GN_GROUPS = 8        # models/diffusion.py:19
GN_EPS = 1e-6        # models/diffusion.py:20


def _group_norm(channels: int) -> nn.GroupNorm:
    return nn.GroupNorm(GN_GROUPS, channels, eps=GN_EPS, affine=True)


class Residual_Block(nn.Module):
    """Parameter holder for models/diffusion.py:12-56 (3 GroupNorms, conv0 without bias, conv1 with bias;
    norm[2] has zero-initialised weight and no bias)."""

    def __init__(self, *, channels, kernel_size=3):
        super().__init__()
        if kernel_size != 3:
            raise NotImplementedError("ddim_audio_b200: Residual_Block kernels exist for kernel_size=3 only (audio.yml:49)")
        self.channels = channels
        self.norm = nn.Sequential(_group_norm(channels), _group_norm(channels), _group_norm(channels))
        nn.init.zeros_(self.norm[2].weight)
        self.norm[2].register_parameter("bias", None)
        self.conv = nn.Sequential(
            nn.Conv2d(channels, channels, kernel_size, 1, kernel_size // 2, bias=False),
            nn.Conv2d(channels, channels, kernel_size, 1, kernel_size // 2, bias=True),
        )


class Upsample(nn.Module):
    """models/diffusion.py:59-67."""

    def __init__(self, *, in_channels, out_channels):
        super().__init__()
        self.conv = nn.ConvTranspose2d(in_channels, out_channels, kernel_size=4, stride=2, padding=1)


class Downsample(nn.Module):
    """models/diffusion.py:70-78."""

    def __init__(self, *, in_channels, out_channels):
        super().__init__()
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size=4, stride=2, padding=1)


def sinusoid_table(length: int, channel: int) -> torch.Tensor:
    """models/diffusion.py:81-92 applied to a zero table: sin on even columns, cos on odd ones (fp32, host)."""
    pos = torch.arange(length, dtype=torch.float32).unsqueeze(1)
    freq = torch.exp(torch.arange(0, channel, 2, dtype=torch.float32) * (-math.log(10000.0) / channel))
    ang = pos * freq
    table = torch.zeros(length, channel, dtype=torch.float32)
    table[:, 0::2] = torch.sin(ang)
    table[:, 1::2] = torch.cos(ang)
    return table


class BetaEmbedding(nn.Module):
    """models/diffusion.py:95-120: sinusoid table buffer ``te`` + three Linears named ``weight.{0,1,2}``."""

    def __init__(self, seq_length, channel_sz):
        super().__init__()
        pos_ch, emb_ch = 128, 512
        self.register_buffer("te", sinusoid_table(seq_length, pos_ch))
        self.weight = nn.Sequential(nn.Linear(pos_ch, emb_ch), nn.Linear(emb_ch, emb_ch), nn.Linear(emb_ch, channel_sz))


class TransformerEmbedding(nn.Module):
    """models/diffusion.py:123-145."""

    def __init__(self, in_channels, config):
        super().__init__()
        self.LayerNorm = nn.LayerNorm(in_channels, eps=config.kwargs.layer_norm_eps)
        self.projection = nn.Linear(in_channels, config.channels)
        self.dropout_p = float(config.kwargs.hidden_dropout_prob)


class _Node(nn.Module):
    """Bare container used to spell the HF FNetEncoder key names (hf modeling_fnet.py:180-286)."""


def _fnet_layer(hidden: int, ffn: int, eps: float) -> nn.Module:
    lay = _Node()
    lay.fourier = _Node()
    lay.fourier.output = _Node()
    lay.fourier.output.LayerNorm = nn.LayerNorm(hidden, eps=eps)
    lay.intermediate = _Node()
    lay.intermediate.dense = nn.Linear(hidden, ffn)
    lay.output = _Node()
    lay.output.dense = nn.Linear(ffn, hidden)
    lay.output.LayerNorm = nn.LayerNorm(hidden, eps=eps)
    return lay


class Transformer_Module(nn.Module):
    """models/diffusion.py:148-167.  The reference ``exec``/``eval``s config strings to build HF's FNetEncoder; the
    only arithmetic those strings can select here is FNet with gelu_new, so they are checked, not executed."""

    def __init__(self, io_channels, config):
        super().__init__()
        kw = config.kwargs
        if "FNetEncoder" not in str(getattr(config, "module", "FNetEncoder")):
            raise NotImplementedError("ddim_audio_b200: transformers.module must be FNetEncoder (audio.yml:31)")
        if getattr(kw, "hidden_act", "gelu_new") != "gelu_new":
            raise NotImplementedError("ddim_audio_b200: FNet hidden_act must be gelu_new (audio.yml:37)")
        if kw.hidden_size != 512 or config.channels != kw.hidden_size:
            raise NotImplementedError("ddim_audio_b200: FNet kernels are built for hidden_size == channels == 512 (audio.yml:34,41)")
        self.embedding = TransformerEmbedding(io_channels, config)
        self.encoder = _Node()
        self.encoder.layer = nn.ModuleList(
            [_fnet_layer(kw.hidden_size, kw.intermediate_size, kw.layer_norm_eps) for _ in range(kw.num_hidden_layers)]
        )
        self.compute_out = nn.Linear(config.channels, io_channels)
        self.ln_eps = float(kw.layer_norm_eps)
        self.hidden = int(kw.hidden_size)


def _fourier_operands(S: int, H: int, device) -> tuple:
    """DFT matrices for ``ddimb_fourier_mix`` (layouts documented in csrc/gemm_host.cu), built in float64 on the
    host and split into bf16 hi / lo planes: Re(F_S x F_H) = C_S (x C_H) - S_S (x S_H)."""
    Sp = (S + 31) // 32 * 32
    kh = torch.arange(H, dtype=torch.float64)
    ang_h = 2.0 * math.pi * torch.outer(kh, kh) / H
    wh = torch.cat([torch.cos(ang_h), torch.sin(ang_h)], dim=0)                     # [2H, H]
    wh_hi = wh.to(torch.bfloat16)
    wh_lo = (wh - wh_hi.double()).to(torch.bfloat16)
    wh_op = torch.cat([wh_hi, wh_hi, wh_lo], dim=1).contiguous()                    # matched with x [hi | lo | hi]
    ks = torch.arange(S, dtype=torch.float64)
    ang_s = 2.0 * math.pi * torch.outer(ks, ks) / S
    blocks = []
    for m in (torch.cos(ang_s), -torch.sin(ang_s)):
        pad = torch.zeros(S, Sp, dtype=torch.float64)
        pad[:, :S] = m
        blocks.append(pad)
    hi = [b.to(torch.bfloat16) for b in blocks]
    lo = [(b - h.double()).to(torch.bfloat16) for b, h in zip(blocks, hi)]
    a2 = torch.cat(hi + lo, dim=1).contiguous()                                     # [S, 4*Sp]
    return wh_op.to(device), a2.to(device), Sp


def _gather_torch(w: torch.Tensor, dims, strides, off: int) -> torch.Tensor:
    """dst[i0][i1][i2][i3] = w.flat[off + sum_k i_k strides_k] (negative strides = flips): the torch statement of one operand layout."""
    idx = torch.full((1, 1, 1, 1), int(off), dtype=torch.int64, device=w.device)
    for k, (d, s_) in enumerate(zip(dims, strides)):
        shape = [1, 1, 1, 1]
        shape[k] = int(d)
        idx = idx + (torch.arange(int(d), dtype=torch.int64, device=w.device) * int(s_)).reshape(shape)
    return w.reshape(-1)[idx.reshape(-1)].reshape([int(d) for d in dims])


class _RepackPlan:
    """Everything ``Model._pack(train=True)`` derived from the fp32 masters, as data: strided-gather records for ONE ``ddimb_repack``
    launch (include/ddim_b200.h) and the argument lists of the weight-image calls.  ``refresh`` rewrites all copies in place."""

    CHUNK = 16384

    def __init__(self):
        self.records, self.images = [], []
        self.table, self.nchunks, self.total, self.param_ptrs = None, 0, 0, ()

    def add(self, w, dst, dims, strides, off, f32):
        self.records.append((w, dst, tuple(int(d) for d in dims), tuple(int(x) for x in strides), int(off), bool(f32)))

    def add_image(self, kind, src, img, Cin, Cout, W):
        self.images.append((int(kind), src, img, int(Cin), int(Cout), int(W)))

    def finish(self, model):
        import numpy as np

        dt = np.dtype([("src", "<u8"), ("dst", "<u8"), ("start", "<i8"), ("n", "<i8"), ("d1", "<i4"), ("d2", "<i4"), ("d3", "<i4"),
                       ("f32", "<i4"), ("s0", "<i8"), ("s1", "<i8"), ("s2", "<i8"), ("s3", "<i8")])
        assert dt.itemsize == 80
        rows = []
        for w, dst, dims, strides, off, f32 in self.records:
            n = dst.numel()
            for start in range(0, n, self.CHUNK):
                rows.append((w.data_ptr() + 4 * off, dst.data_ptr(), start, min(self.CHUNK, n - start), dims[1], dims[2], dims[3], int(f32),
                             strides[0], strides[1], strides[2], strides[3]))
            self.total += n
        arr = np.array(rows, dtype=dt)
        dev = self.records[0][1].device
        self.table = torch.from_numpy(arr.view(np.uint8).reshape(-1).copy()).to(dev)
        self.nchunks = len(rows)
        self.param_ptrs = tuple(p.data_ptr() for p in model._params())
        idt = np.dtype([("kind", "<i4"), ("Cin", "<i4"), ("Cout", "<i4"), ("W", "<i4"), ("w", "<u8"), ("img", "<u8")])
        assert idt.itemsize == 32
        self._img_recs = np.array([(k, ci, co, W, _C.ptr(src), _C.ptr(img)) for k, src, img, ci, co, W in self.images], dtype=idt)

    def matches(self, model) -> bool:
        return self.table is not None and self.param_ptrs == tuple(p.data_ptr() for p in model._params())

    def refresh(self):
        _C.call("ddimb_repack", _C.ptr(self.table), self.nchunks, self.total)
        if len(self._img_recs):
            _C.call("ddimb_conv_wimg_batch", self._img_recs.ctypes.data, len(self._img_recs))


class Model(nn.Module):
    """Drop-in for the reference ``Model`` (models/diffusion.py:170-294): same constructor argument (the full config
    namespace), same parameter names and shapes, same ``forward(input[B,C,T,F], t[B]) -> [B,C,T,F]`` fp32."""

    def __init__(self, config):
        self.config = config.model
        cfg = self.config
        assert len(cfg.ch) == len(cfg.krn) == len(cfg.res)
        super().__init__()
        widths = [c for n, c in zip(cfg.res, cfg.ch) for _ in range(n)]
        self.embedding_size = widths + widths[::-1]                                 # :177-183
        self.temb = BetaEmbedding(config.diffusion.num_diffusion_timesteps, sum(self.embedding_size))
        self.down_modules = nn.ModuleList([nn.Conv2d(cfg.channels, cfg.ch[0], 3, 1, 1)])
        ups: List[nn.Module] = [nn.Conv2d(cfg.ch[0], cfg.channels, 3, 1, 1)]
        prev = None
        for c, k, n in zip(cfg.ch, cfg.krn, cfg.res):
            # construction order (down stage, then up stage, per level) fixes the RNG stream of the default init
            stage = [] if prev is None else [Downsample(in_channels=prev, out_channels=c)]
            stage += [Residual_Block(channels=c, kernel_size=k) for _ in range(n)]
            self.down_modules.append(nn.ModuleList(stage))
            stage = [] if prev is None else [Upsample(in_channels=c, out_channels=prev)]
            stage += [Residual_Block(channels=c, kernel_size=k) for _ in range(n)]
            ups.append(nn.ModuleList(stage[::-1]))
            prev = c
        self.up_modules = nn.ModuleList(ups[::-1])                                  # index 0 = deepest level (:226-227)
        self.levels = len(cfg.ch)
        self.token_f = cfg.f_size // (2 ** (self.levels - 1))
        self.transformer = Transformer_Module(cfg.ch[-1] * self.token_f, cfg.transformers)
        if cfg.dtype:
            self._apply_type_string(cfg.dtype)
        self._packed = None
        self._packed_key = None
        self._gen = getattr(self, "_gen", 0) + 1
        self._consts = {}
        self.flat_grad = None       # flat fp32 buffer the last backward wrote every parameter gradient into

    # ------------------------------------------------------------------ dtype strings (audio.yml:26,42)
    def _apply_type_string(self, s: str):
        if not str(s).endswith("FloatTensor"):
            raise NotImplementedError(f"ddim_audio_b200: parameters are fp32 masters; unsupported model.dtype {s!r}")
        self.type(s)            # 'torch.cuda.FloatTensor' also moves the module to the GPU, as in the reference

    # ------------------------------------------------------------------ kernel-ready weights
    def mark_params_dirty(self):
        """Call after writing parameters in a way torch cannot see (``p.data.copy_``, raw-pointer kernels): the packed bf16
        operand copies are rebuilt on the next forward.  ``load_state_dict``, ``train()`` / ``eval()``, ``.to()`` / ``.type()``,
        the fused optimizers' ``step()`` and ``EMAHelper.ema()`` do this themselves."""
        self._gen = getattr(self, "_gen", 0) + 1

    def load_state_dict(self, *args, **kwargs):
        self.mark_params_dirty()
        return super().load_state_dict(*args, **kwargs)

    def train(self, mode=True):
        # the reference's sample() is load_state_dict -> ema_helper.ema(model) -> model.eval() (runners/diffusion.py:303-315):
        # a mode switch is where a foreign ``param.data.copy_`` (its own EMAHelper) becomes visible at the latest
        self.mark_params_dirty()
        return super().train(mode)

    def _apply(self, fn, *args, **kwargs):
        self.mark_params_dirty()
        r = super()._apply(fn, *args, **kwargs)
        self.__dict__.pop("_param_list", None)       # .to() / .type() may replace the Parameter objects
        return r

    def _params(self):
        """The module's parameters as a cached list: ``nn.Module.parameters()`` walks the module tree (~3 ms per call for this network)
        and the train step needs the list several times per step.  Parameter OBJECTS only change through module surgery / ``_apply``,
        which drops the cache."""
        pl = self.__dict__.get("_param_list")
        if pl is None:
            pl = self.__dict__["_param_list"] = list(self.parameters())
            self.__dict__["_param_names"] = [n for n, _ in self.named_parameters()]
        return pl

    def _pack_key(self, train):
        """Sampling: two counters (no walk over the 389 parameters per forward).  Training adds the (pointer, version) walk so a
        stock torch optimizer's in-place update is seen too (100 us next to an 85 ms step)."""
        base = (_C.param_epoch[0], self._gen)
        return (base, tuple((p.data_ptr(), p._version) for p in self._params()) if train else None)

    @torch.no_grad()
    def _pack(self, train=False):
        """bf16 K-major operand copies of every GEMM weight (layouts: include/ddim_b200.h).  ``train`` adds the operands of
        the data-gradient GEMMs: 3x3 kernels flipped + transposed, Linear weights transposed; a Downsample's data gradient
        runs the ConvTranspose kernel on the SAME weight tensor and vice versa."""
        bf = torch.bfloat16
        pk = {"train": train}
        F0, nl = self.config.f_size, self.levels

        # Every operand copy is a strided gather of its fp32 master (negative strides = flips).  The first train pack builds the copies
        # with torch AND records (master, copy, dims, strides) plus every weight-image call; later packs of the same parameters rewrite
        # the copies in place with one library launch (``_refresh``): ~600 torch calls were 25 ms of host time per train step.
        plan = _RepackPlan() if train else None

        def gather(w, dims, strides, off=0, out_shape=None, f32=False):
            w = w.detach()
            dst = _gather_torch(w, dims, strides, off).to(torch.float32 if f32 else bf).contiguous()
            if out_shape is not None:
                dst = dst.reshape(out_shape)
            if plan is not None:
                plan.add(w, dst, dims, strides, off, f32)
            return dst

        def c3(w):      # [Co, Ci, 3, 3] -> [Co, (ky*3+kx)*Ci + ci]
            Co, Ci = w.shape[:2]
            return gather(w, (Co, 3, 3, Ci), (Ci * 9, 3, 1, 9), 0, (Co, 9 * Ci))

        def c3d(w):     # data-gradient operand: [Ci, (ky*3+kx)*Co + co] of the flipped kernel
            Co, Ci = w.shape[:2]
            return gather(w, (Ci, 3, 3, Co), (9, -3, -1, Ci * 9), 8, (Ci, 9 * Co))

        def as_conv(w):   # [O, I, 4, 4] -> [O, (ky*4+kx)*I + i]   (ddimb_conv_down4x4 operand)
            O, I = w.shape[:2]
            return gather(w, (O, 4, 4, I), (I * 16, 4, 1, 16), 0, (O, 16 * I))

        def as_convT(w):  # [I, O, 4, 4] -> [(ky*4+kx)*O + o, i]    (ddimb_conv_up4x4 operand)
            I, O = w.shape[:2]
            return gather(w, (4, 4, O, I), (4, 1, 16, O * 16), 0, (16 * O, I))

        def lin(w):       # [N, K] -> bf16 copy
            N, K = w.shape
            return gather(w, (1, 1, N, K), (0, 0, K, 1), 0, (N, K))

        def lin_t(w):     # [N, K] -> [K, N]
            N, K = w.shape
            return gather(w, (1, 1, K, N), (0, 0, 1, K), 0, (K, N))

        def img(kind, w, Cin, Cout, W):
            im = ops.conv_wimg(kind, w, Cin, Cout, W)
            if plan is not None and im is not None:
                plan.add_image(kind, w, im, Cin, Cout, W)
            return im

        def rb(m, prefix, W):
            C = m.channels
            d = dict(g0=m.norm[0].weight, b0=m.norm[0].bias, g1=m.norm[1].weight, b1=m.norm[1].bias, g2=m.norm[2].weight,
                     w0=c3(m.conv[0].weight), w1=c3(m.conv[1].weight), cb1=m.conv[1].bias, C=C, prefix=prefix)
            d["w0_img"], d["w1_img"] = img(ops.KIND_CONV3X3, d["w0"], C, C, W), img(ops.KIND_CONV3X3, d["w1"], C, C, W)
            if train:
                d["w0d"], d["w1d"] = c3d(m.conv[0].weight), c3d(m.conv[1].weight)
                d["w0d_img"], d["w1d_img"] = img(ops.KIND_CONV3X3, d["w0d"], C, C, W), img(ops.KIND_CONV3X3, d["w1d"], C, C, W)
            return d

        down = []
        for si, stage in enumerate(list(self.down_modules)[1:], start=1):
            ent = dict(down=None, blocks=[])
            W = F0 >> (si - 1)                                                       # width of this level
            for mj, m in enumerate(stage):
                prefix = f"down_modules.{si}.{mj}"
                if isinstance(m, Downsample):
                    co, ci = m.conv.weight.shape[:2]
                    ent["down"] = dict(w=as_conv(m.conv.weight), b=m.conv.bias, prefix=prefix)
                    ent["down"]["w_img"] = img(ops.KIND_DOWN, ent["down"]["w"], ci, co, 2 * W)
                    if train:
                        ent["down"]["wd"] = as_convT(m.conv.weight)                  # data gradient: ConvTranspose co -> ci on the level's width
                        ent["down"]["wd_img"] = img(ops.KIND_UP, ent["down"]["wd"], co, ci, W)
                else:
                    ent["blocks"].append(rb(m, prefix, W))
            down.append(ent)
        up = []
        for si, stage in enumerate(list(self.up_modules)[:-1]):
            ent = dict(up=None, blocks=[])
            W = F0 >> (nl - 1 - si)
            for mj, m in enumerate(stage):
                prefix = f"up_modules.{si}.{mj}"
                if isinstance(m, Upsample):
                    ci, co = m.conv.weight.shape[:2]
                    ent["up"] = dict(w=as_convT(m.conv.weight), b=m.conv.bias, prefix=prefix)
                    ent["up"]["w_img"] = img(ops.KIND_UP, ent["up"]["w"], ci, co, W)
                    if train:
                        ent["up"]["wd"] = as_conv(m.conv.weight)                     # data gradient: Conv2d 4x4 s2 co -> ci on the doubled width
                        ent["up"]["wd_img"] = img(ops.KIND_DOWN, ent["up"]["wd"], co, ci, 2 * W)
                else:
                    ent["blocks"].append(rb(m, prefix, W))
            up.append(ent)
        pk["down"], pk["up"] = down, up
        cin, cout = self.down_modules[0], self.up_modules[-1]
        pk["conv_in_img"] = img(ops.KIND_CONV_IN, cin.weight.detach(), cin.weight.shape[1], cin.weight.shape[0], F0)
        pk["conv_out_img"] = img(ops.KIND_CONV_OUT, cout.weight.detach(), cout.weight.shape[1], cout.weight.shape[0], F0)
        tr = self.transformer
        pk["proj_w"] = lin(tr.embedding.projection.weight)
        pk["out_w"] = lin(tr.compute_out.weight)
        pk["fnet"] = [dict(w1=lin(l.intermediate.dense.weight), w2=lin(l.output.dense.weight)) for l in tr.encoder.layer]
        if train:
            pk["proj_wt"] = lin_t(tr.embedding.projection.weight)
            pk["out_wt"] = lin_t(tr.compute_out.weight)
            for d, l in zip(pk["fnet"], tr.encoder.layer):
                d["w1t"] = lin_t(l.intermediate.dense.weight)
                d["w2t"] = lin_t(l.output.dense.weight)
            w = self.up_modules[-1].weight                                           # conv_out [2, 32, 3, 3]
            # conv_in-shaped data-gradient kernel: [ci, co, ky, kx] = w[co, ci, 2 - ky, 2 - kx], kept in fp32
            pk["conv_out_wd"] = gather(w, (32, 2, 3, 3), (9, 288, -3, -1), 8, (32, 2, 3, 3), f32=True)
            pk["conv_out_wd_img"] = img(ops.KIND_CONV_IN, pk["conv_out_wd"], 2, 32, F0)
            # conv_in / conv_out images are built from the fp32 masters themselves
            plan.finish(self)
            pk["_plan"] = plan
        return pk

    def _temb_table(self, pk):
        """BetaEmbedding for every timestep at once (:95-120): the MLP sees only the table row te[t], so no-grad forwards just look
        the result up (each row is computed exactly as the per-batch evaluation computes it).  Built on first use per weight version."""
        if "temb_table" not in pk:
            lin = self.temb.weight
            e = ops.linear_small(self.temb.te, lin[0].weight, lin[0].bias, act=1)
            e = ops.linear_small(e, lin[1].weight, lin[1].bias, act=1)
            pk["temb_table"] = ops.linear_small(e, lin[2].weight, lin[2].bias, act=0)   # [num_timesteps, sum(embedding_size)]
        return pk["temb_table"]

    def _packed_for(self, train):
        base, walk = self._pack_key(train)
        pk = self._packed
        fresh = pk is not None and self._packed_key[0] == base and (not train or (pk["train"] and self._packed_key[1] == walk))
        if not fresh:
            plan = pk.get("_plan") if (pk is not None and train and pk["train"]) else None
            if plan is not None and plan.matches(self):
                plan.refresh()                                   # same parameter storage, new values: rewrite the copies in place
                pk.pop("temb_table", None)
                self._packed_key = (base, walk)
            else:
                self._packed, self._packed_key = self._pack(train), (base, walk)
        return self._packed

    def _const(self, key, build):
        if key not in self._consts:
            self._consts[key] = build()
        return self._consts[key]

    # ------------------------------------------------------------------ forward
    def _res_block(self, h, st_h, pb, temb, temb_off, temb_stride, stats, nxt, res2=None):
        """models/diffusion.py:42-56; returns (h', stats of h').  Levels 0-3 (C = 32 / 64 row-walking kernel, C = 96 / 128 window
        kernel): 3 launches, GroupNorm(+SiLU) fused into the convolutions' operand path; levels 4-5: 5 launches with materialised
        activations."""
        tb = temb[:, temb_off:]
        s0, s1, s2 = stats[nxt()], stats[nxt()], stats[nxt()]
        if ops.conv3x3_gn_supported(pb["C"], h.shape[2]):
            c0 = ops.conv3x3_gn(h, st_h, pb["g0"], pb["b0"], GN_EPS, 2, pb["w0"], bias_b=tb, bias_b_stride=temb_stride, act=1, stats=s0, wimg=pb["w0_img"])
            c1 = ops.conv3x3_gn(c0, s0, pb["g1"], pb["b1"], GN_EPS, 1, pb["w1"], bias=pb["cb1"], act=1, stats=s1, wimg=pb["w1_img"])
            out = ops.gn_apply(c1, s1, pb["g2"], None, 2, GN_EPS, res=h, stats_out=s2, out=c0, res2=res2)
            return out, s2
        a = ops.gn_apply(h, st_h, pb["g0"], pb["b0"], 1, GN_EPS)                                      # silu(GN0(h))
        c0 = ops.conv3x3(a, pb["w0"], bias_b=tb, bias_b_stride=temb_stride, act=1, stats=s0, wimg=pb["w0_img"])
        n1 = ops.gn_apply(c0, s0, pb["g1"], pb["b1"], 0, GN_EPS, out=a)                               # GN1(silu(conv0 + temb))
        c1 = ops.conv3x3(n1, pb["w1"], bias=pb["cb1"], act=1, stats=s1, out=c0, wimg=pb["w1_img"])     # silu(conv1 + bias)
        out = ops.gn_apply(c1, s1, pb["g2"], None, 2, GN_EPS, res=h, stats_out=s2, out=n1, res2=res2)  # h + GN2(.)
        return out, s2

    @torch.no_grad()
    def _forward_impl(self, x, t):
        if not x.is_cuda:
            raise RuntimeError("ddim_audio_b200.Model.forward needs CUDA tensors: there is no CPU path")
        cfg = self.config
        B, Cin, T, F = x.shape
        nl = self.levels
        if Cin != cfg.channels or F != cfg.f_size or T % (2 ** (nl - 1)) != 0:
            raise RuntimeError(f"ddim_audio_b200: unsupported input shape {tuple(x.shape)} (need C={cfg.channels}, F={cfg.f_size}, T % {2 ** (nl - 1)} == 0)")
        pk = self._packed_for(False)
        dev = x.device
        x = x.contiguous().float()
        t = t.to(device=dev, dtype=torch.int64).contiguous()

        # ---- timestep embedding (:107-120), fp32: one row lookup into the table packed for this weight version
        temb = ops.gather_rows(self._temb_table(pk), t)                             # [B, sum(embedding_size)]
        temb_stride = temb.shape[1]
        offs = [0]
        for w in self.embedding_size:
            offs.append(offs[-1] + w)
        rb_i = 0

        n_stats = 2 + 2 * (nl - 1) + 3 * len(self.embedding_size)
        stats = torch.zeros((n_stats, B, GN_GROUPS, 2), dtype=torch.float64, device=dev)
        counter = [0]

        def nxt():
            counter[0] += 1
            return counter[0] - 1

        # ---- down path (:254-264)
        conv0 = self.down_modules[0]
        st = stats[nxt()]
        h = ops.conv_in(x, conv0.weight, conv0.bias, st, wimg=pk["conv_in_img"])
        hidden = [h]
        for l, ent in enumerate(pk["down"]):
            if ent["down"] is not None:
                st = stats[nxt()]
                h = ops.conv_down4x4(h, ent["down"]["w"], ent["down"]["b"], st, wimg=ent["down"]["w_img"])
            for pb in ent["blocks"]:
                h, st = self._res_block(h, st, pb, temb, offs[rb_i], temb_stride, stats, nxt)
                rb_i += 1
            hidden.append(h)

        # ---- FNet bottleneck (:267-279)
        tr = self.transformer
        S, Wt, Cl = h.shape[1], h.shape[2], h.shape[3]
        D, H = Cl * Wt, tr.hidden
        rows = B * S
        te = self._const(("te", S, D, dev), lambda: sinusoid_table(2 ** math.ceil(math.log2(S)) if S > 1 else 1, D)[:S].contiguous().to(dev))
        tok = ops.tokens_ln(h, te, tr.embedding.LayerNorm.weight, tr.embedding.LayerNorm.bias, tr.ln_eps)
        xr = ops.linear(tok, pk["proj_w"], tr.embedding.projection.bias, out_f32=True)          # residual stream fp32 [rows, H]
        wh, a2, Sp = self._const(("dft", S, H, dev), lambda: _fourier_operands(S, H, dev))
        ut = self._const(("ut", B, S, H, dev), lambda: torch.zeros((B, 2 * H, 2 * Sp), dtype=torch.bfloat16, device=dev))
        xbf = torch.empty((rows, 2 * H), dtype=torch.bfloat16, device=dev)
        hbf = torch.empty((rows, H), dtype=torch.bfloat16, device=dev)
        ops.split_bf16(xr, xbf, True)
        for lay, pw in zip(tr.encoder.layer, pk["fnet"]):
            y = ops.fourier_mix(xbf, 2 * H, wh, a2, ut, B, S, H, 3)
            ln = lay.fourier.output.LayerNorm
            hr = ops.layernorm512(xr, y, ln.weight, ln.bias, tr.ln_eps, hbf, False)
            u = ops.linear(hbf, pw["w1"], lay.intermediate.dense.bias, act=2)
            v = ops.linear(u, pw["w2"], lay.output.dense.bias, out_f32=True)
            ln = lay.output.LayerNorm
            xr = ops.layernorm512(v, hr, ln.weight, ln.bias, tr.ln_eps, xbf, True)
        tok_out = ops.linear(xbf, pk["out_w"], tr.compute_out.bias, out_f32=True, M=rows, K=H)   # [rows, D]
        st = stats[nxt()]
        h = ops.tokens_to_nhwc(tok_out, hidden.pop(), st)                                          # + first additive skip (:284)

        # ---- up path (:282-292)
        n_blocks = len(self.embedding_size)
        for ent in pk["up"]:
            for pb in ent["blocks"]:
                # the very last block also adds the final additive skip (:284 before conv_out): conv(h + skip) then reads ONE tensor
                # (conv_out is bound by the TMA delivery rate of its two 64-byte-row input planes)
                last = rb_i == n_blocks - 1
                h, st = self._res_block(h, st, pb, temb, offs[rb_i], temb_stride, stats, nxt, res2=hidden[0] if last else None)
                rb_i += 1
            if ent["up"] is not None:
                st = stats[nxt()]
                h = ops.conv_up4x4(h, ent["up"]["w"], ent["up"]["b"], skip=hidden.pop(), stats=st, wimg=ent["up"]["w_img"])
        conv_last = self.up_modules[-1]
        hidden.pop()
        return ops.conv_out(h, None, conv_last.weight, conv_last.bias, wimg=pk["conv_out_img"])

    def forward(self, input, t):
        if torch.is_grad_enabled() and any(p.requires_grad for p in self._params()):
            # training: one autograd node over the whole network (models/training.py); backward runs in library kernels
            from .training import DenoiserFn
            if not input.is_cuda:
                raise RuntimeError("ddim_audio_b200.Model.forward needs CUDA tensors: there is no CPU path")
            cfg = self.config
            B, Cin, T, F = input.shape
            if Cin != cfg.channels or F != cfg.f_size or T % (2 ** (self.levels - 1)) != 0:
                raise RuntimeError(f"ddim_audio_b200: unsupported input shape {tuple(input.shape)}")
            x = input.detach().contiguous().float()
            tt = t.to(device=x.device, dtype=torch.int64).contiguous()
            return DenoiserFn.apply(self, x, tt, *self._params())
        return self._forward_impl(input, t)
