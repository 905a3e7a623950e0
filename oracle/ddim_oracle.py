"""CPU restatement of the klae01/ddim-audio hot path.  TEST INFRASTRUCTURE ONLY.

This file is the *oracle*: a from-scratch restatement, in plain CPU PyTorch (fp32 by
default, fp64 on request), of the arithmetic the reference performs on its one hot path.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it; nothing under ``ddim_audio_b200/`` does, and the
product path raises if the CUDA library is missing instead of falling back to this.

Parity pin: the reference ships no tests, golden vectors or fixtures, so
the oracle is pinned against *outputs of the reference itself*, produced in the build
container by ``oracle/make_golden.py`` (which imports ``/root/reference`` read-only) and
committed under ``tests/golden/``; ``tests/test_oracle_golden.py`` replays them.

Every function cites the reference lines it follows (paths relative to /root/reference;
``hf:`` = site-packages/transformers, version 5.5.0 in this image — the reference leaves the
dependency un-pinned, README.md:8).
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# configuration (configs/audio.yml:1-108) -- only the keys the hot path reads
# --------------------------------------------------------------------------------------
AUDIO_CFG = dict(
    channels=2,              # audio.yml:44
    f_size=256,              # audio.yml:1,46
    t_size=1024,             # audio.yml:4,45
    ch=(32, 64, 96, 128, 192, 256),   # audio.yml:48
    krn=(3, 3, 3, 3, 3, 3),           # audio.yml:49
    res=(2, 2, 3, 3, 3, 3),           # audio.yml:50
    num_groups=8,            # models/diffusion.py:19
    gn_eps=1e-6,             # models/diffusion.py:20
    hidden=512,              # audio.yml:34,41
    ffn=2048,                # audio.yml:36
    layers=12,               # audio.yml:35
    ln_eps=1e-6,             # audio.yml:40
    pos_ch=128,              # models/diffusion.py:98
    emb_ch=512,              # models/diffusion.py:99
    num_timesteps=1000,      # audio.yml:60
    beta_start=1e-4,         # audio.yml:58
    beta_end=2e-2,           # audio.yml:59
    sampling_t_size=8192,    # audio.yml:77
)


def embedding_sizes(cfg=AUDIO_CFG) -> List[int]:
    """models/diffusion.py:177-183 -- per-res-block temb widths, down then mirrored."""
    down = [c for r, c in zip(cfg["res"], cfg["ch"]) for _ in range(r)]
    return down + down[::-1]


def token_width(cfg=AUDIO_CFG) -> int:
    """models/diffusion.py:228-232 -- ch[-1] * f_size / 2**(levels-1)."""
    return cfg["ch"][-1] * (cfg["f_size"] // (2 ** (len(cfg["ch"]) - 1)))


# --------------------------------------------------------------------------------------
# state_dict layout (models/diffusion.py:170-235, 12-56, 59-78, 81-120, 123-167)
# --------------------------------------------------------------------------------------
def state_dict_spec(cfg=AUDIO_CFG) -> "OrderedDict[str, Tuple[int, ...]]":
    """Names and shapes of the 389 entries of ``Model(config).state_dict()`` for audio.yml,
    in the reference's registration order."""
    spec: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    ch, res = cfg["ch"], cfg["res"]
    emb = sum(embedding_sizes(cfg))
    spec["temb.te"] = (cfg["num_timesteps"], cfg["pos_ch"])
    for i, (o, n) in enumerate([(cfg["emb_ch"], cfg["pos_ch"]), (cfg["emb_ch"], cfg["emb_ch"]), (emb, cfg["emb_ch"])]):
        spec[f"temb.weight.{i}.weight"] = (o, n)
        spec[f"temb.weight.{i}.bias"] = (o,)

    def rb(prefix, c):
        for j in range(3):
            spec[f"{prefix}.norm.{j}.weight"] = (c,)
            if j < 2:
                spec[f"{prefix}.norm.{j}.bias"] = (c,)
        spec[f"{prefix}.conv.0.weight"] = (c, c, 3, 3)
        spec[f"{prefix}.conv.1.weight"] = (c, c, 3, 3)
        spec[f"{prefix}.conv.1.bias"] = (c,)

    spec["down_modules.0.weight"] = (ch[0], cfg["channels"], 3, 3)
    spec["down_modules.0.bias"] = (ch[0],)
    for l, (c, r) in enumerate(zip(ch, res)):
        k = 0
        if l > 0:
            spec[f"down_modules.{l+1}.0.conv.weight"] = (c, ch[l - 1], 4, 4)
            spec[f"down_modules.{l+1}.0.conv.bias"] = (c,)
            k = 1
        for j in range(r):
            rb(f"down_modules.{l+1}.{k+j}", c)
    # up_modules is built [conv_out, level0, ..., level5] then reversed (:209-227): index 0 = deepest
    nl = len(ch)
    for u in range(nl):
        l = nl - 1 - u
        c, r = ch[l], res[l]
        for j in range(r):
            rb(f"up_modules.{u}.{j}", c)
        if l > 0:
            spec[f"up_modules.{u}.{r}.conv.weight"] = (c, ch[l - 1], 4, 4)   # ConvTranspose2d: [in, out, kh, kw]
            spec[f"up_modules.{u}.{r}.conv.bias"] = (ch[l - 1],)
    spec[f"up_modules.{nl}.weight"] = (cfg["channels"], ch[0], 3, 3)
    spec[f"up_modules.{nl}.bias"] = (cfg["channels"],)
    tw, h, f = token_width(cfg), cfg["hidden"], cfg["ffn"]
    spec["transformer.embedding.LayerNorm.weight"] = (tw,)
    spec["transformer.embedding.LayerNorm.bias"] = (tw,)
    spec["transformer.embedding.projection.weight"] = (h, tw)
    spec["transformer.embedding.projection.bias"] = (h,)
    for i in range(cfg["layers"]):
        p = f"transformer.encoder.layer.{i}"
        spec[f"{p}.fourier.output.LayerNorm.weight"] = (h,)
        spec[f"{p}.fourier.output.LayerNorm.bias"] = (h,)
        spec[f"{p}.intermediate.dense.weight"] = (f, h)
        spec[f"{p}.intermediate.dense.bias"] = (f,)
        spec[f"{p}.output.dense.weight"] = (h, f)
        spec[f"{p}.output.dense.bias"] = (h,)
        spec[f"{p}.output.LayerNorm.weight"] = (h,)
        spec[f"{p}.output.LayerNorm.bias"] = (h,)
    spec["transformer.compute_out.weight"] = (tw, h)
    spec["transformer.compute_out.bias"] = (tw,)
    return spec


def sinusoid_table(length: int, channel: int, dtype=torch.float32) -> torch.Tensor:
    """models/diffusion.py:81-92 (Add_Encoding on a zero table): sin on even columns, cos on
    odd columns, frequency exp(-ln(1e4) * 2i / channel)."""
    position = torch.arange(length, dtype=dtype).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, channel, 2, dtype=dtype) * (-math.log(10000.0) / channel))
    x = position * div_term
    te = torch.zeros(length, channel, dtype=dtype)
    te[..., 0::2] += torch.sin(x)
    te[..., 1::2] += torch.cos(x)
    return te


def make_state_dict(seed: int = 0, cfg=AUDIO_CFG, live_norm2: bool = True) -> "OrderedDict[str, torch.Tensor]":
    """Deterministic synthetic weights with the reference's names/shapes.

    The reference's own init (torch default Conv/Linear init, models/diffusion.py:25 zeroes
    every norm.2.weight) makes each Residual_Block the identity (SURVEY.md §4 tier 2), so
    parity runs use these weights instead: fan-in-scaled uniform for conv/linear weights,
    small biases, GroupNorm/LayerNorm gains around 1 and -- with ``live_norm2`` -- a non-zero
    norm.2.weight so the convs matter.  The generator is consumed in spec order, so the same
    seed gives bit-identical weights on any machine with this torch build; the golden
    fixtures (oracle/make_golden.py) load exactly this dict into the real reference Model.
    """
    g = torch.Generator().manual_seed(seed)
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for name, shape in state_dict_spec(cfg).items():
        if name == "temb.te":
            sd[name] = sinusoid_table(*shape)
            continue
        leaf = name.rsplit(".", 1)[1]
        is_norm = ".norm." in name or "LayerNorm" in name
        if is_norm and leaf == "weight":
            if ".norm.2." in name:
                t = 0.5 * torch.randn(shape, generator=g) if live_norm2 else torch.zeros(shape)
            else:
                t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif leaf == "bias":
            t = 0.05 * torch.randn(shape, generator=g)
        else:
            if name.startswith("up_modules") and name.endswith(".conv.weight"):
                fan_in = shape[0] * 4          # ConvTranspose2d k4 s2: 4 taps of C_in reach each output
            else:
                fan_in = int(np.prod(shape[1:]))
            bound = math.sqrt(3.0 / fan_in)
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        sd[name] = t.to(torch.float32)
    return sd


# --------------------------------------------------------------------------------------
# schedule (runners/diffusion.py:32-62, 91-128)
# --------------------------------------------------------------------------------------
def beta_schedule(cfg=AUDIO_CFG) -> np.ndarray:
    """runners/diffusion.py:46-49 -- linear betas in float64."""
    return np.linspace(cfg["beta_start"], cfg["beta_end"], cfg["num_timesteps"], dtype=np.float64)


def alphas_cumprod(cfg=AUDIO_CFG) -> torch.Tensor:
    """runners/diffusion.py:108-115 -- [1, 1-beta] cast to fp32 FIRST, cumprod in fp32, drop
    the leading 1."""
    betas = beta_schedule(cfg)
    alphas = np.concatenate([[1], 1.0 - betas], axis=-1)
    alphas = torch.from_numpy(alphas).to(torch.float32)
    return alphas.cumprod(dim=0)[1:]


def uniform_seq(num_timesteps: int, timesteps: int) -> range:
    """runners/diffusion.py:482-484."""
    return range(0, num_timesteps, num_timesteps // timesteps)


def select_index_for(timesteps: int, sequence: int):
    """runners/diffusion.py:388-394."""
    if sequence in (-1, 0):
        return range(timesteps)
    index = np.linspace(1, timesteps, sequence, dtype=np.int32)
    return set((timesteps - index).tolist())


# --------------------------------------------------------------------------------------
# denoiser forward (models/diffusion.py:237-294)
# --------------------------------------------------------------------------------------
def _gn(x, w, b, cfg):
    return F.group_norm(x, cfg["num_groups"], w, b, cfg["gn_eps"])


def residual_block(sd, p: str, x: torch.Tensor, temb: torch.Tensor, cfg=AUDIO_CFG) -> torch.Tensor:
    """models/diffusion.py:42-56: GN0 -> SiLU -> conv0 (+temb) -> SiLU -> GN1 -> conv1(+bias) ->
    SiLU -> GN2 (weight only) -> + input."""
    h = F.silu(_gn(x, sd[f"{p}.norm.0.weight"], sd[f"{p}.norm.0.bias"], cfg))
    h = F.conv2d(h, sd[f"{p}.conv.0.weight"], None, padding=1) + temb[..., None, None]
    h = _gn(F.silu(h), sd[f"{p}.norm.1.weight"], sd[f"{p}.norm.1.bias"], cfg)
    h = F.conv2d(h, sd[f"{p}.conv.1.weight"], sd[f"{p}.conv.1.bias"], padding=1)
    h = _gn(F.silu(h), sd[f"{p}.norm.2.weight"], None, cfg)
    return x + h


def beta_embedding(sd, t: torch.Tensor) -> torch.Tensor:
    """models/diffusion.py:107-120: te[t] -> Linear -> SiLU -> Linear -> SiLU -> Linear."""
    x = sd["temb.te"].index_select(0, t)
    x = F.silu(F.linear(x, sd["temb.weight.0.weight"], sd["temb.weight.0.bias"]))
    x = F.silu(F.linear(x, sd["temb.weight.1.weight"], sd["temb.weight.1.bias"]))
    return F.linear(x, sd["temb.weight.2.weight"], sd["temb.weight.2.bias"])


def gelu_new(x: torch.Tensor) -> torch.Tensor:
    """hf:activations.py:59-66 (NewGELUActivation, tanh form)."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * torch.pow(x, 3.0))))


def fnet_layer(sd, p: str, x: torch.Tensor, cfg=AUDIO_CFG) -> torch.Tensor:
    """hf:models/fnet/modeling_fnet.py:142-258 in eval mode: h = LN(x + Re(fftn(x, dim=(1,2))));
    y = LN(h + W2 gelu_new(W1 h))."""
    eps, hdim = cfg["ln_eps"], (cfg["hidden"],)
    f = torch.fft.fftn(x, dim=(1, 2)).real
    h = F.layer_norm(x + f, hdim, sd[f"{p}.fourier.output.LayerNorm.weight"], sd[f"{p}.fourier.output.LayerNorm.bias"], eps)
    u = gelu_new(F.linear(h, sd[f"{p}.intermediate.dense.weight"], sd[f"{p}.intermediate.dense.bias"]))
    v = F.linear(u, sd[f"{p}.output.dense.weight"], sd[f"{p}.output.dense.bias"])
    return F.layer_norm(v + h, hdim, sd[f"{p}.output.LayerNorm.weight"], sd[f"{p}.output.LayerNorm.bias"], eps)


def transformer_module(sd, x: torch.Tensor, cfg=AUDIO_CFG) -> torch.Tensor:
    """models/diffusion.py:130-145 (TransformerEmbedding, eval: dropout off) + :157-167."""
    s, w = x.shape[1], x.shape[2]
    size = 2 ** math.ceil(math.log2(s))                                   # :133-134
    te = sinusoid_table(size, w, dtype=x.dtype)[:s]                       # :135-140
    x = F.layer_norm(x + te, (w,), sd["transformer.embedding.LayerNorm.weight"],
                     sd["transformer.embedding.LayerNorm.bias"], cfg["ln_eps"])
    x = F.linear(x, sd["transformer.embedding.projection.weight"], sd["transformer.embedding.projection.bias"])
    for i in range(cfg["layers"]):
        x = fnet_layer(sd, f"transformer.encoder.layer.{i}", x, cfg)
    return F.linear(x, sd["transformer.compute_out.weight"], sd["transformer.compute_out.bias"])


def model_forward(sd, x: torch.Tensor, t: torch.Tensor, cfg=AUDIO_CFG, taps: Optional[dict] = None) -> torch.Tensor:
    """models/diffusion.py:237-294, eval mode.  ``x`` is [B, C, T, F]; ``t`` int64 [B].
    ``taps`` (optional dict) receives named intermediates for block-level parity tests."""
    ch, res = cfg["ch"], cfg["res"]
    temb = iter(torch.split(beta_embedding(sd, t), embedding_sizes(cfg), dim=-1))   # :248-250
    hidden = []
    h = F.conv2d(x, sd["down_modules.0.weight"], sd["down_modules.0.bias"], padding=1)
    hidden.append(h)
    for l, r in enumerate(res):                                                       # :254-264
        k = 0
        if l > 0:
            h = F.conv2d(h, sd[f"down_modules.{l+1}.0.conv.weight"], sd[f"down_modules.{l+1}.0.conv.bias"], stride=2, padding=1)
            k = 1
        for j in range(r):
            h = residual_block(sd, f"down_modules.{l+1}.{k+j}", h, next(temb), cfg)
        hidden.append(h)
        if taps is not None:
            taps[f"down{l}"] = h
    y = torch.permute(h, (0, 2, 1, 3))                                                # :273
    shape = y.shape
    y = transformer_module(sd, y.reshape(*shape[:2], -1), cfg)                        # :275-276
    h = torch.permute(y.reshape(*shape), (0, 2, 1, 3))                                # :277-278
    if taps is not None:
        taps["mid"] = h
    nl = len(ch)
    for u in range(nl + 1):                                                           # :282-292
        h = h + hidden.pop()
        if u == nl:
            h = F.conv2d(h, sd[f"up_modules.{nl}.weight"], sd[f"up_modules.{nl}.bias"], padding=1)
            break
        l = nl - 1 - u
        for j in range(res[l]):
            h = residual_block(sd, f"up_modules.{u}.{j}", h, next(temb), cfg)
        if l > 0:
            h = F.conv_transpose2d(h, sd[f"up_modules.{u}.{res[l]}.conv.weight"], sd[f"up_modules.{u}.{res[l]}.conv.bias"], stride=2, padding=1)
        if taps is not None:
            taps[f"up{l}"] = h
    return h


class OracleModel:
    """Callable ``model(x, t)`` over a state_dict, for generalized_steps / the loss."""

    def __init__(self, sd, cfg=AUDIO_CFG, dtype=torch.float32):
        self.cfg = cfg
        self.sd = {k: v.to(dtype) for k, v in sd.items()}
        self.dtype = dtype

    def __call__(self, x, t):
        with torch.no_grad():
            return model_forward(self.sd, x.to(self.dtype), t.long(), self.cfg)


# --------------------------------------------------------------------------------------
# DDIM sampler (functions/denoising.py:10-52)
# --------------------------------------------------------------------------------------
def ddim_coefficients(alpha: torch.Tensor, seq: Sequence[int], eta: float = 0.0):
    """Host-side float64 scalars exactly as functions/denoising.py:12,22-24,36-40 forms them:
    alpha list = [1.0] + fp32 ᾱ as Python floats; per step (i, j) in reversed(seq, seq_next):
    at, at_next, c1, c2.  Returns a list of dicts in execution order."""
    al = [1.0] + alpha.to("cpu", torch.float32).numpy().tolist()
    seq = list(seq)
    seq_next = [-1] + seq[:-1]
    out = []
    for i, j in zip(reversed(seq), reversed(seq_next)):
        at, at_next = al[int(i) + 1], al[int(j) + 1]
        c1 = eta * ((1 - at / at_next) * (1 - at_next) / (1 - at)) ** 0.5
        c2 = ((1 - at_next) - c1 ** 2) ** 0.5
        out.append(dict(i=int(i), j=int(j), at=at, at_next=at_next, c1=c1, c2=c2))
    return out


def generalized_steps(x, seq, model, alpha, select_index, noise: Optional[Iterable[torch.Tensor]] = None, **kwargs):
    """functions/denoising.py:10-52 on CPU.  Differences from the literal text, all forced by
    running off-GPU and stated in SURVEY.md §8c: the working copy is a clone (on CUDA the
    reference aliases its input), snapshots are clones (``.to('cpu')`` of a CPU tensor is the
    same object), and the N(0,1) draw of :42 can be injected through ``noise`` (an iterable
    of [B,C,T,F] tensors, one per step) so eta>0 runs are reproducible across devices."""
    eta = kwargs.get("eta", 0)
    n = x.size(0)
    xs, x0_preds = [x.clone()], []
    xt = x.clone().to(torch.float32)
    t = torch.zeros(n, dtype=torch.long)
    noise_it = iter(noise) if noise is not None else None
    L = len(list(seq))
    for index, c in enumerate(ddim_coefficients(alpha, seq, eta)):
        t[...] = c["i"]
        et = model(xt, t.long()).to(torch.float32)
        xt.add_(et, alpha=-((1 - c["at"]) ** 0.5)).div_(c["at"] ** 0.5)            # :27
        sel = select_index is None or index in select_index or index - L in select_index
        if sel:
            x0_preds.append(xt.clone())                                               # :29-34
        z = next(noise_it) if noise_it is not None else torch.randn_like(x)          # :42 (always drawn)
        xt.mul_(c["at_next"] ** 0.5).add_(et, alpha=c["c2"]).add_(z, alpha=c["c1"])  # :41-43
        if sel:
            xs.append(xt.clone())                                                     # :45-50
    return xs, x0_preds


# --------------------------------------------------------------------------------------
# training loss (functions/losses.py:4-18)
# --------------------------------------------------------------------------------------
def q_sample(x0, t, e, a):
    """functions/losses.py:12-13."""
    a = a.index_select(0, t).view(-1, 1, 1, 1)
    return x0 * a.sqrt() + e * (1.0 - a).sqrt()


def noise_estimation_loss(model, x0, t, e, a, keepdim=False):
    """functions/losses.py:4-18."""
    x = q_sample(x0, t, e, a)
    output = model(x, t.long())
    per = (e - output).square().sum(dim=(1, 2, 3))
    return per if keepdim else per.mean(dim=0)


def loss_and_grads(sd, x0, t, e, a, cfg=AUDIO_CFG, dtype=torch.float32):
    """functions/losses.py:4-18 followed by ``loss.backward()`` (runners/diffusion.py:154) in eval mode (dropout off):
    returns (loss, {parameter name: gradient}) from torch autograd over the restated forward."""
    params = {k: v.detach().to(dtype).clone().requires_grad_(k != "temb.te") for k, v in sd.items()}
    x = q_sample(x0.to(dtype), t, e.to(dtype), a.to(dtype))
    out = model_forward(params, x, t.long(), cfg)
    loss = (e.to(dtype) - out).square().sum(dim=(1, 2, 3)).mean(dim=0)
    loss.backward()
    return loss.detach(), {k: p.grad for k, p in params.items() if p.grad is not None}


def grad_fingerprint(name: str, g: torch.Tensor):
    """(L2 norm, sum, projection on a seeded N(0,1) direction) of one gradient tensor, float64: what tests/golden/grad_*.npz
    stores per parameter (full gradients are 47 M floats)."""
    seed = sum((i + 1) * ord(c) for i, c in enumerate(name)) % (2 ** 31)
    d = torch.randn(g.numel(), generator=torch.Generator().manual_seed(seed), dtype=torch.float64)
    gd = g.detach().double().flatten().cpu()
    return float(gd.norm()), float(gd.sum()), float(gd @ d)


def antithetic_t(n: int, num_timesteps: int, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """runners/diffusion.py:141-142."""
    t = torch.randint(low=0, high=num_timesteps, size=((n + 1) // 2,), generator=generator)
    return torch.cat([t, num_timesteps - t - 1], dim=0)[:n]


# --------------------------------------------------------------------------------------
# seeded synthetic inputs shared by tests / smoke / bench
# --------------------------------------------------------------------------------------
def synthetic_inputs(batch: int, t_size: int, seed: int = 1234, cfg=AUDIO_CFG):
    g = torch.Generator().manual_seed(seed)
    shape = (batch, cfg["channels"], t_size, cfg["f_size"])
    x = torch.randn(shape, generator=g)
    e = torch.randn(shape, generator=g)
    t = antithetic_t(batch, cfg["num_timesteps"], g)
    return x, e, t


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


# --------------------------------------------------------------------------------------
# denoise_2d post-filter on every returned x0
# External/UPU/signal/denoise.py:45-70 (torch branch); call site runners/diffusion.py:400-401
# --------------------------------------------------------------------------------------
def lower_median(x: torch.Tensor, dim: int) -> torch.Tensor:
    """``Tensor.median(dim, keepdim=True)[0]``: for an even count torch returns the LOWER of the two middle values,
    i.e. element ``(n - 1) // 2`` of the sorted slice (denoise.py:54-55 relies on it)."""
    n = x.shape[dim]
    return torch.sort(x, dim=dim)[0].narrow(dim, (n - 1) // 2, 1)


def denoise_2d(x: torch.Tensor, iters: int = 5) -> torch.Tensor:
    """denoise.py:45-70 for ``x[*, D, H, W]``: subtract the median over W, then the median over H of the result (:54-55);
    then 5 times (:63-69): ``denom = 1 / (1e-9 + ||r||_2 over D)``, ``num = r * denom`` (formed ONCE per iteration),
    and for axis W then H: ``r -= mean_axis(num) / mean_axis(denom)``."""
    r = x - lower_median(x, -1)
    r = r - lower_median(r, -2)
    for _ in range(iters):
        denom = 1 / (1e-9 + torch.linalg.norm(r, ord=2, dim=-3, keepdim=True))
        num = r * denom
        for axis in (-1, -2):
            a = num.mean(dim=axis, keepdim=True) / denom.mean(dim=axis, keepdim=True)
            r = r - a
    return r


def denoise_inputs(batch: int, t_size: int, seed: int, cfg=AUDIO_CFG, f_size: Optional[int] = None) -> torch.Tensor:
    """Seeded x0-like input for the post-filter: unit noise plus the per-row / per-column offsets and the outliers the
    filter exists to remove (so the medians and the five mean iterations all do real work)."""
    g = torch.Generator().manual_seed(seed)
    F_ = cfg["f_size"] if f_size is None else f_size
    C = cfg["channels"]
    x = torch.randn(batch, C, t_size, F_, generator=g)
    x = x + 0.7 * torch.randn(batch, C, t_size, 1, generator=g) + 1.3 * torch.randn(batch, C, 1, F_, generator=g)
    spikes = (torch.rand(batch, C, t_size, F_, generator=g) < 0.01).float() * 25.0 * torch.randn(batch, C, t_size, F_, generator=g)
    return x + spikes + 0.25


# --------------------------------------------------------------------------------------
# EMA of the parameters -- models/ema.py:16-23, runners/diffusion.py:182-183
# --------------------------------------------------------------------------------------
def ema_update(shadow: Dict[str, torch.Tensor], params: Dict[str, torch.Tensor], mu: float) -> Dict[str, torch.Tensor]:
    """models/ema.py:19-23: ``shadow = (1 - mu) * param + mu * shadow`` (two rounded products, then the sum), per tensor."""
    return {k: (1.0 - mu) * params[k] + mu * shadow[k] for k in shadow}


def ema_case(seed: int, steps: int = 3):
    """Seeded parameter trajectories for the EMA fixtures: ragged sizes (scalar, odd, > one 16 K chunk)."""
    g = torch.Generator().manual_seed(seed)
    shapes = {"a.weight": (7, 5, 3, 3), "a.bias": (7,), "b.weight": (130, 257), "c.scale": (1,), "d.weight": (3, 16384 + 5)}
    traj = [{k: torch.randn(s, generator=g) for k, s in shapes.items()} for _ in range(steps + 1)]
    return shapes, traj


# --------------------------------------------------------------------------------------
# optimizer tail of the train step -- runners/diffusion.py:155-173
# --------------------------------------------------------------------------------------
def clip_grad_norm(grads: Dict[str, torch.Tensor], max_norm: float):
    """torch.nn.utils.clip_grad_norm_ (L2) as called at runners/diffusion.py:155-160: returns (total_norm, clipped grads)."""
    total = torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(g) for g in grads.values()]))
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    return total, {k: g * coef for k, g in grads.items()}


def adamw_step(p, g, m, v, step: int, lr: float, betas, eps: float, weight_decay: float):
    """torch.optim.AdamW single-tensor arithmetic (functions/__init__.py:15-23 constructs it): decoupled decay, lerp first
    moment, bias corrections as Python floats.  Returns new (p, m, v)."""
    b1, b2 = betas
    p = p * (1 - lr * weight_decay)
    m = m + (g - m) * (1 - b1)
    v = v * b2 + (1 - b2) * g * g
    bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
    denom = v.sqrt() / (bc2 ** 0.5) + eps
    return p + (-(lr / bc1)) * (m / denom), m, v


def adabelief_step(p, g, m, v, step: int, lr: float, betas, eps: float, weight_decay: float):
    """External/step-clip-optimizer/clip_opt/adabelief/AdaBelief.py:112-142 with the reference's configuration
    (functions/__init__.py:36-41: weight_decouple, not fixed_decay, no amsgrad / rectify; audio.yml clip_step null)."""
    b1, b2 = betas
    m = m * b1 + (1 - b1) * g                                   # :113
    r = g - m                                                   # :114
    v = v * b2 + (1 - b2) * r * r                               # :115-117
    bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step                   # :119-120
    denom = v.sqrt() / (bc2 ** 0.5) + eps                       # :130
    if weight_decay != 0:
        p = p * (1.0 - weight_decay)                            # :134-137 (rate = [1, lr][fixed_decay] * weight_decay)
    return p + (-lr / bc1) * (m / denom), m, v                  # :138


def optim_case(seed: int, steps: int = 3):
    """Seeded parameters and per-step gradients for the optimizer fixtures (ragged sizes, one tensor > one 64 K chunk)."""
    g = torch.Generator().manual_seed(seed)
    shapes = {"a.weight": (7, 5, 3, 3), "a.bias": (7,), "b.weight": (130, 257), "c.scale": (1,), "d.weight": (2, 65536 + 3)}
    params = {k: torch.randn(s, generator=g) for k, s in shapes.items()}
    grads = [{k: 0.3 * torch.randn(s, generator=g) * (1.0 + 4.0 * (i == 1)) for k, s in shapes.items()} for i in range(steps)]
    return params, grads


def optim_probe(t: torch.Tensor) -> torch.Tensor:
    """What the optimizer fixtures store of a tensor: all of it, or every 7th element above 20 K elements."""
    t = t.detach().reshape(-1)
    return t[::7] if t.numel() > 20000 else t
